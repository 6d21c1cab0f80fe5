"""x-slab decomposition.

CPU (`not gpu`): the host-side logic — slab boundaries, rebalancing arithmetic, and the neighbour
routing of the halo messages over a real 2- and 3-process `gloo` group.
GPU: the whole protocol with every rank's handle in ONE process on one device (messages are device
copies instead of NCCL), against the single-handle run: the results must be bit-identical.
"""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_partition_is_contiguous_and_balanced():
    from sph_b200.slabs import partition
    for n, w in ((101, 2), (64 * 1024 * 1024, 8), (10, 3), (7, 7)):
        b = partition(n, w)
        assert b[0] == 0 and b[-1] == n and len(b) == w + 1
        sizes = np.diff(b)
        assert sizes.min() >= 1 and sizes.max() - sizes.min() <= 1


def test_rebalance_shifts_restore_equal_counts():
    from sph_b200.slabs import partition, rebalance_shifts
    counts = [120, 80, 100, 100]
    shifts = rebalance_shifts(counts)           # particles crossing each internal edge, left -> right
    new = counts[:]
    for e, s in enumerate(shifts):
        new[e] -= s
        new[e + 1] += s
    assert new == list(np.diff(partition(sum(counts), len(counts))))
    assert rebalance_shifts([100, 100, 100]) == [0, 0]


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _routing_worker(rank, world, port, q):
    import torch
    import torch.distributed as dist
    sys.path.insert(0, ROOT)
    from sph_b200.slabs import LEFT, RIGHT, agree_max, neighbour_exchange
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    try:
        send = {LEFT: torch.full((5,), 10.0 * rank - 1), RIGHT: torch.full((5,), 10.0 * rank + 1)}
        recv = {LEFT: torch.zeros(5), RIGHT: torch.zeros(5)}
        neighbour_exchange(dist, rank, world, send, recv)
        ok = True
        if rank > 0:                      # my LEFT buffer holds what rank-1 sent to ITS right
            ok &= bool((recv[LEFT] == 10.0 * (rank - 1) + 1).all())
        else:
            ok &= bool((recv[LEFT] == 0).all())
        if rank < world - 1:
            ok &= bool((recv[RIGHT] == 10.0 * (rank + 1) - 1).all())
        else:
            ok &= bool((recv[RIGHT] == 0).all())
        flag = torch.zeros(6, dtype=torch.int32)
        # one rank saw a change, another an error code, each its own h maximum: everybody learns the maxima
        ok &= agree_max(dist, flag, [13 if rank == 0 else 0, 0, 1 if rank == world - 1 else 0, 100 + rank, 7 * rank, rank % 2]) == [13, 0, 1, 99 + world, 7 * (world - 1), 1 if world > 1 else 0]
        ok &= agree_max(dist, flag, [0] * 6) == [0] * 6
        q.put((rank, ok))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_halo_routing_over_gloo(world):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_routing_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert sorted(res) == [(r, True) for r in range(world)]


# ------------------------------------------------------------------------------------------ GPU

def _cfg(n, pc):
    from sph_b200 import Config
    return Config(n_part=n, mass=4.0 / (n - 1), pressure_calc=pc, t_i=0.005 * 100 / (n - 1))


@pytest.mark.gpu
@pytest.mark.parametrize("sweep_mode", [0, 1, 2])      # auto, work lists, serial sweep: the same bits
@pytest.mark.parametrize("n,pc,varh,world", [(20001, 1, True, 2), (20001, 1, True, 3), (30000, 0, False, 2),
                                             (6001, 0, True, 4)])
def test_slabs_in_one_process_match_single_handle_bit_for_bit(n, pc, varh, world, sweep_mode):
    from sph_b200 import Simulation
    from sph_b200.slabs import make_local_slabs
    cfg = _cfg(n, pc)
    fixed_h = 0.2 * 100 / (n - 1)
    one = Simulation(cfg, variable_h=varh, fixed_h=fixed_h, sweep_mode=1)
    one.init_lattice()
    many = make_local_slabs(cfg, world, variable_h=varh, fixed_h=fixed_h, halo_count=128, sweep_mode=sweep_mode)
    many.init_lattice()
    for st in range(0, 9):
        if st:
            one.step(1)
            many.step(1)
        if st in (0, 1, 4, 8):
            f1, t1, g1 = one.fields()
            f2, t2, g2, ids = many.fields()
            assert g1 == g2 and f1.shape == f2.shape
            # alive particles: everything; wall ghosts: all but the dead acc/du_dt of t = 0
            assert np.array_equal(f1[:n], f2[:n]), "step %d: %s" % (st, np.argwhere(f1[:n] != f2[:n])[:4])
            assert np.array_equal(f1[n:, [0, 1, 2, 4, 6, 7]], f2[n:, [0, 1, 2, 4, 6, 7]])
    assert many.rounds_last >= 2
    c = many.ranks[-1].counters()
    if sweep_mode == 2:
        # the serial sweep walked every slab, its left halo included, and what it computed for the halo particles
        # was what their owner computed: no extra round was needed
        assert c.sweep_steps == 8 and many.rounds_last == 2
    if sweep_mode == 1:
        assert c.sweep_steps == 0
    one.close()
    many.close()


@pytest.mark.gpu
def test_too_thin_a_halo_is_reported_by_every_rank():
    """The halo-width verdict is taken from flags agreed over all ranks: a rank raising alone would leave the
    others waiting in the next exchange.  Here the ranks share a process; each is asked separately."""
    from sph_b200 import SphbError
    from sph_b200.slabs import make_local_slabs
    cfg = _cfg(40001, 1)
    many = make_local_slabs(cfg, 3, halo_count=12)       # enough for the kernel itself (~5 particles) ...
    many.init_lattice()
    many.step(2)
    # ... but not for a smoothing length five times the lattice's, wherever in the run it sits
    rec, ng = many.ranks[0].download()
    rec = rec[:many.ranks[0].counters().n_alive].copy()
    rec["h"][len(rec) // 2] *= 5.0
    many.ranks[0].upload(rec, 0)
    many._all("step_begin")
    many._attempt()
    agreed, _ = many._flags()
    raised = 0
    for r in many.ranks:
        with pytest.raises(SphbError) as e:
            r.agree(agreed)
        assert e.value.code == 13
        raised += 1
    assert raised == 3
    many.close()


@pytest.mark.gpu
def test_too_thin_a_halo_is_reported():
    from sph_b200 import SphbError
    from sph_b200.slabs import make_local_slabs
    cfg = _cfg(4001, 1)
    many = make_local_slabs(cfg, 2, halo_count=3)       # the kernel reaches ~5 particles
    with pytest.raises(SphbError) as e:
        many.init_lattice()
        many.step(1)
    assert e.value.code == 13
    many.close()


@pytest.mark.gpu
@pytest.mark.parametrize("native", ["1", "0"])
def test_slabs_over_nccl_match_single_handle(native):
    """One process per GPU, NCCL send/recv (needs >= 2 GPUs on the box; skipped otherwise): the
    exchange inside libsphb (native) and the torch.distributed one."""
    import subprocess
    import torch
    ngpu = torch.cuda.device_count()
    if ngpu < 2:
        pytest.skip("needs at least 2 GPUs")
    world = min(ngpu, 4)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world),
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()),
           os.path.join(ROOT, "tests", "dist_slab_check.py")]
    res = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=900,
                         env=dict(os.environ, SLAB_NATIVE=native))
    assert res.returncode == 0, res.stdout[-3000:]
    assert "DIFFERS" not in res.stdout


@pytest.mark.gpu
def test_particles_crossing_a_slab_edge_change_owner():
    """Two particles either side of the slab edge are thrown at each other so that they swap places in one
    step: both ranks merge-split the particles around the edge (no extra message), the particle counts stay
    equal, and the run stays bit-identical to the single-handle run (which re-sorts)."""
    from sph_b200 import Simulation
    from sph_b200.slabs import make_local_slabs
    n, world = 4000, 2
    cfg = _cfg(n, 1)
    ref = Simulation(cfg)
    ref.init_lattice()
    rec, ng = ref.download()
    rec = rec[:n].copy()                 # alive particles only, array order = x order = global index
    rec["id"] = np.arange(n)
    rec["vel"][n // 2 - 1] = 5.0         # last particle of slab 0 ...
    rec["vel"][n // 2] = -5.0            # ... and first of slab 1 approach by more than one spacing per step
    ref.upload(rec.copy(), 0)
    many = make_local_slabs(cfg, world, halo_count=64)
    for r in many.ranks:
        r.upload(rec[r.i0:r.i1].copy(), 0)
    for st in (1, 2, 3):
        ref.step(1)
        many.step(1)
        f1 = ref.fields()[0]
        f2, t2, g2, ids = many.fields()
        assert np.array_equal(ids, np.arange(n))
        assert np.array_equal(f1[:n], f2[:n]), "step %d: %s" % (st, np.argwhere(f1[:n] != f2[:n])[:4])
    # the two particles now live on the other rank
    own0 = set(many.ranks[0].download()[0]["id"][: n // 2].tolist())
    assert (n // 2) in own0 and (n // 2 - 1) not in own0
    assert ref.counters().sorts >= 1
    ref.close()
    many.close()
