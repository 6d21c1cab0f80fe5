"""x-slab domain decomposition across the GPUs of one node.

The particle set, sorted by x, is cut into `world` contiguous slabs of (nearly) equal particle
count; rank r owns global indices [partition(n, world)[r], partition(n, world)[r + 1]).  Only the
outer ranks make wall ghosts.  Interactions are compact (2.5 max h), so each rank only needs the
`halo_count` particles next to each of its edges from its neighbour; they fill the pad slots of the
device layout (where ghosts live next to a wall), and every kernel runs unchanged.

Per step (libsphb does the device work of each phase, see include/sphb.h `sphb_slab_*`):

    step_begin                 kick-drift, order check, wall ghosts
    exchange BASE              x, m, v_half, u_half, h_old, rho_old, P, idx of the edge particles
    density                    fused density / root-find / omega of the own particles
    exchange RECORDS           owner-computed (h, rho, omega) records of the edge particles
    force_first, refine        first sweep + local convergence of the ordered sweep
    exchange ITERATE, refine   current fully kicked (v, u) of the edge particles; a rank whose copy
                               differed re-queues the tiles in reach and converges again
    exchange ITERATE           check: did anything still move?
    poll + agree               the step's ONE host synchronisation: MAX over ranks of (device error,
                               x-order broken by the drift, halo still changing) -> raise / sort and
                               repeat the attempt / more rounds / done
    step_end

There is no data-path collective: every message goes to a direct neighbour (NCCL send/recv through
torch.distributed, or a device copy when the ranks share a process).  The result is bit-identical to
the single-GPU run (same kernels, same summation order, exact halo state).

Transports: `DistTransport` (one process per GPU, torch.distributed) and `LocalTransport` (all ranks in
one process on one device: used by the single-GPU tests of this protocol and by nothing else).
"""
from __future__ import annotations

import ctypes as C
from typing import List

from .api import Config, Simulation, SphbError

MSG_BASE, MSG_RECORDS, MSG_ITERATE = 0, 1, 2
NFLAGS = 6   # device error, order broken, halo changed, h max, thinnest halo, work lists busy (sphb_slab_poll)
LEFT, RIGHT = -1, 1


def partition(n: int, world: int) -> List[int]:
    """Slab boundaries: rank r owns [b[r], b[r+1]); sizes differ by at most one particle."""
    base, rem = divmod(n, world)
    b = [0]
    for r in range(world):
        b.append(b[-1] + base + (1 if r < rem else 0))
    return b


def rebalance_shifts(counts: List[int]) -> List[int]:
    """Particles to move across each internal edge so that all slabs hold (nearly) the same count
    again: shifts[e] > 0 moves that many particles from rank e to rank e + 1, < 0 the other way.
    Pure host arithmetic on the all-gathered counts (SURVEY.md §8(e) "rebalancing")."""
    world = len(counts)
    target = partition(sum(counts), world)
    shifts, have = [], 0
    for e in range(world - 1):
        have += counts[e]
        shifts.append(have - target[e + 1])
    return shifts


class SlabRank(Simulation):
    """One rank's handle plus its message buffers (torch tensors on the rank's device)."""

    def __init__(self, config: Config, rank: int, world: int, n_global: int, halo_count: int = 256,
                 device: int = -1, **kw):
        import torch
        b = partition(n_global, world)
        self.rank, self.world, self.n_global = rank, world, n_global
        self.i0, self.i1 = b[rank], b[rank + 1]
        local = Config(n_part=self.i1 - self.i0, mass=config.mass, pressure_calc=config.pressure_calc,
                       limit=config.limit, v_0=config.v_0, h_factor=config.h_factor, t_i=config.t_i)
        kw.setdefault("pad_capacity", max(4096, halo_count))
        super().__init__(local, device=device, **kw)
        rc = self.L.sphb_slab_configure(self.h, rank, world, halo_count, self.i0, n_global)
        self._check(rc)
        dev = torch.device("cuda", torch.cuda.current_device() if device < 0 else device)
        self.device = dev
        self.send, self.recv = {}, {}
        for what in (MSG_BASE, MSG_RECORDS, MSG_ITERATE):
            n = self.L.sphb_slab_message_doubles(self.h, what)
            for side in self.sides():
                self.send[(what, side)] = torch.zeros(n, dtype=torch.float64, device=dev)
                self.recv[(what, side)] = torch.zeros(n, dtype=torch.float64, device=dev)

    def sides(self):
        return [s for s in (LEFT, RIGHT) if 0 <= self.rank + s < self.world]

    def pack(self, what, side):
        self._check(self.L.sphb_slab_pack(self.h, what, side, C.c_void_p(self.send[(what, side)].data_ptr())))

    def unpack(self, what, side):
        self._check(self.L.sphb_slab_unpack(self.h, what, side, C.c_void_p(self.recv[(what, side)].data_ptr())))

    def call(self, name, *args):
        self._check(getattr(self.L, "sphb_slab_" + name)(self.h, *args))

    def poll(self):
        flags = (C.c_int * NFLAGS)()
        self._check(self.L.sphb_slab_poll(self.h, flags))
        return list(flags)

    def agree(self, flags_max):
        """The element-wise MAX over all ranks of the polled flags: halo-width verdict (the same on every rank) and
        the sweep strategy of the next step."""
        arr = (C.c_int * NFLAGS)(*[int(v) for v in flags_max])
        self._check(self.L.sphb_slab_agree(self.h, arr))


def _fields_by_index(ranks):
    """Records of the given ranks' handles: alive particles sorted by their global index (a slab's records
    travel in x order with the index in `id`, and particles may have changed slab), then the wall ghosts."""
    import numpy as np
    from .api import FIELDS
    alive, ghosts = [], []
    for r in ranks:
        rec, g = r.download()
        na = len(rec) - g
        alive.append(rec[:na].copy())
        ghosts.append(rec[na:].copy())
    a = np.concatenate(alive)
    a = a[np.argsort(a["id"], kind="stable")]
    gh = np.concatenate(ghosts)
    rec = np.concatenate([a, gh])
    f = np.stack([rec[name] for name in FIELDS], axis=1)
    t = rec["type"].astype(np.int32)
    return f, t, len(gh), a["id"].astype(np.int64)


class LocalTransport:
    """All ranks live in this process (one device): a message is a device-to-device copy."""

    def __init__(self, ranks: List[SlabRank]):
        self.ranks = {r.rank: r for r in ranks}

    def exchange(self, what):
        for r in self.ranks.values():
            for side in r.sides():
                # what I send to my right neighbour arrives as its message from the LEFT
                self.ranks[r.rank + side].recv[(what, -side)].copy_(r.send[(what, side)])

    def agree(self, flags):
        return list(flags)


def neighbour_exchange(dist, rank: int, world: int, send: dict, recv: dict):
    """One grouped send/recv with the direct neighbours: send[side] goes to rank + side and arrives
    there as recv[-side].  Tensors live wherever the process group's backend wants them (CUDA for
    nccl; CPU for the gloo tests of this routing)."""
    ops = []
    for side in (LEFT, RIGHT):
        peer = rank + side
        if 0 <= peer < world:
            ops.append(dist.P2POp(dist.isend, send[side], peer))
            ops.append(dist.P2POp(dist.irecv, recv[side], peer))
    if ops:
        for w in dist.batch_isend_irecv(ops):
            w.wait()


def agree_max(dist, flag_tensor, values):
    """Element-wise MAX over all ranks of a few small non-negative ints (one all-reduce)."""
    for k, v in enumerate(values):
        flag_tensor[k] = int(v)
    dist.all_reduce(flag_tensor, op=dist.ReduceOp.MAX)
    return [int(v) for v in flag_tensor.tolist()]


class DistTransport:
    """One process per GPU: NCCL send/recv between direct neighbours (torch.distributed)."""

    def __init__(self, rank_obj: SlabRank):
        import torch
        import torch.distributed as dist
        self.r, self.dist = rank_obj, dist
        self.flag = torch.zeros(NFLAGS, dtype=torch.int32, device=rank_obj.device)
        self.host_flag = torch.zeros(NFLAGS, dtype=torch.int32).pin_memory()

    def exchange(self, what):
        r = self.r
        neighbour_exchange(self.dist, r.rank, r.world, {s: r.send[(what, s)] for s in r.sides()},
                           {s: r.recv[(what, s)] for s in r.sides()})

    def agree(self, flags):
        self.host_flag[:] = self.host_flag.new_tensor([int(v) for v in flags])
        self.flag.copy_(self.host_flag, non_blocking=True)
        self.dist.all_reduce(self.flag, op=self.dist.ReduceOp.MAX)
        return [int(v) for v in self.flag.tolist()]


class SlabSimulation:
    """The protocol above over a set of ranks held by THIS process (one with DistTransport, all of them
    with LocalTransport).  Mirrors the Simulation calls the bench and the tests use."""

    def __init__(self, ranks: List[SlabRank], transport, max_rounds: int = 32, fixed_rounds: int = 1):
        import torch
        self.ranks, self.t, self.max_rounds, self.fixed_rounds = ranks, transport, max_rounds, fixed_rounds
        self.rounds_last = 0
        # every handle, the message copies and the NCCL calls are ordered on ONE stream
        self.stream = torch.cuda.Stream(device=ranks[0].device)
        for r in ranks:
            r.set_stream(self.stream.cuda_stream)
        self._torch = torch

    # ---- plumbing
    def _exchange(self, what):
        for r in self.ranks:
            for side in r.sides():
                r.pack(what, side)
        self.t.exchange(what)
        for r in self.ranks:
            for side in r.sides():
                r.unpack(what, side)

    def _all(self, name, *args):
        for r in self.ranks:
            r.call(name, *args)

    # ---- the path
    def init_lattice(self):
        """init_particles (setup.cpp:143-228) across the slabs."""
        with self._torch.cuda.stream(self.stream):
            self._init_lattice()

    def step(self, nsteps=1):
        with self._torch.cuda.stream(self.stream):
            self._step(nsteps)

    def init_developed(self, t_star=0.5):
        """sphb_init_developed across the slabs."""
        with self._torch.cuda.stream(self.stream):
            for r in self.ranks:
                r._check(r.L.sphb_slab_init_developed_fill(r.h, float(t_star)))
            for r in self.ranks:
                r.ghosts()
            self._exchange(MSG_BASE)
            for r in self.ranks:
                r.density(include_ghosts=True)
            self._all("records")
            self._exchange(MSG_RECORDS)
            self._all("forces")

    def _init_lattice(self):
        adiabatic = self.ranks[0].config.pressure_calc == 1
        self._all("init_lattice")
        if adiabatic:
            self._exchange(MSG_BASE)
            for r in self.ranks:
                r.density(include_ghosts=False)
            self._all("prepass")
        for r in self.ranks:
            r.ghosts()
        self._exchange(MSG_BASE)
        for r in self.ranks:
            r.density(include_ghosts=True)
        self._all("records")
        self._exchange(MSG_RECORDS)
        self._all("forces")

    def _attempt(self):
        """Everything of a step after the drift, up to the flags."""
        self._exchange(MSG_BASE)
        self._all("density")
        self._exchange(MSG_RECORDS)
        self._all("force_first")
        self._all("refine")
        # The first ITERATE exchange replaces the neighbours' GUESS by converged values, so it practically
        # always changes something: refine without asking.  The last exchange of the fixed schedule only
        # CHECKS (its unpack raises the halo_changed flag if anything still moved).
        for _ in range(self.fixed_rounds):
            self._exchange(MSG_ITERATE)
            self._all("refine")
        self._exchange(MSG_ITERATE)
        return self.fixed_rounds + 1

    def _flags(self):
        """MAX over all ranks of (device error, order broken, halo changed, largest h): the step's host sync."""
        fl = [0] * NFLAGS
        for r in self.ranks:
            f = r.poll()
            fl = [max(a, b) for a, b in zip(fl, f)]
        mine = fl[0]
        return self.t.agree(fl), mine

    def _step(self, nsteps=1):
        for _ in range(nsteps):
            self._all("step_begin")
            rounds = self._attempt()
            agreed, mine = self._flags()
            err, unsorted, changed = agreed[:3]
            for _repair in range(2):
                if not unsorted:
                    break
                # some rank's drift swapped neighbours: those ranks repair their order (a few transposition
                # steps first, the radix sort if that was not enough), everybody repeats the attempt
                # (nothing of the previous state has been overwritten)
                self._all("resort")
                rounds = self._attempt()
                agreed, mine = self._flags()
                err, unsorted, changed = agreed[:3]
            if unsorted:
                raise SphbError(11, "x-order still broken after the sort")
            while changed and not err:
                if rounds > self.max_rounds:
                    raise SphbError(10, "x-slab ordered sweep did not settle in %d rounds" % self.max_rounds)
                self._all("refine")
                self._exchange(MSG_ITERATE)
                rounds += 1
                agreed, mine = self._flags()
                err, unsorted, changed = agreed[:3]
            if err:
                for r in self.ranks:
                    r.call("raise")          # raises SphbError with the device's message on the rank that has it
                raise SphbError(err, "a neighbouring slab reported device error %d" % err)
            # the widest kernel support anywhere in the run must fit inside its thinnest halo: judged from the agreed
            # flags, so every rank raises (or not) together
            for r in self.ranks:
                r.agree(agreed)
            self.rounds_last = rounds
            self._all("step_end")

    # ---- pass-throughs used by bench.py / tests
    def set_profiling(self, on=True):
        for r in self.ranks:
            r.set_profiling(on)

    def phase_times(self, reset=False):
        return self.ranks[0].phase_times(reset)

    def counters(self):
        return self.ranks[0].counters()

    def fields(self):
        """([n, 9] doubles, type[n], n_ghost, global index[n_alive]) of the local ranks: alive particles in
        ascending global index, then the wall ghosts."""
        return _fields_by_index(self.ranks)

    def close(self):
        for r in self.ranks:
            r.close()


class NativeSlabSimulation:
    """One process per GPU; the whole protocol runs inside libsphb (sphb_slab_step: grouped
    ncclSend/ncclRecv on the handle's stream, one host synchronisation per step).  Python only
    distributes the NCCL unique id through the already initialised torch.distributed group."""

    def __init__(self, rank_obj: SlabRank):
        import torch
        import torch.distributed as dist
        self.ranks = [rank_obj]
        r = rank_obj
        self.stream = torch.cuda.Stream(device=r.device)
        r.set_stream(self.stream.cuda_stream)
        ident = [None]
        if r.rank == 0:
            buf = C.create_string_buffer(128)
            r._check(r.L.sphb_nccl_unique_id(buf))
            ident[0] = buf.raw
        dist.broadcast_object_list(ident, src=0)
        r._check(r.L.sphb_slab_comm_init(r.h, ident[0]))
        self._torch = torch

    def init_lattice(self):
        r = self.ranks[0]
        r._check(r.L.sphb_slab_init_particles(r.h))

    def init_developed(self, t_star=0.5):
        r = self.ranks[0]
        r._check(r.L.sphb_slab_init_developed(r.h, float(t_star)))

    def step(self, nsteps=1):
        r = self.ranks[0]
        r._check(r.L.sphb_slab_step(r.h, nsteps))

    @property
    def rounds_last(self):
        r = self.ranks[0]
        return r.L.sphb_slab_rounds_last(r.h)

    def set_profiling(self, on=True):
        self.ranks[0].set_profiling(on)

    def phase_times(self, reset=False):
        return self.ranks[0].phase_times(reset)

    def counters(self):
        return self.ranks[0].counters()

    def fields(self):
        return _fields_by_index(self.ranks)

    def close(self):
        self.ranks[0].close()


def make_simulation(config: Config, variable_h=True, fixed_h=0.2, device=-1, rank=0, world=1,
                    halo_count=256, native=True, **kw):
    """world == 1: one device-resident Simulation.  world > 1 (one process per GPU under
    torch.distributed): this rank's slab, exchanging halos inside libsphb (native) or through
    torch.distributed point-to-point calls."""
    if world == 1:
        return Simulation(config, variable_h=variable_h, fixed_h=fixed_h, device=device, **kw)
    r = SlabRank(config, rank, world, config.n_part, halo_count=halo_count, device=device,
                 variable_h=variable_h, fixed_h=fixed_h, **kw)
    if native:
        return NativeSlabSimulation(r)
    return SlabSimulation([r], DistTransport(r))


def make_local_slabs(config: Config, world: int, variable_h=True, fixed_h=0.2, device=-1, halo_count=256, **kw):
    """All `world` slabs in this process on one device (protocol tests on a single GPU)."""
    ranks = [SlabRank(config, r, world, config.n_part, halo_count=halo_count, device=device,
                      variable_h=variable_h, fixed_h=fixed_h, **kw) for r in range(world)]
    return SlabSimulation(ranks, LocalTransport(ranks))
