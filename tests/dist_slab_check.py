"""Run under torch.distributed.run (one process per GPU): every rank advances its x-slab over NCCL
halo exchanges and compares it, bit for bit, with a single-handle run of the whole problem that it
computes on its own GPU.  Exit code 0 = identical.  Started by tests/test_slabs.py when >= 2 GPUs
are visible, and by hand:  python -m torch.distributed.run --nproc-per-node 2 tests/dist_slab_check.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist
    from sph_b200 import Config, Simulation
    from sph_b200.slabs import make_simulation
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    n = int(os.environ.get("SLAB_CHECK_N", "200001"))
    cfg = Config(n_part=n, mass=4.0 / (n - 1), pressure_calc=1, t_i=0.005 * 100 / (n - 1))
    mode = int(os.environ.get("SLAB_SWEEP", "0"))      # 0 auto, 1 work lists, 2 serial sweep
    one = Simulation(cfg, device=local, sweep_mode=mode)
    one.init_lattice()
    native = os.environ.get("SLAB_NATIVE", "1") == "1"
    sim = make_simulation(cfg, device=local, rank=rank, world=world, halo_count=128, native=native, sweep_mode=mode)
    sim.init_lattice()
    bad = 0
    for st in range(0, 7):
        if st:
            one.step(1)
            sim.step(1)
        if st in (0, 1, 3, 6):
            f1 = one.fields()[0]
            f2, _, g2, ids = sim.fields()
            r = sim.ranks[0]
            na = len(ids)
            same = np.array_equal(f1[ids], f2[:na])
            print("rank %d step %d: slab [%d, %d) %s, rounds %d" % (rank, st, r.i0, r.i1,
                                                                   "identical" if same else "DIFFERS", sim.rounds_last),
                  "(native)" if native else "(torch.distributed)", flush=True)
            bad += 0 if same else 1
    t = torch.tensor([bad], device="cuda")
    dist.all_reduce(t)
    sim.close()
    one.close()
    dist.barrier()
    dist.destroy_process_group()
    return 1 if int(t.item()) else 0


if __name__ == "__main__":
    sys.exit(main())
