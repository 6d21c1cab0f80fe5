"""ms/step and refinement statistics as the run progresses (positions lose the lattice's dyadic structure)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from sph_b200 import Config, Simulation
n = int(sys.argv[1]) if len(sys.argv) > 1 else 64 * 1024 * 1024
blocks = int(sys.argv[2]) if len(sys.argv) > 2 else 8
tol = float(sys.argv[3]) if len(sys.argv) > 3 else 0.0
cfg = Config(n_part=n, mass=4.0 / (n - 1), pressure_calc=1, t_i=0.005 * 100 / (n - 1))
sim = Simulation(cfg, sweep_rel_tol=tol)
sim.init_lattice()
sim.set_profiling(True)
for b in range(blocks):
    sim.phase_times(reset=True)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    sim.step(20)
    torch.cuda.synchronize(); dt = (time.perf_counter() - t0) / 20 * 1e3
    c = sim.counters()
    ph = sim.phase_times()
    print("steps %3d-%3d: %.2f ms/step  first %.2f refine %.2f density %.2f | tiles swept %d evals %d sweeps %d" % (
        b * 20, b * 20 + 20, dt, ph["force_first"][0] / 20, ph["force_refine"][0] / 20, ph["density"][0] / 20,
        c.refine_tiles, c.refine_tile_evals, c.sweeps_last), "pairs %.2f scans %.2f nbr %.2f" % (c.force_pairs_mean, c.density_scans_mean, c.nbr_mean), flush=True)
