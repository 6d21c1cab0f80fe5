"""Per-step cost in a developed flow: run N particles to t = t_end and print phase times along the way."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from sph_b200 import Config, Simulation
n = int(sys.argv[1]) if len(sys.argv) > 1 else 200001
t_end = float(sys.argv[2]) if len(sys.argv) > 2 else 0.5
dt = 0.005 * 100 / (n - 1)
cfg = Config(n_part=n, mass=4.0 / (n - 1), pressure_calc=1, t_i=dt)
sim = Simulation(cfg)
sim.init_lattice()
sim.set_profiling(True)
nsteps = int(round(t_end / dt))
blocks = 10
per = nsteps // blocks
for b in range(blocks):
    sim.phase_times(reset=True)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    sim.step(per)
    torch.cuda.synchronize(); w = (time.perf_counter() - t0) / per * 1e3
    c = sim.counters(); ph = sim.phase_times()
    f = sim.fields()[0][:n]
    h = f[:, 4]
    print("t=%.3f  %.4f ms/step  first %.4f refine %.4f density %.4f | h ratio %.2f pairs %.2f scans %.2f tiles %d evals %d sorts %d" % (
        (b + 1) * per * dt, w, ph["force_first"][0] / per, ph["force_refine"][0] / per, ph["density"][0] / per,
        h.max() / h.min(), c.force_pairs_mean, c.density_scans_mean, c.refine_tiles, c.refine_tile_evals, c.sorts), flush=True)
