"""Where does the refinement phase spend its time?  Prints global sweeps per step and phase times."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from sph_b200 import Config, Simulation
n = int(sys.argv[1]) if len(sys.argv) > 1 else 8 * 1024 * 1024
cfg = Config(n_part=n, mass=4.0 / (n - 1), pressure_calc=1, t_i=0.005 * 100 / (n - 1))
sim = Simulation(cfg)
sim.init_lattice()
sim.step(5)
sim.set_profiling(True)
sim.phase_times(reset=True)
c0 = sim.counters()
sim.step(20)
c1 = sim.counters()
ph = sim.phase_times()
print("n=%d sweeps/step=%.2f tiles=%d evals=%d" % (n, (c1.sweeps_total - c0.sweeps_total) / 20.0, c1.refine_tiles, c1.refine_tile_evals), {k: round(v[0] / 20, 4) for k, v in ph.items()})
