import sys, os
sys.path.insert(0, '/root/repo')
import torch
from sph_b200 import Config, Simulation
n = 16*1024*1024
cfg = Config(n_part=n, mass=4.0 / (n - 1), pressure_calc=1, t_i=0.005 * 100 / (n - 1))
sim = Simulation(cfg)
sim.init_lattice()
sim.set_profiling(True)
sim.step(95)
for st in range(95, 150):
    sim.phase_times(reset=True)
    sim.step(1)
    c = sim.counters(); ph = sim.phase_times()
    if st % 3 == 0 or ph["density"][0] > 0.7:
        print(st, "density %.3f scans %.3f newton_hist %s hmax %.6e" % (ph["density"][0], c.density_scans_mean, c.newton_hist[:4], c.h_max))
