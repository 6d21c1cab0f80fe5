#!/usr/bin/env python
"""bench.py — particle-steps/s of the per-timestep particle loop (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload NAME] [--impl reference]

A "step" is one pass of SPHSimulation::step_forward (sph_simulation.cpp:55-108) over the whole
particle set: kick-drift, ghosts, fused density + variable-h root-find + omega, fused
acceleration + du/dt with the reference's index-ordered sweep, second kick.  Inputs are the
synthetic colliding-stream lattice of setup.cpp:145-185 scaled to N,
generated on the device by sphb_init_lattice.

One JSON line on stdout (rank 0):
  value      whole-job particle-steps/s, state resident in HBM, CUDA events on the launching stream
  e2e        same metric through the host-facing C-ABI route with HOST buffers every step:
             sphb_upload (pinned AoS Particle[] -> device) + sphb_step(1) + sphb_download
  roofline   dominant kernel (the longer of the fused density pass and the first force sweep): algorithmic
             bytes per launch ÷ its CUDA-event duration, against MEASURED_PEAKS.json's HBM copy bandwidth;
             `traffic` = ncu dram bytes of that kernel (profiles/), next to the whole-step fractions
  cpu_baseline  the plain-C O(N k) port of the reference (oracle/) timed on one host core on a
             bounded sample of the same workload
--impl reference times the reference's own code (oracle/_ref, built from /root/reference by
oracle/Makefile) on the host cores instead; nothing of sph_b200 runs on that arm.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "particle_steps_per_sec"
UNIT = "particle-steps/s"

# name -> (N, pressure_calc, variable_h); h for fixed-h runs is 5 lattice spacings.
# Decimal particle counts are the default family: their lattice positions are not exactly equidistant in floating
# point, so every particle of the undisturbed streams feels a rounding-level kick and the index-ordered sweep has
# work EVERYWHERE, as it has in any developed flow.  With N = 2^k the positions are dyadic, the undisturbed streams
# feel |a| ~ 1e-13 and only the ~100 tiles around the collision front and the walls are re-swept: those workloads
# are kept as the labelled best case ("..._pow2").
WORKLOADS = {
    "colliding_streams_adiabatic_varh_n64m": (64_000_000, 1, True),           # BASELINE configs[3], the target line
    "colliding_streams_adiabatic_varh_n16m": (16_000_000, 1, True),           # configs[2]
    "colliding_streams_isothermal_fixedh_n1m": (1_000_000, 0, False),         # configs[1]
    "colliding_streams_adiabatic_varh_n256m": (256_000_000, 1, True),         # configs[4]: per GPU with --scaling weak
    "colliding_streams_adiabatic_varh_n1m": (1_000_000, 1, True),
    # synthetic developed start state (sphb_init_developed): the exact solution of the same problem at t* = 0.5 — wall
    # rarefactions, undisturbed streams, two shocks around compressed gas — sampled with equal-mass particles
    # (second start state of config 5); every particle is being accelerated from the first step
    "colliding_streams_adiabatic_varh_developed_n64m": (64_000_000, 1, True, "developed"),
    "colliding_streams_adiabatic_varh_developed_n16m": (16_000_000, 1, True, "developed"),
    "colliding_streams_adiabatic_varh_developed_n256m": (256_000_000, 1, True, "developed"),
    "colliding_streams_adiabatic_varh_n64m_pow2": (64 * 1024 * 1024, 1, True),    # best case of the ordered sweep
    "colliding_streams_adiabatic_varh_n16m_pow2": (16 * 1024 * 1024, 1, True),
    "colliding_streams_isothermal_fixedh_n1m_pow2": (1024 * 1024, 0, False),
}
DEFAULT_WORKLOAD = "colliding_streams_adiabatic_varh_n64m"
FP64_ROOF_TFLOPS = 37.2    # assumed: 148 SMs x 64 FP64 lanes x 2 flop x 1.965 GHz (MEASURED_PEAKS.json has no FP64 figure)


def workload_kwargs(n, pc, varh):
    return dict(n_part=n, mass=4.0 / (n - 1), pressure_calc=pc, limit=2.0, v_0=1.0, h_factor=2.0,
                t_i=0.005 * 100 / (n - 1), variable_h=varh, constant_h=0.2 * 100 / (n - 1))


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                 "-lms", "10"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thr = threading.Thread(target=self._pump, daemon=True)
            self.thr.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, pw = [], [], set(), []
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1])); pw.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": sorted(reasons)}


# ------------------------------------------------------------------------------------------
# CPU legs (the only places bench.py may execute oracle/)

def cpu_port_baseline(pc, varh, sample_n=600_001, steps=4, init="lattice"):
    """The O(N k) plain-C port (oracle/sph_oracle.c), one thread, same generator at sample_n.  For the developed
    workload the start state of the sample comes from the same generator as the GPU arm's (sphb_init_developed at
    sample_n, downloaded): the port is timed on it, it does not produce it."""
    from oracle.sph_oracle import Oracle
    kw = workload_kwargs(sample_n, pc, varh)
    if init == "developed":
        from sph_b200 import Config, Simulation
        g = Simulation(Config(n_part=sample_n, mass=kw["mass"], pressure_calc=pc, t_i=kw["t_i"]), variable_h=varh,
                       fixed_h=kw["constant_h"])
        g.init_developed(0.5)
        f, _, ng = g.fields()
        g.close()
        o = Oracle(**kw, init=False)
        o.set_state(f, ng)
    else:
        o = Oracle(**kw)
    t0 = time.perf_counter()
    o.step(steps)
    dt = time.perf_counter() - t0
    o.close()
    return {"value": sample_n * steps / dt, "unit": UNIT, "cores": 1, "kind": "port",
            "sample": "same %s at N=%d, %d steps, %.1f s; O(N k) port of the "
                      "reference (the literal reference is O(N^3) and cannot run at this N)" % (
                          "developed start state" if init == "developed" else "lattice generator", sample_n, steps, dt)}


def _ref_worker(args):
    n, pc, varh, steps, warm = args
    from oracle import ref_literal
    kw = workload_kwargs(n, pc, varh)
    r = ref_literal.RefLiteral(flavour="ofast", **kw)
    if warm:
        r.step(warm)
    t0 = time.perf_counter()
    r.step(steps)
    dt = time.perf_counter() - t0
    r.close()
    return dt


def run_reference(args, name, n_full, pc, varh):
    """The reference's own sources (oracle/_ref, -Ofast like its Makefile) on the host cores.  It is
    single-threaded and O(N^3) per step in variable-h mode (calculators.cpp:117), so a bounded sample of
    the workload is the same lattice at small N; one independent replica per host core is run so that
    every core the box has is used.  Falls back to the O(N k) port where oracle/_ref is absent."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    from oracle import ref_literal
    steps, warm = max(1, args.steps), max(0, args.warmup)
    cores = os.cpu_count() or 1
    if ref_literal.available(varh, "ofast"):
        import multiprocessing as mp
        n_s = 401 if varh else 3201          # ~0.3 s / step / core either way
        steps_s, warm_s = min(steps, 6), min(warm, 1)
        with mp.get_context("spawn").Pool(cores) as pool:
            t0 = time.perf_counter()
            dts = pool.map(_ref_worker, [(n_s, pc, varh, steps_s, warm_s)] * cores)
            wall = time.perf_counter() - t0
        dt = max(dts)
        value = cores * n_s * steps_s / dt
        single = n_s * steps_s / min(dts)
        kind = "reference"
        sample = ("reference sources compiled unmodified (-Ofast), %d independent single-threaded replicas (one per host core) of "
                  "the same lattice at N=%d, %d steps each, slowest replica %.2f s (wall %.1f s); value = sum over replicas, "
                  "single_replica_value = one of them; the reference step is O(N^%d), so its rate falls as N grows and N=%d "
                  "is the largest sample that keeps this arm within minutes" % (cores, n_s, steps_s, dt, wall, 3 if varh else 2, n_s))
        ms = dt / steps_s * 1e3
        steps_out = steps_s
        n_ran = n_s
    else:
        b = cpu_port_baseline(pc, varh)
        value, kind, sample, cores = b["value"], "port", b["sample"], 1
        single = value
        ms, steps_out = None, 3
        n_ran = 600_001
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": steps_out, "warmup": warm, "ms_per_step": ms, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": name, "n_particles": n_ran, "n_particles_of_the_gpu_arm": n_full,
                       "note": "the reference cannot run the GPU arm's N (O(N^3) per step): same generator, same physics, at the N named here"},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample,
                             "single_replica_value": single, "replicas": cores if kind == "reference" else 1},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)
    return 0


# ------------------------------------------------------------------------------------------

def slab_parity(args, world, rank, local, n=200_001, steps=6):
    """Evidence that the multi-GPU path computes the same thing: on the SAME ranks and NCCL transport as the timed run,
    a small problem is advanced as x-slabs and, on every rank, as one single-GPU handle; every field of every particle
    a rank owns must be bit-identical, and the all-reduced conservation totals must equal the single handle's."""
    import numpy as np
    import torch
    import torch.distributed as dist
    from sph_b200 import Config, Simulation
    from sph_b200.slabs import make_simulation
    cfg = Config(n_part=n, mass=4.0 / (n - 1), pressure_calc=1, t_i=0.005 * 100 / (n - 1))
    one = Simulation(cfg, device=local, sweep_mode=args.sweep_mode)
    one.init_lattice()
    sl = make_simulation(cfg, device=local, rank=rank, world=world, halo_count=128, native=not args.torch_exchange,
                         sweep_mode=args.sweep_mode)
    sl.init_lattice()
    bad = 0
    for st in range(steps + 1):
        if st:
            one.step(1)
            sl.step(1)
        if st in (0, 1, steps):
            f1 = one.fields()[0]
            f2, _, _, ids = sl.fields()
            bad += 0 if np.array_equal(f1[ids], f2[:len(ids)]) else 1
    d1 = one.diagnostics()
    d2 = sl.ranks[0].diagnostics()
    t = torch.tensor([float(bad), d2["mass"], d2["momentum"], d2["kinetic"], d2["thermal"]], device="cuda", dtype=torch.float64)
    dist.all_reduce(t)
    tot = dict(mass=float(t[1]), momentum=float(t[2]), kinetic=float(t[3]), thermal=float(t[4]))
    rel = max(abs(tot[k] - d1[k]) / max(abs(d1[k]), 1e-30) for k in ("mass", "kinetic", "thermal"))
    serial_steps = sl.counters().sweep_steps
    sl.close()
    one.close()
    ok = int(t[0].item()) == 0 and rel < 1e-12
    return {"verdict": "identical" if ok else "differs", "n_particles": n, "steps": steps, "ranks_with_a_difference": int(t[0].item()),
            "totals_all_reduced": tot, "totals_single_gpu": d1, "totals_max_rel_diff": rel, "serial_sweep_steps": serial_steps,
            "compared": "every field of every owned particle at steps 0, 1, %d, bit for bit, on each rank" % steps}


def run_ours(args, name, n, pc, varh):
    import numpy as np
    import torch
    import torch.distributed as dist
    from sph_b200 import PARTICLE_DTYPE, Config, Simulation

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — libsphb has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("bench.py --gpus %d must be launched under torch.distributed.run" % args.gpus)

    limit = 2.0
    if args.scaling == "weak":
        # The named N is per GPU (8 x 256,000,000 stays below 2^31, the index field), and the TUBE grows with the GPU
        # count (limit = 2 x GPUs): spacing, smoothing lengths and time step per GPU stay what they are on one GPU.
        # Packing 2e9 particles into the reference's 4 units of length is not an option: h = 2 spacings = 4e-9 would lie
        # below the reference's absolute CALC_EPSILON = 1e-8 (define.hpp:25), which its ghost set-up
        # (ghost_particles.cpp:37-43) and root-find bracket (smoothing_length.cpp:223-230,165) treat as "zero".
        n = n * world
        limit = 2.0 * world
    kw = workload_kwargs(n, pc, varh)
    kw["mass"] *= limit / 2.0
    kw["t_i"] *= limit / 2.0
    kw["constant_h"] *= limit / 2.0
    cfg = Config(n_part=n, mass=kw["mass"], pressure_calc=pc, limit=limit, t_i=kw["t_i"])
    from sph_b200.slabs import make_simulation   # 1 rank: a plain Simulation; >1: this rank's x-slab
    sim = make_simulation(cfg, variable_h=varh, fixed_h=kw["constant_h"], device=local, rank=rank, world=world,
                          halo_count=args.halo, native=not args.torch_exchange, sweep_rel_tol=args.sweep_tol,
                          sweep_mode=args.sweep_mode, sweep_group=args.sweep_group, sweep_segment=args.sweep_segment,
                          sweep_burn_in=args.sweep_burn_in)
    if world > 1:
        stream = sim.stream                      # handle, halo messages and NCCL calls share it
    else:
        stream = torch.cuda.Stream()
        sim.set_stream(stream.cuda_stream)
    if args.init == "developed":
        sim.init_developed(0.5)
    else:
        sim.init_lattice()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        sim.step(1)
    barrier()
    launches0 = sim.counters().launches
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    with torch.cuda.stream(stream):
        ev0.record(stream)
        for _ in range(args.steps):
            sim.step(1)
        ev1.record(stream)
    barrier()
    ms_total = ev0.elapsed_time(ev1)
    clocks = sampler.stop() if rank == 0 else None
    c = sim.counters()
    launches = c.launches - launches0
    if world > 1:
        t = torch.tensor([ms_total], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total = float(t.item())
        lt = torch.tensor([launches], device="cuda", dtype=torch.int64)
        dist.all_reduce(lt, op=dist.ReduceOp.SUM)
        launches = int(lt.item())
    ms_step = ms_total / args.steps
    value = n * args.steps / (ms_total * 1e-3)
    sweeps_mean = (c.sweeps_total / max(c.steps, 1))

    # ---- per-kernel durations (CUDA events on the launching stream, live, after the timed region)
    prof_steps = min(args.steps, 10)
    sim.set_profiling(True)
    sim.phase_times(reset=True)
    for _ in range(prof_steps):
        sim.step(1)
    barrier()
    phases = sim.phase_times(reset=True)
    sim.set_profiling(False)
    c2 = sim.counters()
    n_local = c2.n_alive
    if world > 1:     # statistics below are per rank; take the busiest phase times over ranks
        for k in list(phases):
            t = torch.tensor([phases[k][0]], device="cuda", dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            phases[k] = (float(t.item()), phases[k][1])
    hist = c2.newton_hist
    tot = max(sum(hist), 1)
    newton_mean = sum(k * v for k, v in enumerate(hist)) / tot if varh else 0.0
    nbr_rho = c2.nbr_mean
    nbr_f = max(c2.force_pairs_mean - 1.0, 0.0)
    # neighbour-window scans actually made per particle (1 + Newton iterations, minus the last one wherever the
    # final Newton step was below 1e-12 and the previous sums were reused)
    passes = c2.density_scans_mean if c2.density_scans_mean > 0 else ((1.0 + newton_mean) if varh else 1.0)
    # Bytes per particle-step.  SURVEY 8(d)'s figure B_ns counts every neighbour read ("bytes touched per particle x
    # neighbours"): density 56 + 8 B (x_j) per neighbour per pass of the root-find, force 124 + 56 B per pair, kick 64.
    # The kernels serve neighbours from shared memory, so what can actually cross the HBM interface is the
    # COMPULSORY traffic: density reads x, h, rho (24 B) and writes h, rho, the two force records, the gradient sum
    # and the packed counts (92 B) = 116 B; the force kernel reads x, idx, both records, v, u, gradient sum, stored
    # pressure (108 B) and writes a, du/dt, P, v, u (40 B) = 148 B; kick-drift 72 B.  roofline.frac uses those
    # (ncu's dram__bytes, `traffic`, agree with them); the B_ns figures are reported beside it, not as the fraction.
    bytes_density = (56.0 if varh else 24.0) + 8.0 * passes * nbr_rho
    rec = 56.0 if varh else 40.0
    bytes_force = (124.0 if pc == 1 else 92.0) + rec * nbr_f
    bytes_kick = 64.0
    b_ns = bytes_kick + bytes_density + bytes_force
    comp = {"density": 116.0, "force_first": 148.0, "kick_drift": 72.0}
    peak, peak_src = measured_peaks()
    d_ms = phases["density"][0] / max(phases["density"][1], 1)
    f_ms = phases["force_first"][0] / max(phases["force_first"][1], 1)
    r_ms = phases["force_refine"][0] / max(phases["force_refine"][1], 1)
    k_ms = phases["kick_drift"][0] / max(phases["kick_drift"][1], 1)
    step_prof_ms = sum(v[0] for v in phases.values()) / prof_steps
    dom = max((("density", d_ms, bytes_density), ("force_first", f_ms, bytes_force)), key=lambda t: t[1])
    serial = c2.sweep_group > 0
    kname = "k_density" if dom[0] == "density" else ("k_force_sweep" if serial else "k_force_first")
    achieved = n_local * comp[dom[0]] / (dom[1] * 1e-3) / 1e9 if dom[1] > 0 else 0.0
    achieved_b_ns = n_local * dom[2] / (dom[1] * 1e-3) / 1e9 if dom[1] > 0 else 0.0
    traffic = fp64_pct = None
    tsrc = None
    for tf in ("traffic_r02.json", "traffic_r01.json"):
        try:   # per launch, from the committed ncu --set full capture, scaled per particle to this launch
            tr = json.load(open(os.path.join(ROOT, "profiles", tf)))
            traffic = (tr[kname]["dram_bytes_read"] + tr[kname]["dram_bytes_write"]) / tr[kname].get("n_particles", tr.get("n_particles")) * n_local
            fp64_pct = tr[kname].get("fp64_pipe_pct")
            tsrc = "profiles/%s: %s" % (tf, tr.get("source", ""))
            break
        except Exception:
            continue
    roofline = {"bound": "hbm", "kernel": kname,
                "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                "traffic_source": tsrc, "peak_source": peak_src,
                "bytes_per_particle": comp[dom[0]], "launch_ms": dom[1],
                "bytes_basis": "compulsory HBM traffic of the kernel (inputs read once + outputs written once); "
                               "neighbour reads are served from shared memory",
                "fp64_pipe_frac": (fp64_pct / 100.0) if fp64_pct is not None else None,
                "fp64_roof_tflops_assumed": FP64_ROOF_TFLOPS,
                "note": "the kernel sits between the roofs: it moves its compulsory bytes once (traffic ~ bytes_per_particle x N) and "
                        "is limited by dependent FP64 chains / instruction issue, not by HBM (profiles/r02_summary.md)",
                "survey_b_ns": {"bytes_per_particle_kernel": dom[2], "achieved_kernel_gbs": achieved_b_ns,
                                "frac_kernel": achieved_b_ns / peak,
                                "bytes_per_particle_step": b_ns, "frac_step": value / world * b_ns / 1e9 / peak,
                                "note": "SURVEY 8(d) formula: counts every neighbour read as HBM traffic, can exceed 1; not the roofline fraction"},
                "step_frac_compulsory": value / world * ((116.0 + 148.0 + 72.0)) / 1e9 / peak,
                "phase_ms": {"kick_drift": k_ms, "density": d_ms, "force_first": f_ms, "force_refine": r_ms,
                             "sort": phases["sort"][0] / prof_steps, "ghosts": phases["ghosts"][0] / prof_steps,
                             "step_sum": step_prof_ms},
                "kernel_share_of_step": dom[1] / step_prof_ms if step_prof_ms > 0 else None}

    # ---- end to end through the C-ABI with HOST buffers, every step: upload + step + download
    e2e = None
    if args.e2e_steps > 0:
        # every rank moves ITS particles (the whole set on one GPU, its slab otherwise) through host memory
        e2e_steps = max(1, min(args.steps, args.e2e_steps))
        hs = sim if world == 1 else sim.ranks[0]     # the object that owns this rank's handle
        n_loc = n_local
        nbuf = n_loc + 2 * 4096
        try:
            host = torch.empty(nbuf * PARTICLE_DTYPE.itemsize, dtype=torch.uint8, pin_memory=True)
        except RuntimeError:     # the box cannot pin that much host memory: pageable buffers (slower copies)
            host = torch.empty(nbuf * PARTICLE_DTYPE.itemsize, dtype=torch.uint8)
        rec_arr = host.numpy().view(PARTICLE_DTYPE)
        out, ng = hs.download(rec_arr)
        barrier()
        h2d = d2h = 0
        with torch.cuda.stream(stream):
            ev0.record(stream)
            for _ in range(e2e_steps):
                npart = len(out)
                hs.upload(rec_arr[:npart], ng)
                h2d += npart * PARTICLE_DTYPE.itemsize
                sim.step(1)
                out, ng = hs.download(rec_arr)
                d2h += len(out) * PARTICLE_DTYPE.itemsize
            ev1.record(stream)
        barrier()
        e2e_ms = ev0.elapsed_time(ev1)
        if world > 1:
            t = torch.tensor([e2e_ms, float(h2d), float(d2h)], device="cuda", dtype=torch.float64)
            tmax = t.clone()
            dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
            e2e_ms, h2d, d2h = float(tmax[0].item()), int(t[1].item()), int(t[2].item())
        e2e = {"value": n * e2e_steps / (e2e_ms * 1e-3), "unit": UNIT, "steps": e2e_steps,
               "h2d_bytes_per_step": h2d // e2e_steps, "d2h_bytes_per_step": d2h // e2e_steps,
               "ms_per_step": e2e_ms / e2e_steps,
               "route": "sphb_upload(pinned Particle[]) + sphb_step(1) + sphb_download(Particle[]), all ranks"}
    rounds_last = getattr(sim, 'rounds_last', 0)
    sim.close()

    parity = slab_parity(args, world, rank, local) if world > 1 else None

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu = cpu_port_baseline(pc, varh, init=args.init)

    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank != 0:
        return 0
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": args.scaling,
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": name, "n_particles": n, "pressure_calc": "adiabatic" if pc == 1 else "isothermal",
                       "smoothing_length": "variable (Newton/bisection per particle)" if varh else "fixed, 5 spacings",
                       "start_state": ("synthetic developed state: exact colliding-streams solution at t* = 0.5, equal-mass sampling "
                                       "(sphb_init_developed)" if args.init == "developed" else "t = 0 lattice of init_particles (setup.cpp:143-228)"),
                       "lattice": ("N = 2^k: dyadic positions, undisturbed streams feel no kick: best case of the ordered sweep"
                                   if (n // world if args.scaling == "weak" else n) & ((n // world if args.scaling == "weak" else n) - 1) == 0
                                   else "generic N: every particle feels a rounding-level kick, the ordered sweep has work everywhere"),
                       "order": ("reference (index-ordered in-place sweep, bitwise fixed point)" if args.sweep_tol == 0 else
                                 "reference (index-ordered in-place sweep, converged to %.1e relative)" % args.sweep_tol), "sweeps_per_step": sweeps_mean,
                       "n_particles_per_gpu": n // world, "limit": limit,
                       "parallelism": ("x-slabs x%d, %d-particle halos over ncclSend/ncclRecv, %d ITERATE exchanges in the last step"
                                       % (world, args.halo, rounds_last)) if world > 1 else "single GPU",
                       "sweep": {"serial_steps": c2.sweep_steps, "group": c2.sweep_group, "segment": c2.sweep_segment,
                                 "burn_in": c2.sweep_burn_in, "flagged": c2.sweep_flagged, "verify_fail": c2.sweep_verify_fail,
                                 "refine_tiles": c2.refine_tiles, "refine_tile_evals": c2.refine_tile_evals},
                       "l2": "state arrays (%.1f GB) larger than L2, no flush needed" % (n * 8 * 20 / 1e9)},
            "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches), "roofline": roofline, "cpu_baseline": cpu}
    if parity is not None:
        line["slab_parity"] = parity["verdict"]
        line["slab_parity_detail"] = parity
    print(json.dumps(line), flush=True)
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD, choices=sorted(WORKLOADS))
    ap.add_argument("--n", type=int, default=0, help="override the particle count of the workload")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--e2e-steps", type=int, default=3, help="steps of the end-to-end (host buffer) leg; 0 skips it")
    ap.add_argument("--sweep-tol", type=float, default=0.0,
                    help="0 (default): the ordered sweep is iterated to its bitwise fixed point; > 0: relative tolerance")
    ap.add_argument("--scaling", default="strong", choices=["strong", "weak"],
                    help="--gpus N > 1: strong = the named N in total, weak = the named N per GPU")
    ap.add_argument("--sweep-mode", type=int, default=0, help="ordered sweep: 0 auto, 1 work lists, 2 serial sweep")
    ap.add_argument("--sweep-group", type=int, default=0)
    ap.add_argument("--sweep-segment", type=int, default=0)
    ap.add_argument("--sweep-burn-in", type=int, default=0)
    ap.add_argument("--torch-exchange", action="store_true",
                    help="x-slab runs: drive the halo exchange from Python (torch.distributed) instead of inside libsphb")
    ap.add_argument("--halo", type=int, default=256, help="particles exchanged per slab edge (x-slab runs)")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "ours":
        args.warmup = 3
    n, pc, varh = WORKLOADS[args.workload][:3]
    args.init = WORKLOADS[args.workload][3] if len(WORKLOADS[args.workload]) > 3 else "lattice"
    name = args.workload
    if args.n:
        n = args.n
        name += "(n=%d)" % n
    if args.impl == "reference":
        return run_reference(args, name, n, pc, varh)
    return run_ours(args, name, n, pc, varh)


if __name__ == "__main__":
    sys.exit(main())
