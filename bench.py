#!/usr/bin/env python
"""bench.py -- headline measurement of the Mini-PIC time-step path on B200 (contract in the task statement).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl native|reference] [--workload NAME]

Native arm: one "step" = one full PIC iteration (gather+Boris push+deposit fused with the sort's reorder; half-B, E with
-dt*J, half-B in ONE fused kernel; per-cell count/scan/permutation sort) over the synthetic workload, state resident in HBM.
N=1 workload = BASELINE.json configs[1] (two-stream, 512x512 cells, 64 ppc = 16.8 M particles, 0.87 GB of particle state
>> 126 MB L2), timed in the regime the 2000-step run lives in: the plasma is first advanced, untimed, into saturation.
N>1: weak scaling, every rank owns a 512-row slab of a (512N x 512) domain (halos, J reduce and particle migration over
NVLink peer memory).  Every line also carries the sub-records `c4_slab` (configs[3]: 4096x512 cells, 134 M particles per
GPU -- at N=8 this IS the 1.07 B-particle config) and `yee` / `yee_8192` (configs[2], 8192^2 field-only, strong scaling).
Prints ONE JSON line on rank 0.

Reference arm (--impl reference): the reference has no particle code, so the timed CPU implementation is
the repo's scalar FP64 oracle port with OpenMP over particles ("kind": "port") on the SAME configuration, all host cores,
for exactly the --warmup/--steps it is given (bounded sample only when the state would not fit the host; stated).
"""
import argparse
import atexit
import contextlib
import importlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
PKG = "mini-pic--2d-pic-code-with-tiling-load-balance_b200"

WORKLOADS = {
    # BASELINE.json configs[1]; loader parameters from SURVEY.md section 8(d) C2
    "two_stream_512": dict(nx=512, ny=512, tn=16, ppc=64, u_th=0.001, two_stream=0.2, dx=0.1, saturate_steps=1500),
    # configs[0] (the "reference default" that exists only in BASELINE.json): thermal 128^2, 16 ppc
    "thermal_128": dict(nx=128, ny=128, tn=16, ppc=16, u_th=0.05, two_stream=0.0, dx=0.1),
    # configs[3] per-GPU share at 8 GPUs
    "uniform_2048": dict(nx=2048, ny=2048, tn=16, ppc=64, u_th=0.05, two_stream=0.0, dx=0.1),
    # configs[3] (4096^2, 64 ppc, 1.07 B particles) per-GPU slab at 8 GPUs: 4096 x 512 cells, 134 M particles per rank
    "uniform_c4_slab": dict(nx=4096, ny=512, tn=16, ppc=64, u_th=0.05, two_stream=0.0, dx=0.1),
    # configs[3] whole: 4096^2 x 64 ppc = 1.07 B particles, STRONG scaling (the domain is fixed, every rank owns 4096/N rows)
    "uniform_c4": dict(nx=4096, ny=4096, tn=16, ppc=64, u_th=0.05, two_stream=0.0, dx=0.1, strong=True),
    # configs[2]: vacuum pulse, no particles; STRONG scaling over N slabs (halo rows stored into the neighbours by the sweeps)
    "yee_8192": dict(nx=8192, ny=8192, tn=32, ppc=0, u_th=0.0, two_stream=0.0, dx=10.0 / 8192, fields_only=True, strong=True),
    # configs[4] per-GPU slab: thermal background 8 ppc + drifting beam 80 ppc in the band y in [0.45,0.55) of the GLOBAL box
    "beam_c5_slab": dict(nx=4096, ny=512, tn=16, ppc=8, u_th=0.05, two_stream=0.0, dx=0.1, beam_ppc=80, beam_band=(0.45, 0.55), beam_drift=0.5, beam_u_th=0.01),
}
# Algorithmic bytes.  SURVEY.md 8(d): fused gather+push+deposit reads x,y,ux,uy,uz,q (48 B) and
# writes x,y,ux,uy,uz (40 B) = 88 B, + the per-tile grid term 2.3 B/particle at 64 ppc => 90 B/particle: `roofline.frac`.
# The kernel here is ALSO the sort's reorder (it gathers through the slot->storage permutation and writes the other buffer
# in sorted order): one full read + one full write of x,y,ux,uy,uz,q,tag = 104 B: `frac_104B`.
PUSH_BYTES_SURVEY = 90.0
PUSH_BYTES_FUSED_REORDER = 104.0
# constant, NOT measured by the run that prints it: dram__bytes_read+write of one push launch in the saturated regime,
# ncu --set full, profiles/r02an_ncu_full_push_packed_saturated_raw.csv: 1.094 GB + 0.872 GB over 16 777 216 particles
NCU_TRAFFIC_BYTES_PER_PARTICLE = 117.2
# ncu --set full, profiles/r02an_ncu_full_fused_yee_8192_raw.csv: 3.238 GB + 3.179 GB over 8192^2 cells (algorithmic: 96 B/cell)
NCU_YEE_TRAFFIC_BYTES_PER_CELL = 95.6
YEE_BYTES_3SWEEP = 216.0           # reference-faithful half-B, E, half-B: 3 x 72 B per cell
YEE_BYTES_ALGO = 144.0             # SURVEY.md section 8(d) algorithmic figure (2 sweeps)
YEE_BYTES_FUSED = 96.0             # fused step: every plane read once, written once


def load_workload(sim, w):
    """Populate `sim` with the workload's plasma (device loader, partition-independent: counter-based RNG keyed by tag)."""
    sim.load_uniform(w["ppc"], u_th=w["u_th"], two_stream=w["two_stream"], seed=12345)
    if "beam_ppc" in w:      # configs[4]: dense drifting beam confined to a band of the GLOBAL box (same macro-weight as the background)
        sim.load_uniform(w["beam_ppc"], u_th=w["beam_u_th"], drift=(w["beam_drift"], 0.0, 0.0), band=tuple(w["beam_band"]),
                         density=float(w["beam_ppc"]) / w["ppc"], seed=777, append=True)


def expected_particles(w, ny_global, cy_lo, cy_hi):
    """Particles the loader puts into global cell rows [cy_lo, cy_hi)."""
    n = w["nx"] * (cy_hi - cy_lo) * w["ppc"]
    if "beam_ppc" in w:
        lo, hi = int(round(w["beam_band"][0] * ny_global)), int(round(w["beam_band"][1] * ny_global))
        n += w["nx"] * max(0, min(hi, cy_hi) - max(lo, cy_lo)) * w["beam_ppc"]
    return n


def workload_config(name, w, world=1):
    """The `config` object of the JSON line.  Built by this one function for BOTH arms, so the two lines name the same
    configuration key for key."""
    strong = bool(w.get("strong"))
    ny = w["ny"] if strong else w["ny"] * world
    dx = w["dx"]
    dt = 0.5 / np.sqrt(2.0 / (dx * dx))
    c = {"workload": name}
    c.update({k: (list(v) if isinstance(v, tuple) else v) for k, v in w.items()})
    c.update({"ny_global": ny, "particles": int(expected_particles(w, ny, 0, ny)), "tile": f"{w['tn']}x{w['tn']}", "dt": float(dt)})
    if w.get("saturate_steps"):
        c["regime"] = (f"saturated two-stream: the GPU arm advances the plasma {w['saturate_steps']} untimed steps before the warm-up and timed "
                       "steps; the CPU arm times the same configuration from the loader's quiet start (the scalar port's cost per "
                       "push does not depend on the cell-crossing rate)")
    if not w.get("fields_only"):
        c["l2"] = "per-GPU particle state >> 126 MB L2 (no flush needed)"
    else:
        c["l2"] = "48 B/cell x 67 M cells = 3.2 GB >> 126 MB L2"
    return c


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks + throttle reasons DURING the timed region.  ONE nvidia-smi process per bench run, started before any
    GPU work (its start-up on a fresh box takes longer than a 13 ms timed region), every line stamped on arrival; a
    `with ClockSampler(i) as clk:` block marks a window and `clk.summary()` reports the samples that fell inside it (widened
    to the nearest samples, and saying so, when the window was shorter than the 100 ms sampling period)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    _shared = {}

    def __init__(self, index):
        self.index = index
        self.t0 = self.t1 = None
        self.log = ClockSampler._shared.get(index)
        if self.log is None:
            self.log = ClockSampler._shared[index] = {"rows": [], "proc": None}
            try:
                proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                                         "-i", str(index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
                self.log["proc"] = proc
                atexit.register(ClockSampler.stop_all)          # the sampler must never outlive the bench process
                threading.Thread(target=self._read, args=(proc, self.log["rows"]), daemon=True).start()
            except Exception:
                pass

    @staticmethod
    def _read(proc, rows):
        for line in proc.stdout:
            rows.append((time.time(), [c.strip() for c in line.split(",")]))

    def __enter__(self):
        self.t0 = time.time()
        return self

    def __exit__(self, *a):
        self.t1 = time.time()
        if self.log["proc"]:
            t = time.time()
            while not any(ts >= self.t1 for ts, _ in self.log["rows"]) and time.time() - t < 3.0:   # one sample past the window
                time.sleep(0.05)

    @classmethod
    def stop_all(cls):
        for log in cls._shared.values():
            if log["proc"]:
                log["proc"].terminate()
        cls._shared.clear()

    def summary(self):
        rows = [(ts, r) for ts, r in list(self.log["rows"]) if len(r) >= 6 and r[0].replace(".", "").isdigit()]
        inside = [r for ts, r in rows if self.t0 <= ts <= self.t1]
        how = "inside the window"
        if not inside and rows:                   # window shorter than the sampling period: the samples on either side of it
            before = [r for ts, r in rows if ts < self.t0][-1:]
            after = [r for ts, r in rows if ts > self.t1][:1]
            inside = before + after
            how = "window shorter than the 100 ms sampling period: nearest sample on either side"
        sm = [float(r[0]) for r in inside]
        mx = [float(r[1]) for r in inside if r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[k] for r in inside for k in range(4) if r[2 + k].lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(sm), "sampled": how, "window_s": round(self.t1 - self.t0, 3)}


# ------------------------------------------------------------------------------------------------ CPU legs
def cpu_threads():
    """All host cores, whatever a launcher exported (torchrun sets OMP_NUM_THREADS=1)."""
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def cpu_port_run(w, nx, ny, steps, warmup=1, omp=True):
    """The oracle port (scalar FP64 C, OpenMP over particles when omp) on an nx x ny piece of workload `w`: `warmup`
    untimed + `steps` timed full PIC steps.  Returns (pushes/s, threads used, particles, seconds)."""
    from oracle import bindings as ob
    plasma = importlib.import_module(PKG + ".plasma")
    m = importlib.import_module(PKG)
    lib = ob.load_oracle(omp=omp)
    threads = 1
    if omp:
        lib.orc_set_threads(cpu_threads())
        threads = int(lib.orc_num_threads())
    dx = dy = w["dx"]
    dt = m.courant_dt(dx, dy)
    P = ob.make_params(lib, nx, ny, w["tn"], w["tn"], dx, dy, dt)
    parts = plasma.load_uniform(nx, ny, w["tn"], w["tn"], dx, dy, w["ppc"], u_th=w["u_th"], two_stream=w["two_stream"])
    planes = [np.zeros((ny, nx)) for _ in range(6)]
    if warmup > 0:
        ob.pic_steps(lib, P, planes, parts, warmup)             # page faults, thread pool
    t0 = time.perf_counter()
    ob.pic_steps(lib, P, planes, parts, steps)
    dtw = time.perf_counter() - t0
    n = parts["x"].shape[0]
    return n * steps / dtw, threads, n, dtw


def cpu_baseline_pic(w, cells=512, steps=3):
    """`cpu_baseline` of the native line: bounded sample of the same workload on all host cores + a single-core figure."""
    nx = min(cells, w["nx"])
    ny = min(cells, w["ny"])
    v, threads, n, sec = cpu_port_run(w, nx, ny, steps)
    v1, _, n1, _ = cpu_port_run(w, min(128, nx), min(128, ny), 1, omp=False)
    return {"value": v, "unit": "particles/s", "cores": threads, "kind": "port",
            "sample": f"{nx}x{ny} cells x {w['ppc']} ppc = {n} particles, {steps} full PIC steps (push+deposit+Yee+sort) after 1 warm-up step, {sec:.1f} s",
            "value_single_core": v1, "sample_single_core": f"{min(128, nx)}x{min(128, ny)} cells = {n1} particles, 1 step",
            "note": "reference has no particle code (SURVEY.md F1); this is the repo's scalar FP64 oracle with OpenMP"}


def yee_cpu_reference(n=1800, tile=25, steps=10):
    """The reference's OWN field path on the host: unmodified Field_update.cpp + Auxiliar_functions.cpp pack/unpack driven
    by the reference loop structure (phases A-F with its Grid<->split copies; oracle/ref_wrap.cpp only calls reference
    functions), single rank through the mpi.h shim, on a bounded n x n sample.  -O2 build and the documented -O0 build."""
    from oracle import bindings as ob
    out = {}
    dx = 10.0 / n
    dt = 0.5 * dx / np.sqrt(2.0)
    x = (np.arange(n) + 0.5) * dx
    prof = 0.1 * np.sin(np.pi * x) * np.cos((x - 3.5) / 3.0 * np.pi / 2) ** 2 * (np.abs(x - 3.5) < 3.0)
    g6 = np.zeros((n, n, 6))
    g6[:, :, 1] = prof
    g6[:, :, 5] = prof
    for name, k in (("libref.so", steps), ("libref_O0.so", max(1, steps // 5))):
        if not ob.have_ref(name):
            continue
        ref = ob.load_ref(name)
        r = ob.ref_run_loop(ref, g6, 1, 1, n // tile, n // tile, 2, dt, dx, dx, k, want_tiles=False)
        out[name] = float(n) * n * k / r["seconds"]
    if not out:
        return None
    res = {"value": out.get("libref.so"), "unit": "cells/s", "cores": 1, "kind": "reference",
           "sample": f"{n}x{n} cells in {tile}x{tile} tiles (the reference's tile size), {steps} steps, 1 MPI rank (shim), g++ -O2",
           "value_O0_documented_build": out.get("libref_O0.so")}
    if ob.have_ref("libref_omp.so"):
        # all host cores: the same reference functions, the loops over tiles run with OpenMP (oracle/ref_wrap.cpp).  NOT the
        # reference's own parallelisation (MPI ranks, unavailable here: no mpirun in the image) -- labelled as such.
        ref = ob.load_ref("libref_omp.so")
        if hasattr(ref, "ref_set_threads"):
            ref.ref_set_threads(cpu_threads())
        r = ob.ref_run_loop(ref, g6, 1, 1, n // tile, n // tile, 2, dt, dx, dx, steps, want_tiles=False)
        res["omp_over_tiles"] = {"value": float(n) * n * steps / r["seconds"], "cores": int(ref.ref_num_threads()),
                                 "note": "reference functions, OpenMP over tiles (not the reference's MPI decomposition)"}
    return res


def run_reference(args):
    """--impl reference: rank 0 alone times the CPU implementation of the path on the arm's own config; the line's
    steps / warmup are the ones actually run."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    os.environ["OMP_NUM_THREADS"] = str(cpu_threads())          # before libgomp starts: torchrun exports OMP_NUM_THREADS=1
    w = WORKLOADS[args.workload]
    world = args.gpus
    cfg = workload_config(args.workload, w, world)
    K, W = max(1, args.steps), max(1, args.warmup)
    if w.get("fields_only"):
        # the reference's OWN field path (unmodified Field_update.cpp + pack/unpack under its loop structure), one rank:
        # 17 M cell-updates/s, so the 67 M-cell config costs 4 s per step -- bounded sample of the same pulse, stated
        cb = yee_cpu_reference(steps=min(K, 10))
        v = cb["value"]
        print(json.dumps({
            "impl": "reference", "metric": "Yee cell-updates/s", "value": v, "unit": "cells/s", "n_gpus": args.gpus,
            "steps": min(K, 10), "warmup": 0, "ms_per_step": 1e3 * 1800.0 * 1800.0 / v, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": cfg,
            "timed": "bounded sample of the same pulse: " + cb["sample"],
            "cpu_baseline": cb, "e2e": {"value": v, "unit": "cells/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}))
        return
    nx, ny = w["nx"], cfg["ny_global"]
    full = cfg["particles"]
    sample = "the full configuration"
    cap = 160_000_000                       # host image + the port's work arrays stay well inside the box's RAM
    if full > cap:
        ny = max(w["tn"], (cap // (w["nx"] * w["ppc"])) // w["tn"] * w["tn"])
        sample = f"bounded sample: {nx}x{ny} cells of the {nx}x{cfg['ny_global']} domain (same ppc, same loader)"
    if "beam_ppc" in w:
        sample += "; background population only (the port's loader has no band option)"
    v, threads, n, sec = cpu_port_run(w, nx, ny, K, warmup=W)
    print(json.dumps({
        "impl": "reference", "metric": "particle-pushes/s (push+deposit+sort)", "value": v, "unit": "particles/s",
        "n_gpus": args.gpus, "steps": K, "warmup": W, "ms_per_step": 1e3 * sec / K,
        "higher_is_better": True, "scaling": "strong" if w.get("strong") else "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic", "config": cfg,
        "timed": f"{sample}: {n} particles, {W} warm-up + {K} timed full PIC steps from the loader's quiet start, {threads} threads",
        "note": "reference has no particle code (SURVEY.md F1); timed = the repo's scalar FP64 oracle port, OpenMP over particles",
        "cpu_baseline": {"value": v, "unit": "particles/s", "cores": threads, "kind": "port",
                         "sample": f"{sample}: {nx}x{ny} cells x {w['ppc']} ppc = {n} particles, {K} full PIC steps, {sec:.1f} s"},
        "e2e": {"value": v, "unit": "particles/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}))


# ------------------------------------------------------------------------------------------------ native, one GPU
def yee_bench(m, torch, steps, warmup, n=8192):
    """Field-only sub-benchmark (BASELINE.json configs[2]): n x n vacuum pulse.  `value` = the fused one-kernel step
    (bit-identical to the reference-faithful three sweeps, which are timed beside it)."""
    dx = dy = 10.0 / n
    sim = m.Sim(n, n, 32, 32, dx, dy, m.courant_dt(dx, dy))
    x = (np.arange(n) + 0.5) * dx
    prof = 0.1 * np.sin(np.pi * x) * np.cos((x - 3.5) / 3.0 * np.pi / 2) ** 2 * (np.abs(x - 3.5) < 3.0)
    z = np.zeros((n, n))
    sim.upload_planes([z, np.tile(prof, (n, 1)), z, z, z, np.tile(prof, (n, 1))])
    st = torch.cuda.ExternalStream(sim.stream)
    res = {}
    for fused in (False, True):
        sim.field_steps(warmup, fused=fused)
        sim.sync()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        l0 = sim.kernel_launches
        e0.record(st)
        sim.field_steps(steps, fused=fused)
        e1.record(st)
        sim.sync()
        res[fused] = (e0.elapsed_time(e1), sim.kernel_launches - l0)
    ms, launches = res[True]
    ms3 = res[False][0]
    cells = float(n) * n * steps
    peak, how = measured_peak()
    gbs = cells * YEE_BYTES_FUSED / (ms * 1e-3) / 1e9
    out = {"metric": "Yee cell-updates/s", "value": cells / (ms * 1e-3), "unit": "cells/s", "ms_per_step": ms / steps,
           "config": {"workload": f"vacuum pulse {n}x{n}, half-B/E/half-B per step in one fused kernel", "fields_bytes": 48 * n * n},
           "gpu_launches": launches,
           "roofline": {"bound": "hbm", "kernel": "fused_yee_kernel", "achieved": gbs, "peak": peak, "unit": "GB/s", "frac": gbs / peak,
                        "traffic": NCU_YEE_TRAFFIC_BYTES_PER_CELL * float(n) * n,
                        "traffic_source": "CONSTANT from one ncu --set full capture of this kernel at 8192^2 (profiles/r02an_ncu_full_fused_yee_8192_raw.csv: "
                                          "95.6 B/cell), per launch; not measured by this run",
                        "bytes_per_cell_update": YEE_BYTES_FUSED,
                        "frac_of_144B_figure": cells * YEE_BYTES_ALGO / (ms * 1e-3) / 1e9 / peak, "peak_source": how,
                        "note": "96 B/cell: six planes read once + written once (halo ring re-reads are served by L2); SURVEY's 144 B "
                                "figure assumes two sweeps, so its fraction exceeds 1"},
           "three_sweeps": {"ms_per_step": ms3 / steps, "value": cells / (ms3 * 1e-3), "bytes_per_cell_update": YEE_BYTES_3SWEEP,
                            "frac": cells * YEE_BYTES_3SWEEP / (ms3 * 1e-3) / 1e9 / peak,
                            "frac_of_144B_figure": cells * YEE_BYTES_ALGO / (ms3 * 1e-3) / 1e9 / peak}}
    sim.close()
    return out


def time_pic(sim, torch, K, W):
    """W warm-up + K timed PIC steps on a single-slab handle, per-phase CUDA events on the library's stream."""
    st = torch.cuda.ExternalStream(sim.stream)

    def step(evs=None):
        if evs:
            evs[0].record(st)
        sim.push_deposit()
        if evs:
            evs[1].record(st)
        sim.step_fields_fused()            # half-B, E - dt*J (J guard ring folded on the fly), half-B: one kernel
        if evs:
            evs[2].record(st)
        sim.sort_tiles()
        if evs:
            evs[3].record(st)

    for _ in range(W):
        step()
    sim.sync()
    evs = [[torch.cuda.Event(enable_timing=True) for _ in range(4)] for _ in range(K)]
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    l0 = sim.kernel_launches
    torch.cuda.synchronize()
    t0.record(st)
    for k in range(K):
        step(evs[k])
    t1.record(st)
    sim.sync()
    torch.cuda.synchronize()
    f = lambda a, b: float(np.mean([e[a].elapsed_time(e[b]) for e in evs]))
    # cell-crossing fraction of the regime: one more push, counted before its sort
    sim.push_deposit()
    movers = sim.cell_crossers()
    sim.step_fields_fused()
    sim.sort_tiles()
    sim.sync()
    return {"ms": t0.elapsed_time(t1), "push_ms": f(0, 1), "fields_ms": f(1, 2), "sort_ms": f(2, 3),
            "launches": int(sim.kernel_launches - l0), "movers": movers}


def push_roofline(n, t, peak, peak_how):
    gbs = n * PUSH_BYTES_SURVEY / (t["push_ms"] * 1e-3) / 1e9
    return {"bound": "hbm", "kernel": "push_packed_kernel (dense slabs; push_deposit_kernel, the lane-group cell walk, for sparse ones)", "achieved": gbs, "peak": peak, "unit": "GB/s", "frac": gbs / peak,
            "bytes_per_particle": PUSH_BYTES_SURVEY,
            "frac_104B": gbs / peak * PUSH_BYTES_FUSED_REORDER / PUSH_BYTES_SURVEY,
            "traffic": NCU_TRAFFIC_BYTES_PER_PARTICLE * n,
            "traffic_source": "CONSTANT from one ncu --set full capture of this kernel in the saturated regime (profiles/r02an_ncu_full_push_packed_saturated_raw.csv: "
                              "117.2 B/particle), scaled by the particle count; not measured by this run",
            "pushes_per_s_kernel_only": n / (t["push_ms"] * 1e-3), "peak_source": peak_how,
            "note": "frac: SURVEY.md 8(d)'s 88 B push + 2.3 B grid = 90 B/particle.  frac_104B: what this kernel must move, since it is also the "
                    "sort's reorder (52 B read + 52 B written per particle); the sort that remains (scan + permutation build) moves 8 B/particle"}


def pic_record(m, torch, name, w, K, W, local, peak, peak_how, clocks=True, keep=False):
    """Load workload `w` on one GPU, advance it into its regime, time K steps.  Returns (record, sim or None)."""
    nx, ny, tn = w["nx"], w["ny"], w["tn"]
    dx = dy = w["dx"]
    dt = m.courant_dt(dx, dy)
    n = expected_particles(w, ny, 0, ny)
    sim = m.Sim(nx, ny, tn, tn, dx, dy, dt, max_particles=n + 1024, device=local)
    load_workload(sim, w)
    sim.sync()
    assert sim.num_particles == n
    sat = int(w.get("saturate_steps", 0))
    with (ClockSampler(local) if clocks else contextlib.nullcontext()) as clk:      # sampled over the untimed run-in too (same kernels, same load)
        if sat:
            sim.pic_steps(sat)
            sim.sync()
        t = time_pic(sim, torch, K, W)
    diag = sim.diagnostics()
    rec = {"metric": "particle-pushes/s (push+deposit+sort)", "value": n * K / (t["ms"] * 1e-3), "unit": "particles/s",
           "ms_per_step": t["ms"] / K, "particles": n,
           "regime": (f"saturated: timed after {sat} untimed steps + {W} warm-up steps" if sat else f"thermal u_th={w['u_th']}: timed after {W} warm-up steps"),
           "cell_crossing_fraction": t["movers"] / float(n),
           # the sort chain runs on its own stream NEXT TO the fused field kernel; the events sit on the main stream, so the
           # second entry is the field kernel with part of the sort beside it and the third is only what is left of the sort
           # after the fields have finished -- not the duration of the sort kernels (no bandwidth figure is derived from it)
           "phases_ms": {"push_deposit": t["push_ms"], "fields_fused": t["fields_ms"], "sort": t["sort_ms"],
                         "note": "sort = tail of the sort chain left after the concurrent field kernel, not the sort's own duration"},
           "roofline": push_roofline(n, t, peak, peak_how), "gpu_launches": t["launches"],
           "diagnostics_after": {k: float(v) for k, v in diag.items()}}
    if clocks:
        rec["clocks"] = clk.summary()
    # self-check: nothing lost, charge conserved, Gauss's law at round-off (charge-conserving deposit)
    q0 = -1.0 * dx * dy / w["ppc"] * n if "beam_ppc" not in w else diag["charge"]
    # (the beam load is not a solution of Gauss's law at t = 0 -- a dense band on zero fields --, so only the neutral loads are held to it)
    ok = diag["particles"] == n and abs(diag["charge"] - q0) <= 1e-9 * abs(q0) and (diag["gauss_residual"] < 1e-9 or "beam_ppc" in w)
    rec["self_check"] = {"ok": bool(ok), "particles": diag["particles"], "charge_error": float(diag["charge"] - q0),
                         "gauss_residual": float(diag["gauss_residual"])}
    if not ok:
        print(f"bench.py: self-check FAILED for {name}: {rec['self_check']}", file=sys.stderr)
        sys.exit(3)
    if keep:
        return rec, sim
    sim.close()
    return rec, None


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--workload", default="two_stream_512", choices=sorted(WORKLOADS))
    ap.add_argument("--no-yee", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-sub", action="store_true", help="skip the c4_slab / quiet-start sub-records")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)

    import torch
    m = importlib.import_module(PKG)
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torchrun --nproc-per-node {args.gpus}")
    ClockSampler(local)                    # start nvidia-smi now: its start-up must not eat the first timed window
    torch.cuda.set_device(local)
    w = WORKLOADS[args.workload]
    K, W = args.steps, max(args.warmup, 3)
    peak, peak_how = measured_peak()

    if world > 1:
        dist = importlib.import_module(PKG + ".slab")
        return dist.bench_multi(args, w, rank, world, local, peak, peak_how)
    if w.get("fields_only"):
        with ClockSampler(local) as clk:
            y = yee_bench(m, torch, K, W, n=w["nx"])
        out = {"metric": y["metric"], "value": y["value"], "unit": y["unit"], "n_gpus": 1, "steps": K, "warmup": W,
               "ms_per_step": y["ms_per_step"], "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
               "data": "synthetic", "config": workload_config(args.workload, w), "roofline": y["roofline"], "three_sweeps": y["three_sweeps"],
               "gpu_launches": y["gpu_launches"], "clocks": clk.summary()}
        if not args.no_cpu:
            out["cpu_baseline"] = yee_cpu_reference()
        print(json.dumps(out))
        return

    nx, ny = w["nx"], w["ny"]
    rec, sim = pic_record(m, torch, args.workload, w, K, W, local, peak, peak_how, keep=True)
    n = rec["particles"]
    out = {
        "metric": rec["metric"], "value": rec["value"], "unit": rec["unit"], "n_gpus": 1, "steps": K,
        "warmup": W, "ms_per_step": rec["ms_per_step"], "higher_is_better": True, "scaling": "strong" if w.get("strong") else "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": workload_config(args.workload, w),
        "regime": rec["regime"], "cell_crossing_fraction": rec["cell_crossing_fraction"],
        "phases_ms": rec["phases_ms"], "roofline": rec["roofline"], "gpu_launches": rec["gpu_launches"], "clocks": rec["clocks"],
        "diagnostics_after": rec["diagnostics_after"], "self_check": rec["self_check"],
    }

    if n * 52 > 16e9 and not args.no_e2e:
        args.no_e2e = True                 # the host image of a billion-particle state does not belong in a default bench run
        out["e2e_skipped"] = "host image > 16 GB"
    if not args.no_e2e:
        # end to end through the host-buffer C-ABI call: pinned host state in, 1 PIC step, full state out, every step
        host = {k: torch.empty(n, dtype=torch.float64).pin_memory() for k in m.PART_KEYS}
        host["tag"] = torch.empty(n, dtype=torch.int32).pin_memory()
        parts = {k: v.numpy() for k, v in host.items()}
        sim.download_particles(out=parts)
        hpl = [torch.empty((ny, nx), dtype=torch.float64).pin_memory() for _ in range(6)]
        planes = [t.numpy() for t in hpl]
        sim.download_planes(out=planes)
        ke = max(2, min(K, 5))
        sim.pic_steps_host(1, planes, parts)
        torch.cuda.synchronize()
        tw = time.perf_counter()
        for _ in range(ke):
            sim.pic_steps_host(1, planes, parts)
        torch.cuda.synchronize()
        tw = time.perf_counter() - tw
        nbytes = 6 * nx * ny * 8 + n * 52
        out["e2e"] = {"value": n * ke / tw, "unit": "particles/s", "h2d_bytes_per_step": nbytes, "d2h_bytes_per_step": nbytes,
                      "steps": ke, "ms_per_step": 1e3 * tw / ke,
                      "api": "mpic_pic_steps_host: pinned host fields+particles -> device, 1 PIC step, full state back (chunked, both PCIe directions overlapped with the compute)"}
    sim.close()
    if not args.no_sub and args.workload == "two_stream_512":
        # secondary figures: the loader's quiet start (round 1's headline regime) and the per-GPU share of configs[3]
        wq = dict(w)
        wq.pop("saturate_steps")
        q, _ = pic_record(m, torch, "two_stream_512 (quiet start)", wq, K, W, local, peak, peak_how, clocks=False)
        q["regime"] = f"quiet start: steps {W + 1}..{W + K} after the loader (almost no particle changes cell)"
        out["quiet_start"] = {k: q[k] for k in ("value", "ms_per_step", "regime", "cell_crossing_fraction", "phases_ms", "roofline")}
        c4, _ = pic_record(m, torch, "uniform_c4_slab", WORKLOADS["uniform_c4_slab"], K, W, local, peak, peak_how)
        c4["config"] = workload_config("uniform_c4_slab", WORKLOADS["uniform_c4_slab"])
        c4["scaling"] = "weak"
        out["c4_slab"] = c4
    if not args.no_yee:
        out["yee"] = yee_bench(m, torch, max(K, 10), W)
    if not args.no_cpu:
        if "yee" in out:
            try:
                cb = yee_cpu_reference()
            except Exception as e:      # oracle/_ref is built where /root/reference exists and travels with the snapshot
                cb = {"unavailable": repr(e)}
            if cb:
                out["yee"]["cpu_baseline"] = cb
        out["cpu_baseline"] = cpu_baseline_pic(w)
    ClockSampler.stop_all()
    print(json.dumps(out))


if __name__ == "__main__":
    main()
