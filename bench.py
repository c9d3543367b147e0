#!/usr/bin/env python
"""bench.py -- headline measurement of the Mini-PIC time-step path on B200 (contract in the task statement).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl native|reference] [--workload NAME]

Native arm: one "step" = one full PIC iteration (gather+Boris push+deposit fused with the reorder, J fold, half-B, E with
-dt*J, half-B, per-cell count/scan/permutation sort) over the synthetic workload, state resident in HBM.  N=1 workload = BASELINE.json
configs[1] (two-stream, 512x512 cells, 64 ppc = 16.8 M particles, 0.87 GB of particle state >> 126 MB L2).  N>1: weak
scaling, every rank owns a 512-row slab of a (512N x 512) domain (halo exchange, J reduce and particle migration over
NCCL).  Prints ONE JSON line on rank 0.

Reference arm (--impl reference): the reference has no particle code, so the timed CPU implementation is
the repo's scalar FP64 oracle port with OpenMP over particles ("kind": "port"), on a bounded sample of the same workload.
"""
import argparse
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
    "two_stream_512": dict(nx=512, ny=512, tn=16, ppc=64, u_th=0.001, two_stream=0.2, dx=0.1),
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
# push_deposit_kernel is fused with the sort's data movement: it gathers every particle through the slot->storage
# permutation and writes the other buffer in sorted order, so its algorithmic traffic is one full read + one full write
# of the particle state: x,y,ux,uy,uz,q,tag = 52 B each way.  (SURVEY.md section 8(d) counts 88 B for an in-place push
# PLUS 116 B for a separate reorder; the fused kernel replaces both.  `frac_88B` reports the stricter push-only figure.)
PUSH_BYTES_PER_PARTICLE = 104.0
PUSH_ONLY_BYTES = 88.0
NCU_TRAFFIC_BYTES_PER_PARTICLE = 113.5   # measured: (0.964 GB read + 0.940 GB written) / 16.8 M particles, profiles/r01e_ncu_full_push_two_stream.csv
SORT_BYTES_PER_PARTICLE = 8.0      # what is left of the sort: keys read (4) + permutation written (4); scan is per cell
YEE_BYTES_3SWEEP = 216.0           # reference-faithful half-B, E, half-B: 3 x 72 B per cell
YEE_BYTES_ALGO = 144.0             # SURVEY.md section 8(d) algorithmic figure (2 sweeps)


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


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks + throttle reasons DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                                          "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def __exit__(self, *a):
        if self.proc:
            time.sleep(0.15)
            self.proc.terminate()
            self.t.join(timeout=2)

    def summary(self):
        sm = [float(r[0]) for r in self.rows if len(r) >= 6 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 6 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[k] for r in self.rows if len(r) >= 6 for k in range(4) if r[2 + k].lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(sm)}


def cpu_port_pushes_per_s(w, steps=2, sample_cells=256, omp=True):
    """The oracle port (OpenMP over particles, or the scalar build with omp=False) on a bounded sample: a (sample_cells x
    sample_cells) piece of the workload."""
    from oracle import bindings as ob
    plasma = importlib.import_module(PKG + ".plasma")
    m = importlib.import_module(PKG)
    lib = ob.load_oracle(omp=omp)
    nx = ny = min(sample_cells, w["nx"])
    dx = dy = w["dx"]
    dt = m.courant_dt(dx, dy)
    P = ob.make_params(lib, nx, ny, w["tn"], w["tn"], dx, dy, dt)
    parts = plasma.load_uniform(nx, ny, w["tn"], w["tn"], dx, dy, w["ppc"], u_th=w["u_th"], two_stream=w["two_stream"])
    planes = [np.zeros((ny, nx)) for _ in range(6)]
    ob.pic_steps(lib, P, planes, parts, 1)                      # warm-up (page faults, thread pool)
    t0 = time.perf_counter()
    ob.pic_steps(lib, P, planes, parts, steps)
    dtw = time.perf_counter() - t0
    n = parts["x"].shape[0]
    return n * steps / dtw, (lib.orc_num_threads() if omp else 1), f"{nx}x{ny} cells x {w['ppc']} ppc = {n} particles, {steps} full PIC steps (push+deposit+Yee+sort)"


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    w = WORKLOADS[args.workload]
    if w.get("fields_only"):
        # the reference's OWN field path (unmodified Field_update.cpp + pack/unpack under its loop structure), one rank
        vals = [yee_cpu_reference(steps=10) for _ in range(max(1, min(args.steps, 3)))]
        cb = sorted(vals, key=lambda c: c["value"])[len(vals) // 2]
        v = cb["value"]
        print(json.dumps({
            "impl": "reference", "metric": "Yee cell-updates/s", "value": v, "unit": "cells/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * 1800.0 * 1800.0 / v, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": args.workload, "note": "bounded sample of the same pulse: " + cb["sample"]},
            "cpu_baseline": cb, "e2e": {"value": v, "unit": "cells/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}))
        return
    vals = []
    sample = cores = None
    for _ in range(max(1, min(args.steps, 3))):
        v, cores, sample = cpu_port_pushes_per_s(w, steps=2)
        vals.append(v)
    v = float(np.median(vals))
    n_sample = min(256, w["nx"]) ** 2 * w["ppc"]
    print(json.dumps({
        "impl": "reference", "metric": "particle-pushes/s (push+deposit+sort)", "value": v, "unit": "particles/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * n_sample / v,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": args.workload, **w, "note": "reference has no particle code; timed = repo's scalar FP64 oracle port, OpenMP over particles"},
        "cpu_baseline": {"value": v, "unit": "particles/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": "particles/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}))


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
        r = ob.ref_run_loop(ref, g6, 1, 1, n // tile, n // tile, 2, dt, dx, dx, steps, want_tiles=False)
        res["omp_over_tiles"] = {"value": float(n) * n * steps / r["seconds"], "cores": int(ref.ref_num_threads()),
                                 "note": "reference functions, OpenMP over tiles (not the reference's MPI decomposition)"}
    return res


def yee_bench(m, torch, steps, warmup, n=8192):
    """Field-only sub-benchmark (BASELINE.json configs[2]): n x n vacuum pulse, reference-faithful 3-sweep loop."""
    dx = dy = 10.0 / n
    sim = m.Sim(n, n, 32, 32, dx, dy, m.courant_dt(dx, dy))
    x = (np.arange(n) + 0.5) * dx
    prof = 0.1 * np.sin(np.pi * x) * np.cos((x - 3.5) / 3.0 * np.pi / 2) ** 2 * (np.abs(x - 3.5) < 3.0)
    z = np.zeros((n, n))
    sim.upload_planes([z, np.tile(prof, (n, 1)), z, z, z, np.tile(prof, (n, 1))])
    st = torch.cuda.ExternalStream(sim.stream)
    sim.field_steps(warmup)
    sim.sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    l0 = sim.kernel_launches
    e0.record(st)
    sim.field_steps(steps)
    e1.record(st)
    sim.sync()
    ms = e0.elapsed_time(e1)
    cells = float(n) * n * steps
    peak, how = measured_peak()
    out = {"metric": "Yee cell-updates/s", "value": cells / (ms * 1e-3), "unit": "cells/s", "ms_per_step": ms / steps,
           "config": {"workload": f"vacuum pulse {n}x{n}, half-B/E/half-B per step", "fields_bytes": 48 * n * n},
           "gpu_launches": sim.kernel_launches - l0,
           "roofline": {"bound": "hbm", "achieved": cells * YEE_BYTES_3SWEEP / (ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                        "frac": cells * YEE_BYTES_3SWEEP / (ms * 1e-3) / 1e9 / peak,
                        "traffic": (3.225 + 1.568 + 3.225 + 1.572 + 3.225 + 1.568) * 1e9 * (float(n) * n / 8192.0 ** 2),
                        "traffic_source": "ncu --set full dram__bytes_read+write of b_sweep / e_sweep at 8192^2 (profiles/r01e_ncu_full_yee_8192.csv), three sweeps per step: 14.38 GB vs 14.50 GB algorithmic",
                        "bytes_per_cell_update": YEE_BYTES_3SWEEP,
                        "frac_of_144B_figure": cells * YEE_BYTES_ALGO / (ms * 1e-3) / 1e9 / peak, "peak_source": how}}
    sim.close()
    return out


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
    torch.cuda.set_device(local)
    w = WORKLOADS[args.workload]
    K, W = args.steps, max(args.warmup, 3)
    peak, peak_how = measured_peak()

    if world > 1:
        dist = importlib.import_module(PKG + ".slab")
        if w.get("fields_only"):
            return dist.bench_fields(args, w, rank, world, local, peak, peak_how)
        return dist.bench_main(args, w, rank, world, local, peak, peak_how)
    if w.get("fields_only"):
        with ClockSampler(local) as clk:
            y = yee_bench(m, torch, K, W, n=w["nx"])
        out = {"metric": y["metric"], "value": y["value"], "unit": y["unit"], "n_gpus": 1, "steps": K, "warmup": W,
               "ms_per_step": y["ms_per_step"], "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
               "data": "synthetic", "config": {"workload": args.workload, **y["config"], "l2": "48 B/cell x 67 M cells = 3.2 GB >> 126 MB L2"},
               "roofline": y["roofline"], "gpu_launches": y["gpu_launches"], "clocks": clk.summary()}
        if not args.no_cpu:
            out["cpu_baseline"] = yee_cpu_reference()
        print(json.dumps(out))
        return

    nx, ny, tn, ppc = w["nx"], w["ny"], w["tn"], w["ppc"]
    dx = dy = w["dx"]
    dt = m.courant_dt(dx, dy)
    n = expected_particles(w, ny, 0, ny)
    sim = m.Sim(nx, ny, tn, tn, dx, dy, dt, max_particles=n + 1024, device=local)
    load_workload(sim, w)
    sim.sync()
    assert sim.num_particles == n
    st = torch.cuda.ExternalStream(sim.stream)

    def step(evs=None):
        if evs:
            evs[0].record(st)
        sim.push_deposit()
        if evs:
            evs[1].record(st)
        sim.step_b_half()                 # (the E sweep folds the J guard ring on the fly)
        sim.step_e_full()
        sim.step_b_half()
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
    with ClockSampler(local) as clk:
        torch.cuda.synchronize()
        t0.record(st)
        for k in range(K):
            step(evs[k])
        t1.record(st)
        sim.sync()
        torch.cuda.synchronize()
    ms = t0.elapsed_time(t1)
    launches = sim.kernel_launches - l0
    push_ms = float(np.mean([e[0].elapsed_time(e[1]) for e in evs]))
    field_ms = float(np.mean([e[1].elapsed_time(e[2]) for e in evs]))
    sort_ms = float(np.mean([e[2].elapsed_time(e[3]) for e in evs]))
    value = n * K / (ms * 1e-3)
    push_gbs = n * PUSH_BYTES_PER_PARTICLE / (push_ms * 1e-3) / 1e9
    diag = sim.diagnostics()

    out = {
        "metric": "particle-pushes/s (push+deposit+sort)", "value": value, "unit": "particles/s", "n_gpus": 1, "steps": K,
        "warmup": W, "ms_per_step": ms / K, "higher_is_better": True, "scaling": "strong" if w.get("strong") else "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic",
        "config": {"workload": args.workload, **w, "particles": n, "tile": f"{tn}x{tn}", "dt": dt,
                   "regime": f"steps {W + 1}..{W + K} after the loader's quiet start (the whole 2000-step run of configs[1] averages 0.649 ms/step: DESIGN.md section 7.0)",
                   "l2": f"particle state {n * 52 / 1e9:.2f} GB per pass >> 126 MB L2 (no flush needed)"},
        "phases_ms": {"push_deposit": push_ms, "fields_and_fold": field_ms, "sort": sort_ms},
        "roofline": {"bound": "hbm", "kernel": "push_deposit_kernel", "achieved": push_gbs, "peak": peak, "unit": "GB/s",
                     "frac": push_gbs / peak, "traffic": NCU_TRAFFIC_BYTES_PER_PARTICLE * n, "bytes_per_particle": PUSH_BYTES_PER_PARTICLE,
                     "frac_88B": push_gbs / peak * PUSH_ONLY_BYTES / PUSH_BYTES_PER_PARTICLE,
                     "traffic_source": "ncu --set full dram__bytes_read+write of one launch, profiles/r01e_ncu_full_push_two_stream.csv (113.5 B/particle)",
                     "pushes_per_s_kernel_only": n / (push_ms * 1e-3), "peak_source": peak_how,
                     "sort_gbs": n * SORT_BYTES_PER_PARTICLE / (sort_ms * 1e-3) / 1e9,
                     "note": "fused gather+push+deposit+reorder: 52 B read + 52 B written per particle; the sort that remains (scan + permutation build) moves 8 B/particle"},
        "gpu_launches": int(launches), "clocks": clk.summary(),
        "diagnostics_after": {k: float(v) for k, v in diag.items()},
    }

    if n * 52 > 16e9 and not args.no_e2e:
        args.no_e2e = True                 # the host image of a billion-particle state does not belong in a default bench run
        out["e2e_skipped"] = "host image > 16 GB"
    if not args.no_e2e:
        # end to end through the host-buffer C-ABI call: pinned host state in, 1 PIC step, full state out, every step
        host = {k: torch.empty(n, dtype=torch.float64).pin_memory() for k in m.PART_KEYS}
        host["tag"] = torch.empty(n, dtype=torch.int32).pin_memory()
        parts = {k: v.numpy() for k, v in host.items()}
        cur = sim.download_particles()
        for k in parts:
            parts[k][:] = cur[k]
        hpl = [torch.empty((ny, nx), dtype=torch.float64).pin_memory() for _ in range(6)]
        planes = [t.numpy() for t in hpl]
        for a, b in zip(planes, sim.download_planes()):
            a[:] = b
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
                      "api": "mpic_pic_steps_host: pinned host fields+particles -> device, 1 PIC step, full state back"}
    sim.close()
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
        v, cores, sample = cpu_port_pushes_per_s(w, steps=2)
        v1, _, sample1 = cpu_port_pushes_per_s(w, steps=1, sample_cells=128, omp=False)
        out["cpu_baseline"] = {"value": v, "unit": "particles/s", "cores": cores, "kind": "port", "sample": sample,
                               "value_single_core": v1, "sample_single_core": sample1,
                               "note": "reference has no particle code (SURVEY.md F1); this is the repo's scalar FP64 oracle with OpenMP"}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
