"""Why is push_deposit slower in the step loop than alone under ncu?  Times the push kernel (CUDA events on the library
stream) in several regimes and samples clocks/power meanwhile.  Run on the GPU box: python tools/push_gap.py"""
import importlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
m = importlib.import_module("mini-pic--2d-pic-code-with-tiling-load-balance_b200")


class Smi:
    def __init__(self):
        self.rows = []

    def __enter__(self):
        self.p = subprocess.Popen(["nvidia-smi", "--query-gpu=clocks.sm,clocks.mem,power.draw,clocks_event_reasons.sw_power_cap,clocks_event_reasons.hw_slowdown,clocks_event_reasons.sw_thermal_slowdown",
                                   "--format=csv,noheader,nounits", "-lms", "20", "-i", "0"], stdout=subprocess.PIPE, text=True)
        self.t = threading.Thread(target=lambda: [self.rows.append(l.strip()) for l in self.p.stdout], daemon=True)
        self.t.start()
        return self

    def __exit__(self, *a):
        time.sleep(0.05)
        self.p.terminate()
        self.t.join(timeout=2)

    def summary(self):
        sm = [float(r.split(",")[0]) for r in self.rows if r and r[0].isdigit()]
        pw = [float(r.split(",")[2]) for r in self.rows if r and r[0].isdigit()]
        cap = sum("Active" in r.split(",")[3] and "Not" not in r.split(",")[3] for r in self.rows if r and r[0].isdigit())
        return {"n": len(sm), "sm_med": float(np.median(sm)) if sm else None, "sm_min": min(sm) if sm else None,
                "pw_med": float(np.median(pw)) if pw else None, "pw_max": max(pw) if pw else None, "power_cap_samples": cap}


def main():
    nx = ny = 512
    tn, ppc, dx = 16, 64, 0.1
    dt = m.courant_dt(dx, dx)
    n = nx * ny * ppc
    sim = m.Sim(nx, ny, tn, tn, dx, dx, dt, max_particles=n + 1024)
    sim.load_uniform(ppc, u_th=0.001, two_stream=0.2, seed=12345)
    sim.sync()
    st = torch.cuda.ExternalStream(sim.stream)
    flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")

    def rest():
        sim.step_b_half(); sim.step_e_full(); sim.step_b_half(); sim.sort_tiles()

    def run(K, pre=None):
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with Smi() as smi:
            t0.record(st)
            for k in range(K):
                if pre:
                    pre()
                evs[k][0].record(st)
                sim.push_deposit()
                evs[k][1].record(st)
                rest()
            t1.record(st)
            sim.sync()
            torch.cuda.synchronize()
        ms = np.array([a.elapsed_time(b) for a, b in evs])
        return {"push_ms_mean": float(ms.mean()), "push_ms_min": float(ms.min()), "push_ms_first10": [round(float(v), 3) for v in ms[:10]],
                "push_ms_last5": [round(float(v), 3) for v in ms[-5:]], "total_ms_per_step": t0.elapsed_time(t1) / K, "smi": smi.summary()}

    for _ in range(5):
        sim.push_deposit(); rest()
    sim.sync()
    out = {}
    out["A_back_to_back_50"] = run(50)
    if os.environ.get("GAP_FULL"):
        def idle():
            sim.sync(); time.sleep(0.005)
        out["B_idle_before_push"] = run(30, idle)

        def l2flush():
            sim.sync()
            with torch.cuda.stream(st):
                flush.fill_(1)
        out["C_l2_dirty_before_push"] = run(30, l2flush)
    if not os.environ.get("GAP_SHORT"):
        out["E_sustained_800"] = run(800)
        out["F_after_800"] = run(50)
    # thermal plasma: cell occupancy fluctuates (chunks of 32 per cell become ragged)
    sim.load_uniform(ppc, u_th=0.05, two_stream=0.0, seed=7)
    for _ in range(5):
        sim.push_deposit(); rest()
    out["T_thermal_0.05_50"] = run(50)
    if not os.environ.get("GAP_SHORT"):
        out["T_thermal_0.05_after_300"] = (run(300), run(50))[1]
    for k, v in out.items():
        print(k, json.dumps(v))


if __name__ == "__main__":
    main()
