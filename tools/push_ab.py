"""Per-phase device timing of the PIC step in a named regime (A/B tool for the push kernel).
  python tools/push_ab.py <regime> [steps] [nx] [ny]
regimes: quiet (two-stream, steps 6..), sat (two-stream after 1500 untimed steps), thermal (u_th 0.05), c4 (4096x512 thermal)"""
import importlib, os, sys, json
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
m = importlib.import_module("mini-pic--2d-pic-code-with-tiling-load-balance_b200")
regime = sys.argv[1]
K = int(sys.argv[2]) if len(sys.argv) > 2 else 20
nx = int(sys.argv[3]) if len(sys.argv) > 3 else (4096 if regime == "c4" else 512)
ny = int(sys.argv[4]) if len(sys.argv) > 4 else 512
ppc, dx = int(os.environ.get("PPC", "64")), 0.1
sim = m.Sim(nx, ny, 16, 16, dx, dx, m.courant_dt(dx, dx), max_particles=nx * ny * ppc + 1024)
if regime in ("quiet", "sat"):
    sim.load_uniform(ppc, u_th=0.001, two_stream=0.2, seed=12345)
else:
    sim.load_uniform(ppc, u_th=0.05, seed=12345)
sim.pic_steps(1500 if regime == "sat" else 5)
sim.sync()
st = torch.cuda.ExternalStream(sim.stream)
evs = [[torch.cuda.Event(enable_timing=True) for _ in range(4)] for _ in range(K)]
for e in evs:
    e[0].record(st); sim.push_deposit(); e[1].record(st)
    sim.step_b_half(); sim.step_e_full(); sim.step_b_half(); e[2].record(st)
    sim.sort_tiles(); e[3].record(st)
sim.sync()
f = lambda a, b: float(np.mean([e[a].elapsed_time(e[b]) for e in evs]))
n = sim.num_particles
print(json.dumps({"regime": regime, "ppc": ppc, "gl": os.environ.get("MPIC_PUSH_GL", "auto"), "n": n, "push_ms": round(f(0, 1), 4), "yee_ms": round(f(1, 2), 4),
                  "sort_ms": round(f(2, 3), 4), "step_ms": round(f(0, 3), 4), "push_frac104": round(n * 104 / (f(0, 1) * 1e-3) / 6446e9, 4),
                  "push_frac90": round(n * 90 / (f(0, 1) * 1e-3) / 6446e9, 4)}))
