"""Field-only timing at n x n: three sweeps vs the fused kernel.  python tools/yee_ab.py [n] [steps]"""
import importlib, os, sys, json
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
m = importlib.import_module("mini-pic--2d-pic-code-with-tiling-load-balance_b200")
n = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
K = int(sys.argv[2]) if len(sys.argv) > 2 else 20
dx = 10.0 / n
sim = m.Sim(n, n, 32, 32, dx, dx, m.courant_dt(dx, dx))
x = (np.arange(n) + 0.5) * dx
prof = 0.1 * np.sin(np.pi * x)
z = np.zeros((n, n))
sim.upload_planes([z, np.tile(prof, (n, 1)), z, z, z, np.tile(prof, (n, 1))])
st = torch.cuda.ExternalStream(sim.stream)
out = {"n": n, "lib": os.environ.get("MPIC_LIB", "default")}
for fused in (False, True):
    sim.field_steps(3, fused=fused); sim.sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st); sim.field_steps(K, fused=fused); e1.record(st); sim.sync()
    ms = e0.elapsed_time(e1) / K
    out["fused_ms" if fused else "sweeps_ms"] = round(ms, 4)
    out["fused_frac144" if fused else "sweeps_frac144"] = round(n * n * 144 / (ms * 1e-3) / 6446e9, 4)
print(json.dumps(out))
