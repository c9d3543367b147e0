"""Fused Yee steps at 8192^2 for profiling: python tools/prof_yee.py [steps] [n].  Under ncu use -k regex:fused_yee -s 5 -c 1."""
import importlib
import os
import sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
m = importlib.import_module("mini-pic--2d-pic-code-with-tiling-load-balance_b200")
steps = int(sys.argv[1]) if len(sys.argv) > 1 else 10
n = int(sys.argv[2]) if len(sys.argv) > 2 else 8192
dx = 10.0 / n
sim = m.Sim(n, n, 32, 32, dx, dx, m.courant_dt(dx, dx))
x = (np.arange(n) + 0.5) * dx
row = np.tile(0.1 * np.sin(np.pi * x), (n, 1))
z = np.zeros((n, n))
sim.upload_planes([z, row, z, z, z, row])
sim.field_steps(steps, fused=True)
sim.sync()
print("done", sim.kernel_launches)
