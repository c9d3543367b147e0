"""Steady-state PIC steps for profiling: python tools/prof_push.py [steps] [u_th] [two_stream] [nx] [ppc].  Under ncu use
-k regex:push_deposit -s <skip> -c 1 to capture one representative push."""
import importlib
import os
import sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
m = importlib.import_module("mini-pic--2d-pic-code-with-tiling-load-balance_b200")
steps = int(sys.argv[1]) if len(sys.argv) > 1 else 30
u_th = float(sys.argv[2]) if len(sys.argv) > 2 else 0.05
ts = float(sys.argv[3]) if len(sys.argv) > 3 else 0.0
nx = int(sys.argv[4]) if len(sys.argv) > 4 else 512
ppc = int(sys.argv[5]) if len(sys.argv) > 5 else 64
dx = 0.1
sim = m.Sim(nx, nx, 16, 16, dx, dx, m.courant_dt(dx, dx), max_particles=nx * nx * ppc + 1024)
sim.load_uniform(ppc, u_th=u_th, two_stream=ts, seed=12345)
sim.pic_steps(steps)
sim.sync()
print("done", sim.num_particles, sim.kernel_launches)
