# launch list + full ncu capture of the push kernel for the current build
mkdir -p gpurun_out
timeout 300 python bench.py --steps 3 --warmup 3 --no-yee --no-cpu --no-e2e > gpurun_out/b.log 2>&1 || exit 1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 3 --warmup 3 --no-yee --no-cpu --no-e2e > gpurun_out/ncu_l.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"push_deposit" -s 3 -c 1 -o gpurun_out/push_full -f python bench.py --steps 3 --warmup 3 --no-yee --no-cpu --no-e2e > gpurun_out/ncu_f.log 2>&1
ncu -i gpurun_out/push_full.ncu-rep --page raw --csv > gpurun_out/push_full_raw.csv 2>/dev/null
ncu -i gpurun_out/push_full.ncu-rep --page source --csv > gpurun_out/push_full_src.csv 2>/dev/null
timeout 900 ncu --set full --clock-control none -k regex:"b_sweep|e_sweep" -s 6 -c 2 -o gpurun_out/yee_full -f python -c "
import importlib,numpy as np
m=importlib.import_module('mini-pic--2d-pic-code-with-tiling-load-balance_b200')
n=8192; dx=10.0/n
s=m.Sim(n,n,32,32,dx,dx,m.courant_dt(dx,dx)); s.field_steps(5); s.sync()" > gpurun_out/ncu_y.log 2>&1
ncu -i gpurun_out/yee_full.ncu-rep --page raw --csv > gpurun_out/yee_full_raw.csv 2>/dev/null
ls -la gpurun_out | tail -8
