# full-size property tests + extra workloads at N=1 + full ncu capture of the push kernel
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_fullsize.py -m gpu -q -s 2>&1 | tail -40 > gpurun_out/pytest_fullsize.log; cat gpurun_out/pytest_fullsize.log
for wl in uniform_c4_slab beam_c5_slab; do
  timeout 300 python bench.py --workload $wl --steps 10 --warmup 3 --no-yee --no-cpu --no-e2e > gpurun_out/bench_$wl.json 2> gpurun_out/bench_$wl.err; tail -3 gpurun_out/bench_$wl.err; cut -c1-1800 gpurun_out/bench_$wl.json
done
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"push_deposit" -s 3 -c 1 -o gpurun_out/push_full -f python bench.py --steps 3 --warmup 3 --no-yee --no-cpu --no-e2e > gpurun_out/ncu_f.log 2>&1
ncu -i gpurun_out/push_full.ncu-rep --page raw --csv > gpurun_out/push_full_raw.csv 2>/dev/null
ncu -i gpurun_out/push_full.ncu-rep --page source --csv > gpurun_out/push_full_src.csv 2>/dev/null
ls -la gpurun_out | tail -8
