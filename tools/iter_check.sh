# after a kernel change: parity suite, full-size properties, N=1 bench lines of the workloads
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 > gpurun_out/pytest_gpu.log; cat gpurun_out/pytest_gpu.log
timeout 300 python -m pytest tests/test_gpu_fullsize.py -m gpu -q -s -k two_stream 2>&1 | grep "growth rate" > gpurun_out/two_stream_growth.log; cat gpurun_out/two_stream_growth.log
for wl in two_stream_512 thermal_128 uniform_c4_slab beam_c5_slab; do
  timeout 300 python bench.py --workload $wl --steps 20 --warmup 5 --no-yee --no-cpu --no-e2e > gpurun_out/bench_$wl.json 2> gpurun_out/bench_$wl.err; tail -3 gpurun_out/bench_$wl.err
  python -c "
import json
d=json.load(open('gpurun_out/bench_$wl.json'));print(d['config']['workload'],d['config']['particles'],round(d['ms_per_step'],4),'%.3e'%d['value'],d['phases_ms'],round(d['roofline']['frac'],3))"
done
