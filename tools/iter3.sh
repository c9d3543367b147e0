PK=mini-pic--2d-pic-code-with-tiling-load-balance_b200
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 | tee gpurun_out/pytest_gpu.log
run() { timeout 300 python bench.py --workload $1 --steps 20 --warmup 5 --no-yee --no-cpu --no-e2e 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read());print(d['config']['workload'],round(d['ms_per_step'],4),round(d['phases_ms']['push_deposit'],4),round(d['roofline']['frac'],4))"; }
for wl in beam_c5_slab thermal_128 two_stream_512 uniform_c4_slab; do run $wl; done | tee gpurun_out/iter3.log
timeout 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; tail -3 gpurun_out/bench.err; cut -c1-1500 gpurun_out/bench.json
