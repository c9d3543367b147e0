set -x
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r02w_bench_n1.json 2> gpurun_out/r02w_bench_n1.err; echo rc=$?
tail -3 gpurun_out/r02w_bench_n1.err
