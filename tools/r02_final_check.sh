set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -6 > gpurun_out/r02ar_pytest_gpu.log; cat gpurun_out/r02ar_pytest_gpu.log
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -3
timeout 900 python bench.py > gpurun_out/r02ar_bench_n1.json 2> gpurun_out/r02ar_bench.err; tail -c 300 gpurun_out/r02ar_bench_n1.json; tail -3 gpurun_out/r02ar_bench.err
