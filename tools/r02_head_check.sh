# HEAD check: GPU parity suite, smoke, bench lines (native + reference arm)
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -6 > gpurun_out/r02v_pytest_gpu.log; cat gpurun_out/r02v_pytest_gpu.log
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -3
timeout 900 python bench.py > gpurun_out/r02v_bench_n1.json 2> gpurun_out/r02v_bench.err; tail -c 1500 gpurun_out/r02v_bench_n1.json; tail -5 gpurun_out/r02v_bench.err
