set -x
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -8
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r02f_bench_n1.json 2> gpurun_out/r02f_bench_n1.err; echo "bench rc=$?"; tail -3 gpurun_out/r02f_bench_n1.err
timeout 600 python bench.py --impl reference --steps 5 --warmup 2 > gpurun_out/r02f_bench_ref.json 2>gpurun_out/r02f_bench_ref.err; echo "ref rc=$?"
