set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_pic.py tests/test_gpu_driver.py -x -q 2>&1 | tail -4 > gpurun_out/r02aj_pytest.log; cat gpurun_out/r02aj_pytest.log
timeout 600 python bench.py --no-yee --no-cpu --no-sub > gpurun_out/r02aj_bench_n1_e2e.json 2> gpurun_out/r02aj_bench.err; tail -c 700 gpurun_out/r02aj_bench_n1_e2e.json; tail -3 gpurun_out/r02aj_bench.err
timeout 600 python bench.py --workload beam_c5_slab --no-yee --no-cpu --no-sub --no-e2e > gpurun_out/r02aj_beam_n1_adapt.json 2>> gpurun_out/r02aj_bench.err
MPIC_PUSH_GL=1 timeout 600 python bench.py --workload beam_c5_slab --no-yee --no-cpu --no-sub --no-e2e > gpurun_out/r02aj_beam_n1_packed.json 2>> gpurun_out/r02aj_bench.err
tail -c 300 gpurun_out/r02aj_beam_n1_adapt.json; tail -c 300 gpurun_out/r02aj_beam_n1_packed.json
