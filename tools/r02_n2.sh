# N=2 evidence: the multi-process tests on two real devices, the default bench line (two-stream slabs + c4_slab + yee_8192 sub-records),
# and the beam workload with the device-side rebalancer
set -x
mkdir -p gpurun_out
N=${N:-2}
T=${TAG:-r02ac}
nvidia-smi -L | head -8
[ -n "$SKIP_TESTS" ] || timeout 900 python -m pytest tests/test_gpu_slab.py tests/test_gpu_driver.py -x -q 2>&1 | tail -4 > gpurun_out/${T}_pytest_n${N}.log; cat gpurun_out/${T}_pytest_n${N}.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
[ -n "$BEAM_ONLY" ] || timeout 900 $TR --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/${T}_bench_n${N}.json 2> gpurun_out/${T}_bench_n${N}.err
timeout 900 $TR --master-port 29512 bench.py --gpus $N --steps 20 --warmup 5 --workload beam_c5_slab --no-e2e --no-cpu > gpurun_out/${T}_bench_n${N}_beam.json 2> gpurun_out/${T}_bench_n${N}_beam.err; tail -c 1500 gpurun_out/${T}_bench_n${N}_beam.json; tail -5 gpurun_out/${T}_bench_n${N}_beam.err
