# N-GPU check: NCCL slab parity test, then the weak-scaling bench line at N = $1 (default 2)
N=${1:-2}
timeout 600 python -m pytest tests/test_gpu_slab.py -m gpu -x -q -k nccl 2>&1 | tail -3
timeout 300 python bench.py --steps 20 --warmup 5 --no-yee --no-cpu --no-e2e > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; cut -c1-400 gpurun_out/bench_n1.json
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5 --no-e2e > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; tail -c 1500 gpurun_out/bench_n$N.json; tail -5 gpurun_out/bench_n$N.err
