# N-GPU check: slab parity tests (NCCL + CUDA-IPC peer ring), then the weak-scaling bench line at N = $1 (default 2), both transports
N=${1:-2}
timeout 900 python -m pytest tests/test_gpu_slab.py -m gpu -x -q 2>&1 | tail -3
timeout 300 python bench.py --steps 20 --warmup 5 --no-yee --no-cpu --no-e2e > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; cut -c1-330 gpurun_out/bench_n1.json
for tr in peer nccl; do
  MPIC_TRANSPORT=$tr timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5 --no-e2e > gpurun_out/bench_n${N}_$tr.json 2> gpurun_out/bench_n${N}_$tr.err; grep -v "^\*\|OMP_NUM\|^$" gpurun_out/bench_n${N}_$tr.err | tail -5; tail -c 2000 gpurun_out/bench_n${N}_$tr.json | cut -c1-330
done
