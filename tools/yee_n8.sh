mkdir -p gpurun_out
for tr in peer nccl; do
  MPIC_TRANSPORT=$tr timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 8 --workload yee_8192 --steps 20 --warmup 5 > gpurun_out/yee_n8_$tr.json 2> gpurun_out/yee_n8_$tr.err
  grep -v "^\*\|OMP_NUM\|^$\|^W1\|^\[W" gpurun_out/yee_n8_$tr.err | tail -4; grep "^{" gpurun_out/yee_n8_$tr.json | cut -c1-400
done
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29523 bench.py --gpus 8 --workload beam_c5_slab --steps 10 --warmup 3 --no-e2e > gpurun_out/bench_n8_beam2.json 2> gpurun_out/bench_n8_beam2.err
grep "^{" gpurun_out/bench_n8_beam2.json | cut -c1-300
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29525 bench.py --gpus 8 --steps 20 --warmup 5 --no-e2e > gpurun_out/bench_n8_two_stream2.json 2> gpurun_out/bench_n8_two_stream2.err
grep "^{" gpurun_out/bench_n8_two_stream2.json | cut -c1-300
