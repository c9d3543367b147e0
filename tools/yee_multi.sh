N=${1:-4}
mkdir -p gpurun_out
timeout 300 python bench.py --workload yee_8192 --steps 20 --warmup 5 --no-cpu > gpurun_out/yee_n1.json 2> gpurun_out/yee_n1.err; tail -2 gpurun_out/yee_n1.err; cut -c1-400 gpurun_out/yee_n1.json
for n in 2 4 8; do [ $n -le $N ] || continue
  for tr in peer nccl; do
  MPIC_TRANSPORT=$tr timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus $n --workload yee_8192 --steps 20 --warmup 5 > gpurun_out/yee_n${n}_$tr.json 2> gpurun_out/yee_n${n}_$tr.err
  grep -v "^\*\|OMP_NUM\|^$\|^W1\|^\[W" gpurun_out/yee_n${n}_$tr.err | tail -4; grep "^{" gpurun_out/yee_n${n}_$tr.json | cut -c1-400
  done
done
