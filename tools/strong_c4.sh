# strong scaling of configs[3] (4096^2 x 64 ppc = 1.07 B particles): N = 1, 2, 4 on one box (N = 8 is the weak c4-slab line)
mkdir -p gpurun_out
timeout 400 python bench.py --workload uniform_c4 --steps 5 --warmup 3 --no-yee --no-cpu --no-e2e > gpurun_out/strong_n1.json 2> gpurun_out/strong_n1.err; tail -2 gpurun_out/strong_n1.err; cut -c1-300 gpurun_out/strong_n1.json
for N in 2 4; do
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29519 bench.py --gpus $N --workload uniform_c4 --steps 5 --warmup 3 --no-e2e > gpurun_out/strong_n$N.json 2> gpurun_out/strong_n$N.err
  grep -v "^\*\|OMP_NUM\|^$\|^W1\|^\[W" gpurun_out/strong_n$N.err | tail -4; grep "^{" gpurun_out/strong_n$N.json | cut -c1-300
done
