# N-GPU bench lines (peer ring): default workload, C4 per-GPU slab, C5 beam with rebalance
N=${1:-4}
mkdir -p gpurun_out
run() { # name, nproc, extra args
  timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node $2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $2 ${@:3} > gpurun_out/$1.json 2> gpurun_out/$1.err
  echo "== $1 rc=$?"; grep -v "^\*\|OMP_NUM\|^$\|^W1\|^\[W" gpurun_out/$1.err | tail -6; cut -c1-600 gpurun_out/$1.json
}
run bench_n${N}_two_stream $N --steps 20 --warmup 5 --no-e2e
run bench_n${N}_c4slab $N --steps 10 --warmup 3 --no-e2e --workload uniform_c4_slab
run bench_n${N}_beam $N --steps 10 --warmup 3 --no-e2e --workload beam_c5_slab
run bench_n2_two_stream 2 --steps 20 --warmup 5 --no-e2e
run bench_n${N}_two_stream_e2e $N --steps 5 --warmup 3
