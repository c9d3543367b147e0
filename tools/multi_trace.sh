N=${1:-2}
MPIC_PEER_TRACE=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 40 --warmup 5 --no-e2e 2>&1 | grep -E "peer trace|ms_per_step" | cut -c1-420
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 40 --warmup 5 --no-e2e 2>&1 | grep -E "ms_per_step" | cut -c1-220
