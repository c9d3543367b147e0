set -x
nvidia-smi -L
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29655 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r02g_bench_n2.json 2> gpurun_out/r02g_bench_n2.err; echo "rc=$?"
tail -5 gpurun_out/r02g_bench_n2.err
timeout 600 python -m pytest tests/test_gpu_slab.py -x -q -k "two_processes" 2>&1 | tail -3
