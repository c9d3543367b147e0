# A/B of block-size variants x push split
PK=mini-pic--2d-pic-code-with-tiling-load-balance_b200
mkdir -p gpurun_out
for lib in $PK/libminipic_b200*.so; do echo "== $lib"; MPIC_LIB=$PWD/$lib timeout 600 python -m pytest tests/test_gpu_pic.py -m gpu -x -q 2>&1 | tail -1; done | tee gpurun_out/ab_bench2.log
run() { MPIC_PUSH_SPLIT=$3 MPIC_LIB=$PWD/$1 timeout 300 python bench.py --workload $2 --steps 20 --warmup 5 --no-yee --no-cpu --no-e2e 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read());print('$1'[-18:],'split=$3',d['config']['workload'],round(d['ms_per_step'],4),round(d['phases_ms']['push_deposit'],4))"; }
for lib in $PK/libminipic_b200*.so; do for sp in 4 2 1; do run $lib two_stream_512 $sp; done; for sp in 1 2; do run $lib uniform_c4_slab $sp; done; done | tee -a gpurun_out/ab_bench2.log
