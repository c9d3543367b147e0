# A/B of library variants on the same box: parity + bench lines alternating between every libminipic_b200*.so
PK=mini-pic--2d-pic-code-with-tiling-load-balance_b200
mkdir -p gpurun_out
for lib in $PK/libminipic_b200*.so; do echo "== $lib"; MPIC_LIB=$PWD/$lib timeout 600 python -m pytest tests/test_gpu_pic.py -m gpu -x -q 2>&1 | tail -1; done | tee gpurun_out/ab_bench.log
for rep in 1 2; do for wl in two_stream_512 uniform_c4_slab; do for lib in $PK/libminipic_b200*.so; do
  MPIC_LIB=$PWD/$lib timeout 300 python bench.py --workload $wl --steps 20 --warmup 5 --no-yee --no-cpu --no-e2e 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read());print('$lib'[-18:],d['config']['workload'],round(d['ms_per_step'],4),round(d['phases_ms']['push_deposit'],4))"
done; done; done | tee -a gpurun_out/ab_bench.log
