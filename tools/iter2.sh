PK=mini-pic--2d-pic-code-with-tiling-load-balance_b200
mkdir -p gpurun_out
run() { MPIC_LIB=$PWD/$1 timeout 300 python bench.py --workload $2 --steps 20 --warmup 5 --no-yee --no-cpu --no-e2e 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read());print('$1'[-18:],d['config']['workload'],round(d['ms_per_step'],4),round(d['phases_ms']['push_deposit'],4))"; }
for lib in $PK/libminipic_b200*.so; do for wl in beam_c5_slab thermal_128 two_stream_512; do run $lib $wl; done; done | tee gpurun_out/iter2.log
for ch in 1048576 2097152 4194304; do MPIC_PIPE_CHUNK=$ch timeout 300 python bench.py --steps 10 --warmup 3 --no-yee --no-cpu 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read());print('chunk $ch e2e', d['e2e']['ms_per_step'], d['e2e']['value'])"; done | tee -a gpurun_out/iter2.log
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 | tee gpurun_out/pytest_gpu.log
