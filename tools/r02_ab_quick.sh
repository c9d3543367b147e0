# quick A/B: packed (1) vs 32-lane cell walk (32) in the named regimes; appends to gpurun_out/$1
OUT=$1; shift
timeout 600 python -m pytest tests/test_gpu_pic.py -x -q -k "packed or variant or stable or ten_pic" 2>&1 | tail -3
for gl in "$@"; do for r in quiet sat thermal; do MPIC_PUSH_GL=$gl timeout 300 python tools/push_ab.py $r >> gpurun_out/$OUT; done; done
cat gpurun_out/$OUT
