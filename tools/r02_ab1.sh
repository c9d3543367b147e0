set -x
mkdir -p gpurun_out
for gl in 32 16 8; do for r in quiet thermal sat c4; do MPIC_PUSH_GL=$gl timeout 300 python tools/push_ab.py $r >> gpurun_out/r02b_ab.jsonl || exit 1; done; done
cat gpurun_out/r02b_ab.jsonl
for gl in 16 32 8; do MPIC_PUSH_GL=$gl timeout 900 python -m pytest tests/test_gpu_pic.py -x -q 2>&1 | tail -3; done
