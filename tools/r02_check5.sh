set -x
timeout 900 python -m pytest tests/test_gpu_pic.py -x -q 2>&1 | tail -8
for gl in 1 32; do for r in quiet sat thermal c4; do MPIC_PUSH_GL=$gl timeout 300 python tools/push_ab.py $r >> gpurun_out/r02j_ab.jsonl; done; done
cat gpurun_out/r02j_ab.jsonl
