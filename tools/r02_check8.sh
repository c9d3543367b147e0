# packed push with the part-row {f, df/dx} patch and the integer wrap pre-test: parity, then A/B against the previous build
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_pic.py -x -q 2>&1 | tail -4 > gpurun_out/r02y_pytest_pic.log; cat gpurun_out/r02y_pytest_pic.log
V=build/variants/libminipic_nomagic.so
for k in 1 2; do for r in sat quiet thermal c4; do
  timeout 300 python tools/push_ab.py $r | sed 's/^{/{"build": "fd_patch", /' >> gpurun_out/r02y_ab.jsonl
  MPIC_LIB=$V timeout 300 python tools/push_ab.py $r | sed 's/^{/{"build": "before", /' >> gpurun_out/r02y_ab.jsonl
done; done
cat gpurun_out/r02y_ab.jsonl
