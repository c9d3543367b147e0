set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_pic.py tests/test_gpu_slab.py -x -q 2>&1 | tail -4 > gpurun_out/r02ab_pytest.log; cat gpurun_out/r02ab_pytest.log
V=build/variants/libminipic_base.so
for k in 1 2; do for r in sat c4; do
  MPIC_LIB=$V timeout 300 python tools/push_ab.py $r | sed 's/^{/{"build": "previous commit", /' >> gpurun_out/r02ab_ab.jsonl
  timeout 300 python tools/push_ab.py $r | sed 's/^{/{"build": "this", /' >> gpurun_out/r02ab_ab.jsonl
done; done
cat gpurun_out/r02ab_ab.jsonl
