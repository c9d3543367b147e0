set -x
timeout 900 python -m pytest tests/test_gpu_yee.py tests/test_gpu_pic.py -x -q 2>&1 | tail -5
for n in 8192 2048 512; do
  timeout 300 python tools/yee_ab.py $n >> gpurun_out/r02e_yee.jsonl
  for ty in 8 24 32; do MPIC_LIB=build/variants/libminipic_ty$ty.so timeout 300 python tools/yee_ab.py $n >> gpurun_out/r02e_yee.jsonl; done
done
cat gpurun_out/r02e_yee.jsonl
for r in quiet sat; do timeout 300 python tools/push_ab.py $r; done
