# magic-number conversions in the packed push: parity, then A/B against the F2I build; plus the re-run of the fixed tests
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_pic.py -x -q 2>&1 | tail -4 > gpurun_out/r02x_pytest_pic.log; cat gpurun_out/r02x_pytest_pic.log
V=build/variants/libminipic_nomagic.so
for r in sat quiet thermal c4; do
  timeout 300 python tools/push_ab.py $r | sed 's/^{/{"build": "magic", /' >> gpurun_out/r02x_ab.jsonl
  MPIC_LIB=$V timeout 300 python tools/push_ab.py $r | sed 's/^{/{"build": "f2i", /' >> gpurun_out/r02x_ab.jsonl
done
for r in sat; do
  timeout 300 python tools/push_ab.py $r | sed 's/^{/{"build": "magic", /' >> gpurun_out/r02x_ab.jsonl
  MPIC_LIB=$V timeout 300 python tools/push_ab.py $r | sed 's/^{/{"build": "f2i", /' >> gpurun_out/r02x_ab.jsonl
done
cat gpurun_out/r02x_ab.jsonl
timeout 1200 python -m pytest tests/test_gpu_slab.py tests/test_gpu_driver.py -x -q 2>&1 | tail -15 > gpurun_out/r02x_pytest.log; cat gpurun_out/r02x_pytest.log
timeout 600 python -m pytest tests/test_gpu_fullsize.py -x -q -s -k "65536" 2>&1 | tail -8 > gpurun_out/r02x_fullsize.log; cat gpurun_out/r02x_fullsize.log
