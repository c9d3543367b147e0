set -x
timeout 900 python -m pytest tests/test_gpu_pic.py -x -q 2>&1 | tail -15
timeout 900 python -m pytest tests/test_gpu_slab.py tests/test_gpu_driver.py -x -q 2>&1 | tail -15
for r in quiet sat thermal; do timeout 300 python tools/push_ab.py $r; done
