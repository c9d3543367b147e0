mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_fullsize.py -m gpu -q -s 2>&1 | tail -40 > gpurun_out/pytest_fullsize.log; cat gpurun_out/pytest_fullsize.log
