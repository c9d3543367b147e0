# one tuning iteration of the push kernel on the GPU box: parity, then timing in the bimodal / sustained / thermal regimes
timeout 900 python -m pytest tests/test_gpu_pic.py tests/test_gpu_slab.py -m gpu -x -q 2>&1 | tail -3
GAP_SHORT=1 timeout 300 python tools/push_gap.py 2>&1 | sed 's/"push_ms_first10.*total_ms/ total_ms/' | cut -c1-170 | tee gpurun_out/push_gap.log
