# A/B of kernel variants: every libminipic_b200*.so in the package, timing regimes from tools/push_gap.py (GAP_SHORT)
PK=mini-pic--2d-pic-code-with-tiling-load-balance_b200
timeout 600 python -m pytest tests/test_gpu_pic.py tests/test_gpu_slab.py -m gpu -x -q 2>&1 | tail -2
for lib in $PK/libminipic_b200*.so; do
  echo "== $lib"
  MPIC_LIB=$PWD/$lib GAP_SHORT=1 timeout 200 python tools/push_gap.py 2>&1 | cut -c1-300
done | tee gpurun_out/ab_push.log
