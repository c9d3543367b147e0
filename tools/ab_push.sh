set -x
timeout 300 python -m pytest tests/test_gpu_pic.py -m gpu -x -q 2>&1 | tail -3
PK=mini-pic--2d-pic-code-with-tiling-load-balance_b200
for lib in libminipic_b200.so libminipic_b200_c4.so libminipic_b200_c6.so; do
 for sp in 2 4 8; do
  echo "== $lib split=$sp"
  MPIC_LIB=$PWD/$PK/$lib MPIC_PUSH_SPLIT=$sp timeout 100 python bench.py --steps 10 --warmup 3 --no-yee --no-cpu --no-e2e | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['phases_ms'], d['roofline']['frac'])"
 done
done
