PK=mini-pic--2d-pic-code-with-tiling-load-balance_b200
for lib in $PK/libminipic_b200*.so; do
 for sp in ${SPLITS:-4}; do
  echo "== $lib split=$sp"
  MPIC_LIB=$PWD/$lib MPIC_PUSH_SPLIT=$sp timeout 100 python bench.py --steps 10 --warmup 3 --no-yee --no-cpu --no-e2e | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['phases_ms'], d['roofline']['frac'])"
 done
done
