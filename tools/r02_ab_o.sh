V=build/variants
OUT=gpurun_out/$1; shift
run() { env MPIC_LIB=$V/libminipic_$1.so MPIC_PUSH_GL=1 $3 timeout 300 python tools/push_ab.py $2 | sed "s|^{|{\"lib\": \"$1 $3\", |" >> $OUT; }
for v in "$@"; do
  MPIC_LIB=$V/libminipic_$v.so timeout 600 python -m pytest tests/test_gpu_pic.py -x -q -k "packed or stable" 2>&1 | tail -1
  for r in quiet sat thermal; do run $v $r; done
done
cat $OUT
