V=build/variants
OUT=gpurun_out/r02p_ab.jsonl
run() { env MPIC_LIB=$V/libminipic_$1.so MPIC_PUSH_GL=1 $3 timeout 300 python tools/push_ab.py $2 | sed "s|^{|{\"lib\": \"$1 $3\", |" >> $OUT; }
for r in quiet sat; do
  run jp0 $r
  run jp0 $r MPIC_PUSH_SPLIT=4
  run jp0 $r MPIC_PUSH_SPLIT=8
  run jp0w9 $r
  run jp0w9 $r MPIC_PUSH_SPLIT=4
  run jp0b384 $r
  run jp0b384 $r MPIC_PUSH_SPLIT=4
  run jp0b512 $r
done
cat $OUT
