set -x
OLD=build/old/pkg/libminipic_b200.so
V=build/variants
run() { MPIC_LIB=$1 MPIC_PUSH_GL=$2 timeout 300 python tools/push_ab.py $3 | sed "s|^{|{\"lib\": \"$4\", |" >> gpurun_out/r02d_ab.jsonl; }
for r in quiet thermal; do
  run $OLD 0 $r old
  for v in allold wrap mover clamp; do run $V/libminipic_$v.so 32 $r ${v}_32; run $V/libminipic_$v.so 16 $r ${v}_16; done
done
