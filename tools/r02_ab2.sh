set -x
OLD=build/old/pkg/libminipic_b200.so
V=build/variants
run() { MPIC_LIB=$1 MPIC_PUSH_GL=$2 timeout 300 python tools/push_ab.py $3 | sed "s|^{|{\"lib\": \"$4\", |" >> gpurun_out/r02c_ab.jsonl; }
for r in quiet thermal sat; do
  run $OLD 0 $r old
  run "" 16 $r new16
  run "" 32 $r new32
  run $V/libminipic_b384.so 16 $r b384_16
  run $V/libminipic_r72.so 16 $r r72_16
done
cat gpurun_out/r02c_ab.jsonl
