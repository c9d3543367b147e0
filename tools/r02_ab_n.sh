V=build/variants
run() { env MPIC_LIB=$1 MPIC_PUSH_GL=1 $4 timeout 300 python tools/push_ab.py $2 | sed "s|^{|{\"lib\": \"$3\", |" >> gpurun_out/r02n_ab.jsonl; }
for r in quiet sat; do
  run $V/libminipic_nostash.so $r nostash
  run $V/libminipic_nostash.so $r nostash_pad16k MPIC_PUSH_PAD_SMEM=16384
  run $V/libminipic_nostash.so $r nostash_pad8k MPIC_PUSH_PAD_SMEM=8192
  run $V/libminipic_stash.so $r stash
done
cat gpurun_out/r02n_ab.jsonl
