mkdir -p gpurun_out
for ppc in 4 8 16 24; do for gl in 0 1; do
  PPC=$ppc MPIC_PUSH_GL=$gl timeout 300 python tools/push_ab.py thermal 20 1024 1024 >> gpurun_out/r02ak_sparse_ab.jsonl
done; done
cat gpurun_out/r02ak_sparse_ab.jsonl
