mkdir -p gpurun_out
for ppc in 8 16; do for w in 24 1; do
  PPC=$ppc MPIC_PERM_WARP_MIN_PPC=$w timeout 300 python tools/push_ab.py thermal 20 1024 1024 | sed "s/^{/{\"perm_warp_min_ppc\": $w, /" >> gpurun_out/r02al_perm_sparse_ab.jsonl
done; done
cat gpurun_out/r02al_perm_sparse_ab.jsonl
