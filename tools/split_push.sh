for sp in 1 2 3 4 6 8; do
  echo "== split=$sp"
  MPIC_PUSH_SPLIT=$sp GAP_SHORT=1 timeout 200 python tools/push_gap.py 2>&1 | sed 's/"push_ms_first10.*total_ms/ total_ms/' | cut -c1-170
done | tee gpurun_out/split_push.log
