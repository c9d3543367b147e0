# host-side floor of the e2e step with N ranks transferring at once
mkdir -p gpurun_out
N=${N:-8}
for k in 1 2 4 8; do
  [ $k -le $N ] || continue
  if [ $k -eq 1 ]; then timeout 300 python tools/pcie_probe.py >> gpurun_out/r02ai_pcie_probe.jsonl
  else timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $k --master-addr 127.0.0.1 --master-port 2952$k tools/pcie_probe.py >> gpurun_out/r02ai_pcie_probe.jsonl 2>> gpurun_out/r02ai_pcie_probe.err; fi
done
cat gpurun_out/r02ai_pcie_probe.jsonl
