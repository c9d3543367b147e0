set -x
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -15
for r in quiet sat thermal c4; do timeout 300 python tools/push_ab.py $r >> gpurun_out/r02i_ab.jsonl; done
cat gpurun_out/r02i_ab.jsonl
bash tools/r02_launches.sh
