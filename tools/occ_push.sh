for pad in 0 20000 30000 50000 90000; do
  echo "== pad=$pad"
  MPIC_PUSH_PAD_SMEM=$pad timeout 100 python bench.py --steps 10 --warmup 3 --no-yee --no-cpu --no-e2e | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['phases_ms']['push_deposit'])"
done
