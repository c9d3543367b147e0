for sp in 2 3 5 6 7 13; do MPIC_PUSH_SPLIT=$sp timeout 300 python bench.py --steps 20 --warmup 5 --no-yee --no-cpu --no-e2e 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read());print('split $sp', round(d['ms_per_step'],4), round(d['phases_ms']['push_deposit'],4))"; done
