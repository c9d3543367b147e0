# launch list (per-kernel durations, serialised by ncu) of steady-state PIC steps: bash tools/r02_launches.sh OUT.csv [prof_push args]
OUT=${1:-gpurun_out/r02h_launches.csv}; shift
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s ${SKIP:-200} -c 60 --csv --log-file $OUT python tools/prof_push.py "$@" > gpurun_out/ncu_l.log 2>&1
python - $OUT <<'PY'
import csv, sys
rows=list(csv.reader(open(sys.argv[1])))
hdr=[i for i,r in enumerate(rows) if r and r[0]=='ID'][0]
h=rows[hdr]; kn=h.index('Kernel Name'); mv=h.index('Metric Value')
agg={}
for r in rows[hdr+1:]:
    if len(r)>mv:
        name=r[kn].split('(')[0][-60:]; agg.setdefault(name,[]).append(float(r[mv].replace(',','')))
for k,v in agg.items(): print(f"{k:60s} n={len(v):3d} mean_us={sum(v)/len(v)/1000:8.2f}")
PY
