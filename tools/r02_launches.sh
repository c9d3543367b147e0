set -x
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 200 -c 40 --csv --log-file gpurun_out/r02h_launches.csv python tools/prof_push.py 40 0.05 0.0 > gpurun_out/ncu_l.log 2>&1
python - <<'PY'
import csv
rows=list(csv.reader(open('gpurun_out/r02h_launches.csv')))
hdr=[i for i,r in enumerate(rows) if r and r[0]=='ID'][0]
h=rows[hdr]; kn=h.index('Kernel Name'); mv=h.index('Metric Value')
agg={}
for r in rows[hdr+1:]:
    if len(r)>mv:
        name=r[kn].split('(')[0][-60:]; agg.setdefault(name,[]).append(float(r[mv].replace(',','')))
for k,v in agg.items(): print(f"{k:60s} n={len(v):3d} mean_us={sum(v)/len(v)/1000:8.2f}")
PY
