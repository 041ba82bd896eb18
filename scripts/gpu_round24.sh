set -x
mkdir -p gpurun_out
timeout 600 python scripts/ivfpq_latency_probe.py > gpurun_out/ivf_probe.log 2>&1; cat gpurun_out/ivf_probe.log

python - <<'PY'
import csv
rows=list(csv.reader(open('gpurun_out/ivf_launches.csv')))
hi=[i for i,r in enumerate(rows) if 'Kernel Name' in r][0]
h=rows[hi]; kn=h.index('Kernel Name'); mv=h.index('Metric Value'); g=h.index('Grid Size') if 'Grid Size' in h else None
agg={}
for r in rows[hi+2:]:
    if len(r)>mv: agg.setdefault(r[kn][:60]+(' grid '+r[g] if g else ''),[]).append(float(r[mv].replace(',','')))
for n,v in agg.items(): print(f"{n:90s} n={len(v):3d} mean={sum(v)/len(v)/1e3:8.2f} us")
PY
echo done
