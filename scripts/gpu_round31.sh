set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_ivfpq.py -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -30 gpurun_out/pytest_gpu.log
timeout 600 python scripts/ivfpq_latency_probe.py 2>&1 | tail -8
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none --kernel-name-base demangled -k 'regex:dot_bias_small|select_probes|ivfpq_|merge_partials|finalize_cand|refine' -s 200 -c 12 --csv --log-file gpurun_out/ivf_launches.csv python scripts/ivfpq_latency_probe.py > gpurun_out/ivf_ncu.log 2>&1
python - <<'PY'
import csv
rows=list(csv.reader(open('gpurun_out/ivf_launches.csv')))
hi=[i for i,r in enumerate(rows) if 'Kernel Name' in r][0]
h=rows[hi]; kn=h.index('Kernel Name'); mv=h.index('Metric Value'); g=h.index('Grid Size')
for r in rows[hi+2:]:
    if len(r)>mv: print(r[kn].split('(')[0][-28:], r[g], float(r[mv].replace(',',''))/1e3)
PY


echo done
