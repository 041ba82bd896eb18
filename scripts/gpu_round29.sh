set -x
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --cache-control none --import-source on --kernel-name-base demangled -k 'regex:ivfpq_scan' -s 20 -c 1 -o gpurun_out/prof_ivf_scan python scripts/ivfpq_latency_probe.py > gpurun_out/ivf_ncu.log 2>&1
ncu -i gpurun_out/prof_ivf_scan.ncu-rep --page source --csv > gpurun_out/prof_ivf_scan.source.csv 2>/dev/null
ncu -i gpurun_out/prof_ivf_scan.ncu-rep --page details > gpurun_out/prof_ivf_scan.details.txt 2>/dev/null
echo done
