set -x
mkdir -p gpurun_out
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --kernel-name-base demangled -k 'regex:scan_|merge_|rescore|prep_' -c 60 --csv --log-file gpurun_out/launches_bench.csv python bench.py --steps 20 --warmup 3 --no-extras --no-cpu > gpurun_out/ncu_bench.log 2>&1
tail -3 gpurun_out/launches_bench.csv | cut -c1-300
echo done
