# final round-1 records: default bench line, ncu launch list of the bench command, ncu of merge_select_kernel
set -x
mkdir -p gpurun_out
timeout 900 python bench.py > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?"
tail -2 gpurun_out/bench_n1.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --kernel-name-base demangled -c 60 --csv --log-file gpurun_out/launches_bench.csv python bench.py --steps 20 --warmup 3 --no-extras --no-cpu > gpurun_out/ncu_bench.log 2>&1; echo "ncu bench rc=$?"
timeout 300 python scripts/merge_select_probe.py > gpurun_out/probe.log 2>&1; echo "probe rc=$?"; tail -2 gpurun_out/probe.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --kernel-name-base demangled -k 'regex:scan_gemv|merge_|scan_mma|rescore|prep_queries' -c 60 --csv --log-file gpurun_out/launches_merge_select.csv python scripts/merge_select_probe.py > gpurun_out/ncu_probe.log 2>&1; echo "ncu probe rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k 'regex:merge_select' -s 4 -c 1 -o gpurun_out/prof_merge_select python scripts/merge_select_probe.py > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"
ncu -i gpurun_out/prof_merge_select.ncu-rep --page raw --csv > gpurun_out/prof_merge_select.raw.csv 2>/dev/null
ncu -i gpurun_out/prof_merge_select.ncu-rep --page details > gpurun_out/prof_merge_select.details.txt 2>/dev/null
rm -f gpurun_out/prof_merge_select.ncu-rep
echo done
