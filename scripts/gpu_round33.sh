set -x
mkdir -p gpurun_out
python bench.py --steps 20 --warmup 3 --no-extras --no-cpu > gpurun_out/bench_short.json 2> gpurun_out/bench_short.err; echo "rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --kernel-name-base demangled -c 60 --csv --log-file gpurun_out/launches_bench.csv python bench.py --steps 20 --warmup 3 --no-extras --no-cpu > gpurun_out/ncu_bench.log 2>&1
NCU="ncu --set full --clock-control none --import-source on --kernel-name-base demangled"
NVDB_AB_ROUNDS=1 NVDB_AB="i8:mma:1024:10" timeout 600 $NCU -k 'regex:scan_mma2_kernel' -s 3 -c 1 -o gpurun_out/prof_mma2_i8 python scripts/ab_bench.py > gpurun_out/ncu_mma2_i8.log 2>&1
ncu -i gpurun_out/prof_mma2_i8.ncu-rep --page raw --csv > gpurun_out/prof_mma2_i8.raw.csv 2>/dev/null
ncu -i gpurun_out/prof_mma2_i8.ncu-rep --page details > gpurun_out/prof_mma2_i8.details.txt 2>/dev/null
NVDB_AB_ROUNDS=1 NVDB_AB="f16:mma_fast:1024:10" timeout 600 $NCU -k 'regex:scan_mma2_kernel' -s 3 -c 1 -o gpurun_out/prof_mma2_f16 python scripts/ab_bench.py > gpurun_out/ncu_mma2_f16.log 2>&1
ncu -i gpurun_out/prof_mma2_f16.ncu-rep --page raw --csv > gpurun_out/prof_mma2_f16.raw.csv 2>/dev/null
ncu -i gpurun_out/prof_mma2_f16.ncu-rep --page details > gpurun_out/prof_mma2_f16.details.txt 2>/dev/null
grep -E "Duration|SM Frequency" gpurun_out/prof_mma2_i8.details.txt gpurun_out/prof_mma2_f16.details.txt | head
echo done
