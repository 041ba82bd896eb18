set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x -k "exact or host_bench or sharded" > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
NCU="ncu --set full --clock-control none --import-source on --kernel-name-base demangled"
NVDB_AB_ROUNDS=1 NVDB_AB="i8:mma:1024:10" timeout 600 $NCU -k 'regex:scan_mma_kernel<\(int\)3' -s 3 -c 2 -o gpurun_out/prof_mma_i8 python scripts/ab_bench.py > gpurun_out/ncu_mma_i8.log 2>&1
ncu -i gpurun_out/prof_mma_i8.ncu-rep --page raw --csv > gpurun_out/prof_mma_i8.raw.csv 2>/dev/null
ncu -i gpurun_out/prof_mma_i8.ncu-rep --page details > gpurun_out/prof_mma_i8.details.txt 2>/dev/null
NVDB_AB_ROUNDS=1 NVDB_AB="f16:mma_fast:1024:10" timeout 600 $NCU -k 'regex:scan_mma_kernel<\(int\)2' -s 4 -c 1 -o gpurun_out/prof_mma_f16_1limb python scripts/ab_bench.py > gpurun_out/ncu_mma_f16.log 2>&1
ncu -i gpurun_out/prof_mma_f16_1limb.ncu-rep --page raw --csv > gpurun_out/prof_mma_f16_1limb.raw.csv 2>/dev/null
ncu -i gpurun_out/prof_mma_f16_1limb.ncu-rep --page details > gpurun_out/prof_mma_f16_1limb.details.txt 2>/dev/null
python bench.py --steps 200 --warmup 5 --no-extras --no-cpu > gpurun_out/bench_short.json 2> gpurun_out/bench_short.err; head -c 200 gpurun_out/bench_short.json
echo done
