set -x
mkdir -p gpurun_out
NCU="ncu --set full --clock-control none --import-source on --kernel-name-base demangled"
NVDB_AB_ROUNDS=1 NVDB_AB="i8:mma:1024:10" timeout 600 $NCU -k 'regex:scan_mma_kernel<\(int\)3' -s 3 -c 1 -o gpurun_out/prof_mma_i8_v2 python scripts/ab_bench.py > gpurun_out/ncu_mma_i8.log 2>&1
ncu -i gpurun_out/prof_mma_i8_v2.ncu-rep --page raw --csv > gpurun_out/prof_mma_i8_v2.raw.csv 2>/dev/null
ncu -i gpurun_out/prof_mma_i8_v2.ncu-rep --page details > gpurun_out/prof_mma_i8_v2.details.txt 2>/dev/null
ncu -i gpurun_out/prof_mma_i8_v2.ncu-rep --page source --csv > gpurun_out/prof_mma_i8_v2.source.csv 2>/dev/null
NVDB_AB_ROUNDS=1 NVDB_AB="f16:mma_fast:1024:10" timeout 600 $NCU -k 'regex:scan_mma_kernel<\(int\)2' -s 4 -c 1 -o gpurun_out/prof_mma_f16_v2 python scripts/ab_bench.py > gpurun_out/ncu_mma_f16.log 2>&1
ncu -i gpurun_out/prof_mma_f16_v2.ncu-rep --page raw --csv > gpurun_out/prof_mma_f16_v2.raw.csv 2>/dev/null
ncu -i gpurun_out/prof_mma_f16_v2.ncu-rep --page details > gpurun_out/prof_mma_f16_v2.details.txt 2>/dev/null
ncu -i gpurun_out/prof_mma_f16_v2.ncu-rep --page source --csv > gpurun_out/prof_mma_f16_v2.source.csv 2>/dev/null
echo done
