# packed-FP32 (FFMA2) streaming scan: parity + 1 vs 2 CTAs per SM + ncu of the INT8 kernel
set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -8 gpurun_out/pytest_gpu.log
for c in 1 2; do
  NVDB_B200_GEMV_CTAS=$c NVDB_PATHS=gemv NVDB_BATCHES=1,2,4,8 timeout 600 python scripts/bench_sweep.py > gpurun_out/gemv_ctas$c.jsonl 2> gpurun_out/gemv_ctas$c.err
  cat gpurun_out/gemv_ctas$c.jsonl
done
NCU="ncu --set full --clock-control none --import-source on --kernel-name-base demangled"
NVDB_PATHS=gemv NVDB_DTYPES=i8 NVDB_BATCHES=1 timeout 600 $NCU -k 'regex:scan_gemv_tma_kernel<\(int\)3' -s 3 -c 2 -o gpurun_out/prof_gemv_i8 python scripts/bench_sweep.py > gpurun_out/ncu_i8.log 2>&1
tail -3 gpurun_out/ncu_i8.log
ncu -i gpurun_out/prof_gemv_i8.ncu-rep --page raw --csv > gpurun_out/prof_gemv_i8.raw.csv 2>/dev/null
ncu -i gpurun_out/prof_gemv_i8.ncu-rep --page details > gpurun_out/prof_gemv_i8.details.txt 2>/dev/null
echo done
