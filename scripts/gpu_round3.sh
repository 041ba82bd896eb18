# round-1 validation pass of the whole tree + ncu captures of the tensor-core scan and the INT8 streaming scan
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
tail -3 gpurun_out/smoke.log
timeout 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?" >> gpurun_out/bench.err
timeout 900 python scripts/bench_sweep.py > gpurun_out/sweep.jsonl 2> gpurun_out/sweep.err; echo "sweep rc=$?"
NCU="ncu --set full --clock-control none --import-source on --kernel-name-base demangled"
NVDB_BATCHES=1024 timeout 600 $NCU -k regex:scan_mma_kernel -c 6 -o gpurun_out/prof_mma_b1024 python scripts/bench_sweep.py > gpurun_out/ncu_mma.log 2>&1
NVDB_BATCHES=1 timeout 600 $NCU -k 'regex:scan_gemv_tma_kernel<3' -s 3 -c 2 -o gpurun_out/prof_gemv_i8 python scripts/bench_sweep.py > gpurun_out/ncu_i8.log 2>&1
for f in prof_mma_b1024 prof_gemv_i8; do
  ncu -i gpurun_out/$f.ncu-rep --page raw --csv > gpurun_out/$f.raw.csv 2>/dev/null
  ncu -i gpurun_out/$f.ncu-rep --page details > gpurun_out/$f.details.txt 2>/dev/null
done
ls -la gpurun_out
echo done
