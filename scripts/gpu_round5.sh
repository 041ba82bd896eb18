# exact tensor-core path: parity + sweep
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -30 gpurun_out/pytest_gpu.log
NVDB_PATHS=mma_fast,mma_exact NVDB_BATCHES=4,16,128,256,1024 timeout 900 python scripts/bench_sweep.py > gpurun_out/sweep_exact.jsonl 2> gpurun_out/sweep_exact.err
cat gpurun_out/sweep_exact.jsonl; tail -5 gpurun_out/sweep_exact.err
NVDB_K=100 NVDB_DTYPES=f16 NVDB_PATHS=gemv,mma,mma_exact NVDB_BATCHES=1,8,128,1024 timeout 900 python scripts/bench_sweep.py > gpurun_out/sweep_k100.jsonl 2> gpurun_out/sweep_k100.err
cat gpurun_out/sweep_k100.jsonl; tail -5 gpurun_out/sweep_k100.err
echo done
