# heap lists + threshold seeding: parity + sweeps (seed on/off), k = 10 and k = 100
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -30 gpurun_out/pytest_gpu.log
for s in 1 0; do
NVDB_B200_MMA_SEED=$s NVDB_PATHS=mma,mma_fast,mma_exact NVDB_BATCHES=16,128,256,1024 timeout 900 python scripts/bench_sweep.py > gpurun_out/sweep_seed$s.jsonl 2> gpurun_out/sweep_seed$s.err
cat gpurun_out/sweep_seed$s.jsonl; tail -3 gpurun_out/sweep_seed$s.err
done
NVDB_K=100 NVDB_DTYPES=f16 NVDB_PATHS=gemv,mma_fast,mma_exact NVDB_BATCHES=1,8,128,1024 timeout 900 python scripts/bench_sweep.py > gpurun_out/sweep_k100.jsonl 2> gpurun_out/sweep_k100.err
cat gpurun_out/sweep_k100.jsonl; tail -5 gpurun_out/sweep_k100.err
echo done
