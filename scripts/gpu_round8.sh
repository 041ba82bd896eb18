set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -8 gpurun_out/pytest_gpu.log
NVDB_AB="f16:mma_fast:1024:10:pair=0,seed=0;f16:mma_fast:1024:10:pair=0,seed=1;f16:mma_fast:1024:10:pair=1,seed=0;f16:mma_fast:1024:10:pair=1,seed=1;f16:mma_exact:1024:10:pair=0,seed=1;f16:mma_exact:1024:10:pair=1,seed=1;f16:mma_exact:1024:10:pair=1,seed=0;f16:mma_exact:512:10;f16:mma_exact:256:10;f16:mma_exact:128:10;f16:mma_fast:128:10;f16:mma_exact:16:10;f16:gemv:2:10;f16:gemv:4:10;f16:mma_exact:4:10;i8:mma_exact:1024:10:seed=0;i8:mma_exact:1024:10:seed=1;i8:mma:1024:10:seed=1;i8:mma_exact:128:10;f16:mma_exact:1024:100:pair=1;f16:mma_exact:1024:100:pair=0;f16:mma_fast:1024:100:pair=1" timeout 1200 python scripts/ab_bench.py > gpurun_out/ab_round8.jsonl 2> gpurun_out/ab_round8.err
cat gpurun_out/ab_round8.jsonl; tail -3 gpurun_out/ab_round8.err
echo done
