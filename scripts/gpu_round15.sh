set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -12 gpurun_out/pytest_gpu.log
NVDB_AB="i8:mma:1024:10:res=0;i8:mma:1024:10:res=1;i8:mma_exact:1024:10:res=0;i8:mma_exact:1024:10:res=1;i8:mma_exact:128:10:res=0;i8:mma_exact:128:10:res=1;i8:mma:256:10:res=0;i8:mma:256:10:res=1" timeout 900 python scripts/ab_bench.py > gpurun_out/ab_res.jsonl 2> gpurun_out/ab_res.err
cat gpurun_out/ab_res.jsonl | cut -c1-70,95-250; tail -3 gpurun_out/ab_res.err
echo done
