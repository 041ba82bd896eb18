# CTA-pair (cta_group::2) kernel: first run, guarded by short timeouts
set -x
mkdir -p gpurun_out
timeout 240 python -m pytest tests/test_gpu_scan_mma.py -q -x -k "cta_pair" > gpurun_out/pytest_pair.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_pair.log
tail -30 gpurun_out/pytest_pair.log
if grep -q "rc=0" gpurun_out/pytest_pair.log; then
NVDB_AB="f16:mma_fast:1024:10:pair=0;f16:mma_fast:1024:10:pair=1;f16:mma_exact:1024:10:pair=0;f16:mma_exact:1024:10:pair=1;f16:mma:1024:10:pair=0;f16:mma:1024:10:pair=1;f16:mma_exact:256:10:pair=0;f16:mma_exact:256:10:pair=1;f16:mma_exact:1024:10:pair=1,seed=0;f16:mma_exact:128:10:seed=0;f16:mma_exact:128:10:seed=1;f16:mma_fast:1024:100:pair=0;f16:mma_fast:1024:100:pair=1;f16:mma_exact:1024:100:pair=1" timeout 900 python scripts/ab_bench.py > gpurun_out/ab_pair.jsonl 2> gpurun_out/ab_pair.err
cat gpurun_out/ab_pair.jsonl; tail -3 gpurun_out/ab_pair.err
fi
echo done
