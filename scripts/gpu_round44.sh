# odd numbers of query tiles: single-CTA kernels vs one more (empty) tile + CTA-pair kernel
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_scan_mma.py -q -x -k "pair" > gpurun_out/pytest_pair.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_pair.log
tail -4 gpurun_out/pytest_pair.log
NVDB_AB_ROUNDS=5 NVDB_AB="f16:mma_exact:300:10:pair=1;f16:mma_exact:300:10:pair=2;f16:mma_exact:384:10:pair=1;f16:mma_exact:384:10:pair=2;f16:mma_exact:640:10:pair=1;f16:mma_exact:640:10:pair=2;f16:mma_exact:896:10:pair=1;f16:mma_exact:896:10:pair=2;i8:mma_exact:384:10:pair=1;i8:mma_exact:384:10:pair=2;i8:mma_exact:640:10:pair=1;i8:mma_exact:640:10:pair=2;i8:mma_exact:896:10:pair=1;i8:mma_exact:896:10:pair=2;f16:mma_exact:384:100:pair=1;f16:mma_exact:384:100:pair=2" timeout 900 python scripts/ab_bench.py > gpurun_out/ab_pair_pad_odd.jsonl 2> gpurun_out/ab_pair_pad_odd.err; echo "ab rc=$?"
cat gpurun_out/ab_pair_pad_odd.jsonl | cut -c1-220
echo done
