set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_scan_mma.py -m gpu -q -x -k "cta_pair or seeding or auto_path" > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
NVDB_AB_ROUNDS=2 NVDB_AB="f16:mma_exact:1024:10;f16:mma_fast:1024:10;i8:mma_exact:1024:10;f16:mma_exact:128:10" timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_exact.csv python scripts/ab_bench.py > gpurun_out/ncu_exact.log 2>&1
tail -5 gpurun_out/ncu_exact.log
echo done
