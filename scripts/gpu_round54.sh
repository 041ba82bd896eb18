# 2-GPU check after the exact-path change: sharded tests (two processes, peer exchange), bench line, K4-shaped run
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_sharded.py -q > gpurun_out/pytest_sharded_n2.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_sharded_n2.log
tail -5 gpurun_out/pytest_sharded_n2.log
NVDB_GPUS=2 bash scripts/gpu_multi.sh
