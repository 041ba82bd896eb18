set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_ivfpq.py -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -40 gpurun_out/pytest_gpu.log
echo done
