set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_ivfpq.py -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -30 gpurun_out/pytest_gpu.log
timeout 600 python scripts/ivfpq_latency_probe.py 2>&1 | tail -3
echo done
