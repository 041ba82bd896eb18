set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_host_tools.py tests/test_host_bench.py -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -30 gpurun_out/pytest_gpu.log
echo done
