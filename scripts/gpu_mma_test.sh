set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_scan_mma.py -x -q > gpurun_out/pytest_mma.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_mma.log
tail -30 gpurun_out/pytest_mma.log
