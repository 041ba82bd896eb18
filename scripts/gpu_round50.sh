# clustered-distribution parity first (new), then the whole GPU suite
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_clustered.py -q > gpurun_out/pytest_clustered.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_clustered.log
tail -15 gpurun_out/pytest_clustered.log
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -8 gpurun_out/pytest_gpu.log
echo done
