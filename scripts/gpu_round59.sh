# de-flake check: the oversubscribed exchange tests of the host harness, five times; then the whole GPU suite
set -x
mkdir -p gpurun_out
for i in 1 2 3 4 5; do timeout 300 python -m pytest tests/test_host_bench.py -q -m gpu 2>&1 | tail -2; done > gpurun_out/pytest_host_bench_x5.log 2>&1
cat gpurun_out/pytest_host_bench_x5.log
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
echo done
