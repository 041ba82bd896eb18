# radix-select merge: scan / tensor-scan / sharded suites, bench extras (top-100 batch 1)
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_scan.py tests/test_gpu_scan_mma.py tests/test_gpu_sharded.py tests/test_gpu_two_stage.py -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -8 gpurun_out/pytest_gpu.log
timeout 900 python bench.py --no-cpu --no-ivfpq > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?"
tail -3 gpurun_out/bench_n1.err
echo done
