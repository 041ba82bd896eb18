# full GPU suite (incl. the exchange tests with logical ranks on one GPU and the multi-shard harness), default bench line
set -x
mkdir -p gpurun_out
timeout 1700 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -8 gpurun_out/pytest_gpu.log
timeout 900 python bench.py > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?"
head -c 1200 gpurun_out/bench_n1.json; tail -3 gpurun_out/bench_n1.err
echo done
