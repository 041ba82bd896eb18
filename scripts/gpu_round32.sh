set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 600 python scripts/ivfpq_latency_probe.py 2>&1 | tail -7
timeout 1700 python scripts/bench_ivfpq.py > gpurun_out/ivfpq_1M.jsonl 2> gpurun_out/ivfpq_1M.err; echo "rc=$?"
cut -c1-330 gpurun_out/ivfpq_1M.jsonl; tail -5 gpurun_out/ivfpq_1M.err
timeout 1200 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
head -c 300 gpurun_out/bench.json; tail -3 gpurun_out/bench.err
echo done
