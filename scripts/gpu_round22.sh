set -x
mkdir -p gpurun_out
timeout 1700 python scripts/bench_ivfpq.py > gpurun_out/ivfpq_1M.jsonl 2> gpurun_out/ivfpq_1M.err; echo "rc=$?"
cat gpurun_out/ivfpq_1M.jsonl; tail -5 gpurun_out/ivfpq_1M.err
echo done
