set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
NVDB_IVF_OPQ=25 timeout 1700 python scripts/bench_ivfpq.py > gpurun_out/ivfpq_1M_opq.jsonl 2> gpurun_out/ivfpq_1M_opq.err; echo "rc=$?"
cut -c1-260 gpurun_out/ivfpq_1M_opq.jsonl; tail -5 gpurun_out/ivfpq_1M_opq.err
echo done
