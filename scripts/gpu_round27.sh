set -x
mkdir -p gpurun_out
timeout 600 python scripts/ivfpq_latency_probe.py 2>&1 | tail -8
echo done
