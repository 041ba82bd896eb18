set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_scan.py tests/test_gpu_two_stage.py -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
python bench.py --steps 1000 --warmup 20 --no-extras --no-cpu > gpurun_out/bench_short.json 2> gpurun_out/bench_short.err; echo "rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_short.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['e2e'])
PY
echo done
