# bench.py with the clustered extra (must still print its line), then the second-tier width A/B
set -x
mkdir -p gpurun_out
timeout 600 python bench.py > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?"
tail -3 gpurun_out/bench_n1.err
python -c "
import json; d=json.load(open('gpurun_out/bench_n1.json')); print(d['value'], d['e2e']['value'], d['extras'].get('fp16_clustered_exact_b1024'))"
bash scripts/gpu_round55.sh
