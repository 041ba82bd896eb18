set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_scan_mma.py -m gpu -q -x -k "pair or resident" > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -8 gpurun_out/pytest_gpu.log
SPEC="i8:mma:1024:10:res=2;i8:mma:1024:10:res=1;i8:mma_exact:1024:10:res=2;i8:mma_exact:1024:10:res=1;i8:mma_exact:256:10:res=2;i8:mma_exact:256:10:res=1;i8:mma_exact:512:10:res=2;i8:mma_exact:512:10:res=1"
NVDB_AB_ROUNDS=5 NVDB_AB="$SPEC" timeout 900 python scripts/ab_bench.py > gpurun_out/ab_res2.jsonl 2> gpurun_out/ab_res2.err
python - <<'PY'
import json
for l in open("gpurun_out/ab_res2.jsonl"):
    if l.startswith("{"):
        r=json.loads(l); print(r["dtype"],r["path"],r["batch"],r["k"],"pair",r["knobs"]["pair"],"res",r["knobs"]["res"],"min",r["ms_min"],"med",r["ms_median"],r.get("exact_fallbacks_total"))
PY
tail -3 gpurun_out/ab_res2.err
echo done
