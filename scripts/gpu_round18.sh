set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_scan_mma.py tests/test_gpu_two_stage.py -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -8 gpurun_out/pytest_gpu.log
SPEC="i8:mma:1024:10:pair=0;i8:mma:1024:10:pair=2;i8:mma:1024:10:pair=2,res=0;i8:mma_exact:1024:10:pair=0;i8:mma_exact:1024:10:pair=2;f16:mma_exact:1024:10:pair=0;f16:mma_exact:1024:10:pair=2;f16:mma_fast:1024:10:pair=0;f16:mma_fast:1024:10:pair=2;i8:mma_exact:256:10:pair=0;i8:mma_exact:256:10:pair=2;f16:mma_exact:256:10:pair=0;f16:mma_exact:256:10:pair=2;i8:mma_exact:1024:100:pair=0;i8:mma_exact:1024:100:pair=2"
NVDB_AB_ROUNDS=5 NVDB_AB="$SPEC" timeout 900 python scripts/ab_bench.py > gpurun_out/ab_pair_i8.jsonl 2> gpurun_out/ab_pair_i8.err
python - <<'PY'
import json
for l in open("gpurun_out/ab_pair_i8.jsonl"):
    if l.startswith("{"):
        r=json.loads(l); print(r["dtype"],r["path"],r["batch"],r["k"],"pair",r["knobs"]["pair"],"res",r["knobs"]["res"],"min",r["ms_min"],"med",r["ms_median"],r.get("exact_fallbacks_total"))
PY
tail -3 gpurun_out/ab_pair_i8.err
echo done
