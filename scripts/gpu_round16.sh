set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_scan_mma.py tests/test_gpu_two_stage.py -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -8 gpurun_out/pytest_gpu.log
SPEC="i8:mma:1024:10;i8:mma_exact:1024:10;f16:mma_fast:1024:10;f16:mma_exact:1024:10;f16:mma:1024:10;f16:mma_exact:1024:100;i8:mma_exact:128:10;f16:mma_exact:128:10;f16:mma_fast:1024:10:pair=2"
rm -f gpurun_out/ab_elect_new.jsonl gpurun_out/ab_elect_prev.jsonl
for r in 1 2 3; do
  NVDB_B200_LIB=$PWD/nano-vectordb_b200/lib/libnvdb_b200_prev.so NVDB_AB_ROUNDS=3 NVDB_AB="$SPEC" timeout 600 python scripts/ab_bench.py >> gpurun_out/ab_elect_prev.jsonl 2>> gpurun_out/ab_elect.err
  NVDB_AB_ROUNDS=3 NVDB_AB="$SPEC" timeout 600 python scripts/ab_bench.py >> gpurun_out/ab_elect_new.jsonl 2>> gpurun_out/ab_elect.err
done
python - <<'PY'
import json
for tag in ("prev","new"):
    rows=[json.loads(l) for l in open(f"gpurun_out/ab_elect_{tag}.jsonl") if l.startswith("{")]
    agg={}
    for r in rows:
        key=(r.get("dtype"),r.get("path"),r.get("batch"),r.get("k"),json.dumps(r.get("knobs")))
        agg.setdefault(key,[]).append((r.get("ms_min"),r.get("ms_median")))
    for k,v in agg.items(): print(tag,k[:4],k[4][-30:],[x[0] for x in v],[x[1] for x in v])
PY
tail -3 gpurun_out/ab_elect.err
echo done
