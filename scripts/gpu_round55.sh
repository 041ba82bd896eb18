# A/B: width of the second tier's candidate lists (clustered FP16 base, batch 1024)
set -x
mkdir -p gpurun_out
: > gpurun_out/ab_tier2_width.jsonl
for w in 133 141 150 158; do
  echo "{\"k\": 100, \"NVDB_B200_EXACT_TIER2\": $w}" >> gpurun_out/ab_tier2_width.jsonl
  NVDB_B200_EXACT_TIER2=$w NVDB_K=100 NVDB_BATCHES=1024 NVDB_DTYPES=f16 NVDB_KINDS=clustered timeout 300 python scripts/bench_clustered.py >> gpurun_out/ab_tier2_width.jsonl 2> gpurun_out/ab.err || tail -3 gpurun_out/ab.err
done
for w in 24 32 48 64; do
  echo "{\"k\": 10, \"NVDB_B200_EXACT_TIER2\": $w}" >> gpurun_out/ab_tier2_width.jsonl
  NVDB_B200_EXACT_TIER2=$w NVDB_K=10 NVDB_BATCHES=128,1024 NVDB_DTYPES=f16 NVDB_KINDS=clustered timeout 300 python scripts/bench_clustered.py >> gpurun_out/ab_tier2_width.jsonl 2> gpurun_out/ab.err || tail -3 gpurun_out/ab.err
done
cut -c1-330 gpurun_out/ab_tier2_width.jsonl
echo done
