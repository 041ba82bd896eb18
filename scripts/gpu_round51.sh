# clustered vs isotropic base at the K2 / K3 size: timings, certificate fallback rates, exact == streaming scan
set -x
mkdir -p gpurun_out
NVDB_BENCH_ROWS=60000 timeout 300 python scripts/bench_clustered.py > gpurun_out/clustered_small.jsonl 2> gpurun_out/clustered_small.err; echo "small rc=$?"
tail -4 gpurun_out/clustered_small.err; cat gpurun_out/clustered_small.jsonl
timeout 900 python scripts/bench_clustered.py > gpurun_out/clustered_full.jsonl 2> gpurun_out/clustered_full.err; echo "full rc=$?"
tail -4 gpurun_out/clustered_full.err; cat gpurun_out/clustered_full.jsonl
NVDB_K=100 NVDB_BATCHES=1,1024 timeout 900 python scripts/bench_clustered.py > gpurun_out/clustered_full_k100.jsonl 2> gpurun_out/clustered_full_k100.err; echo "k100 rc=$?"
cat gpurun_out/clustered_full_k100.jsonl
echo done
