# second certificate tier of the exact path: tests, then clustered vs isotropic base at the K2 / K3 size
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_scan_mma.py tests/test_gpu_clustered.py tests/test_gpu_sharded.py tests/test_gpu_two_stage.py -q -x > gpurun_out/pytest_tier2.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_tier2.log
tail -25 gpurun_out/pytest_tier2.log
timeout 900 python scripts/bench_clustered.py > gpurun_out/clustered_tier2.jsonl 2> gpurun_out/clustered_tier2.err; echo "full rc=$?"
tail -4 gpurun_out/clustered_tier2.err; cat gpurun_out/clustered_tier2.jsonl
NVDB_K=100 NVDB_BATCHES=1,128,1024 timeout 900 python scripts/bench_clustered.py > gpurun_out/clustered_tier2_k100.jsonl 2> gpurun_out/clustered_tier2_k100.err; echo "k100 rc=$?"
cat gpurun_out/clustered_tier2_k100.jsonl
echo done
