# ncu of one clustered batch-1024 top-100 search through the exact path: launch list (share per kernel) + full capture
# of the tensor-core scan launches (first tier k' = 125, second tier k' = 158)
set -x
mkdir -p gpurun_out
export NVDB_K=100 NVDB_BATCHES=1024 NVDB_DTYPES=f16 NVDB_KINDS=clustered
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --kernel-name-base demangled -k 'regex:scan_|merge_|rescore|prep_|gather_queries|scatter_topk|row_norm' -c 40 --csv --log-file gpurun_out/launches_clustered_k100.csv python scripts/bench_clustered.py > gpurun_out/ncu_list.log 2>&1; echo "ncu list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k 'regex:scan_mma' -c 4 -o gpurun_out/clustered_k100_scan_mma python scripts/bench_clustered.py > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"
ncu -i gpurun_out/clustered_k100_scan_mma.ncu-rep --page details --csv > gpurun_out/clustered_k100_scan_mma_details.csv 2>/dev/null
ls -la gpurun_out/*.ncu-rep
tail -3 gpurun_out/ncu_full.log
echo done
