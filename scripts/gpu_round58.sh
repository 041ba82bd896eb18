# final gate of the session: whole GPU suite, smoke, launch list of one clustered top-100 search (share per kernel)
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"
tail -2 gpurun_out/smoke.log
export NVDB_K=100 NVDB_BATCHES=1024 NVDB_DTYPES=f16 NVDB_KINDS=clustered
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --kernel-name-base demangled -k 'regex:scan_|merge_|rescore|prep_|gather_queries|scatter_topk|row_norm' -c 40 --csv --log-file gpurun_out/launches_clustered_k100.csv python scripts/bench_clustered.py > gpurun_out/ncu_list.log 2>&1; echo "ncu list rc=$?"
echo done
