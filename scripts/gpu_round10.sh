set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -8 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log; tail -2 gpurun_out/smoke.log
NVDB_AB="f16:gemv:1:10:fuse=1;f16:gemv:1:10:fuse=0;i8:gemv:1:10:fuse=1;i8:gemv:1:10:fuse=0;f16:gemv:1:100:fuse=1;f16:gemv:1:100:fuse=0;f16:gemv:2:10:fuse=1;f16:gemv:2:10:fuse=0" timeout 600 python scripts/ab_bench.py > gpurun_out/ab_fuse.jsonl 2> gpurun_out/ab_fuse.err
cat gpurun_out/ab_fuse.jsonl; tail -3 gpurun_out/ab_fuse.err
NVDB_AB_ROUNDS=2 NVDB_AB="f16:mma_exact:1024:10;f16:mma_fast:1024:10;i8:mma_exact:1024:10;f16:mma_exact:128:10" timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k 'regex:scan_|merge_|rescore|prep_' -c 400 --csv --log-file gpurun_out/launches_exact.csv python scripts/ab_bench.py > gpurun_out/ncu_exact.log 2>&1
tail -3 gpurun_out/ncu_exact.log
timeout 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
echo done
