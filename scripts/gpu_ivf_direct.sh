cd $GRAFT_REPO_ROOT
timeout -s KILL 300 python -m pytest tests/test_gpu_ivfpq.py tests/test_host_tools.py -x -q -m gpu 2>&1 | tail -2
E="NVDB_IVF_ROWS=500000 NVDB_IVF_DIM=384 NVDB_IVF_DATA=clustered NVDB_IVF_CLUSTERS=64 NVDB_IVF_REFINE=0,20,50,100"
env $E timeout -s KILL 300 python scripts/bench_ivfpq.py > gpurun_out/r2_ivfpq_direct_out.jsonl 2>&1
tail -2 gpurun_out/r2_ivfpq_direct_out.jsonl | cut -c1-200
