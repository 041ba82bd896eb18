cd $GRAFT_REPO_ROOT
timeout -s KILL 300 python -m pytest tests/test_gpu_ivfpq.py -x -q -m gpu 2>&1 | tail -8
E="NVDB_IVF_ROWS=500000 NVDB_IVF_DIM=384 NVDB_IVF_DATA=clustered NVDB_IVF_CLUSTERS=64 NVDB_IVF_REFINE=0,20,50,100"
env $E timeout -s KILL 300 python scripts/bench_ivfpq.py > gpurun_out/r2_ivfpq_graph_on.jsonl 2>&1
env $E NVDB_B200_IVF_GRAPH=0 timeout -s KILL 300 python scripts/bench_ivfpq.py > gpurun_out/r2_ivfpq_graph_off.jsonl 2>&1
tail -3 gpurun_out/r2_ivfpq_graph_on.jsonl | cut -c1-300
