# exchange tail in the scan's last CTA: sharded tests, bench.py --gpus 2 A/B nccl vs peer, N=1 regression check
set -x
mkdir -p gpurun_out
G=${NVDB_GPUS:-2}
timeout 900 python -m pytest tests/test_gpu_sharded.py -q -x > gpurun_out/pytest_sharded.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_sharded.log
tail -25 gpurun_out/pytest_sharded.log
rm -f gpurun_out/bench_n${G}_*.jsonl
for X in nccl peer nccl peer; do
  NVDB_B200_EXCHANGE=$X timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $G --steps 1000 --warmup 20 --no-extras --no-cpu >> gpurun_out/bench_n${G}_$X.jsonl 2> gpurun_out/bench_n${G}_$X.err; echo "bench $X rc=$?"
  tail -3 gpurun_out/bench_n${G}_$X.err
done
timeout 600 python bench.py --steps 1000 --warmup 20 --no-extras --no-cpu > gpurun_out/bench_n1_quick.json 2> gpurun_out/bench_n1_quick.err; echo "bench n1 rc=$?"
echo done
