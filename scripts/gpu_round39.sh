# peer-memory exchange: tests (logical ranks on one GPU + two processes over CUDA IPC), then bench.py --gpus 2 A/B nccl vs peer
set -x
mkdir -p gpurun_out
G=${NVDB_GPUS:-2}
nvidia-smi -L > gpurun_out/smi_multi.txt 2>&1
nvidia-smi topo -m >> gpurun_out/smi_multi.txt 2>&1
timeout 900 python -m pytest tests/test_gpu_sharded.py -q -x > gpurun_out/pytest_sharded.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_sharded.log
tail -25 gpurun_out/pytest_sharded.log
for X in nccl peer nccl peer; do
  NVDB_B200_EXCHANGE=$X timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $G --steps 1000 --warmup 20 --no-extras --no-cpu >> gpurun_out/bench_n${G}_$X.jsonl 2> gpurun_out/bench_n${G}_$X.err; echo "bench $X rc=$?"
  tail -1 gpurun_out/bench_n${G}_$X.jsonl | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['e2e']['ms_per_step'], d['gpu_launches'], d['config']['parallelism'])"
  tail -3 gpurun_out/bench_n${G}_$X.err
done
echo done
