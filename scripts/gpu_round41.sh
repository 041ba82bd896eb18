# 8 GPUs: two..eight-process exchange test, bench.py --gpus 8 and K4 at full size, NCCL vs peer-memory exchange on the same box
set -x
mkdir -p gpurun_out
G=${NVDB_GPUS:-8}
nvidia-smi -L > gpurun_out/smi_n$G.txt 2>&1
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29511 tests/mp_sharded_peer.py > gpurun_out/mp_peer_n$G.log 2>&1; echo "mp rc=$?" >> gpurun_out/mp_peer_n$G.log
tail -4 gpurun_out/mp_peer_n$G.log
rm -f gpurun_out/bench_n${G}_*.jsonl gpurun_out/k4_n${G}_*.jsonl
for X in nccl peer; do
  NVDB_B200_EXCHANGE=$X timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $G --steps 1000 --warmup 20 --no-extras --no-cpu >> gpurun_out/bench_n${G}_$X.jsonl 2> gpurun_out/bench_n${G}_$X.err; echo "bench $X rc=$?"
  tail -2 gpurun_out/bench_n${G}_$X.err
done
for X in nccl peer; do
  NVDB_B200_EXCHANGE=$X NVDB_K4_STEPS=20 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29519 scripts/bench_k4.py >> gpurun_out/k4_n${G}_$X.jsonl 2> gpurun_out/k4_n${G}_$X.err; echo "k4 $X rc=$?"
  tail -2 gpurun_out/k4_n${G}_$X.err
done
echo done
