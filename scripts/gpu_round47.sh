# K4 strong scaling: the SAME 100 M x 768 FP16 index (153.6 GB) on G GPUs; G=1 holds it in one B200's 180 GB
set -x
mkdir -p gpurun_out
G=${NVDB_GPUS:-1}
ROWS=$((100000000 / G))
nvidia-smi --query-gpu=memory.total,memory.used --format=csv > gpurun_out/k4_strong_n$G.smi
if [ "$G" = "1" ]; then
  NVDB_K4_ROWS_PER_GPU=$ROWS NVDB_K4_STEPS=5 timeout 900 python scripts/bench_k4.py > gpurun_out/k4_strong_n$G.json 2> gpurun_out/k4_strong_n$G.err; echo "k4 rc=$?"
else
  NVDB_K4_ROWS_PER_GPU=$ROWS NVDB_K4_STEPS=10 timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29519 scripts/bench_k4.py > gpurun_out/k4_strong_n$G.json 2> gpurun_out/k4_strong_n$G.err; echo "k4 rc=$?"
fi
cat gpurun_out/k4_strong_n$G.json; tail -5 gpurun_out/k4_strong_n$G.err
echo done
