# multi-GPU validation: bench.py weak scaling line + K4-shaped sharded top-100 (per-GPU shard of the 100 M / 8 case)
set -x
mkdir -p gpurun_out
G=${NVDB_GPUS:-2}
nvidia-smi -L > gpurun_out/smi_multi.txt 2>&1
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $G --steps 300 --warmup 10 > gpurun_out/bench_n$G.json 2> gpurun_out/bench_n$G.err; echo "bench rc=$?"
cat gpurun_out/bench_n$G.json | head -c 1500; tail -3 gpurun_out/bench_n$G.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29519 scripts/bench_k4.py > gpurun_out/k4_n$G.json 2> gpurun_out/k4_n$G.err; echo "k4 rc=$?"
cat gpurun_out/k4_n$G.json; tail -5 gpurun_out/k4_n$G.err
echo done
