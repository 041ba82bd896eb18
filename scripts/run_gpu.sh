#!/bin/bash
# scripts/run_gpu.sh <target> [...] — the GPU-side command lists, one target per kind of evidence (run under gpurun from the
# repo root; everything lands in gpurun_out/). Replaces round 1's numbered one-off scripts.
#   test                 python -m pytest tests -m gpu
#   bench [N]            bench.py, both arms, N GPUs (default 1; N > 1 under torchrun)
#   sanitize <tool>      compute-sanitizer --tool memcheck|racecheck|synccheck|initcheck over scripts/sanitize_case.py
#                        (ONE tool per gpurun call: B200_PROFILING.md)
#   launches <cmd...>    ncu launch list (gpu__time_duration.sum) of a command that has just exited 0 without ncu
#   ncu <regex> <skip> <cmd...>   ncu --set full of the kernels matching <regex> (3 launches after <skip>)
#   mp [N]               tests/mp_sharded_peer.py on N GPUs (one process per GPU; peer == NCCL == unsharded)
set -u
mkdir -p gpurun_out
target=${1:-test}; shift || true
case "$target" in
  test)
    timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
    tail -5 gpurun_out/pytest_gpu.log ;;
  bench)
    N=${1:-1}
    if [ "$N" = 1 ]; then L="python"; else L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"; fi
    timeout 900 $L bench.py --impl reference --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_ref_n$N.json 2> gpurun_out/bench_ref_n$N.err; echo "ref rc=$?"
    timeout 1200 $L bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "ours rc=$?"
    tail -3 gpurun_out/bench_n$N.err; head -c 1500 gpurun_out/bench_n$N.json; echo; head -c 600 gpurun_out/bench_ref_n$N.json; echo ;;
  sanitize)
    tool=${1:-memcheck}; shift || true
    python scripts/sanitize_case.py "$@" > gpurun_out/sanitize_plain.log 2>&1 || { echo "plain run failed"; tail -20 gpurun_out/sanitize_plain.log; exit 1; }
    timeout 1500 compute-sanitizer --tool "$tool" --log-file gpurun_out/sanitizer_$tool.log --print-limit 50 \
      python scripts/sanitize_case.py "$@" > gpurun_out/sanitize_$tool.out 2>&1; echo "sanitizer rc=$?"
    tail -5 gpurun_out/sanitize_$tool.out; grep -c "=========" gpurun_out/sanitizer_$tool.log; tail -15 gpurun_out/sanitizer_$tool.log ;;
  launches)
    "$@" > gpurun_out/launches_plain.log 2>&1 && \
    ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv "$@" > gpurun_out/launches_ncu.log 2>&1
    echo "rc=$?"; tail -3 gpurun_out/launches_ncu.log ;;
  ncu)
    regex=$1; skip=$2; shift 2
    "$@" > gpurun_out/ncu_plain.log 2>&1 && \
    ncu --set full --clock-control none --import-source on -k regex:"$regex" -s "$skip" -c 3 -f -o gpurun_out/prof "$@" > gpurun_out/ncu_full.log 2>&1
    echo "rc=$?"; tail -3 gpurun_out/ncu_full.log ;;
  mp)
    N=${1:-2}
    timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 \
      tests/mp_sharded_peer.py > gpurun_out/mp_sharded_peer_n$N.log 2>&1; echo "rc=$?"; tail -5 gpurun_out/mp_sharded_peer_n$N.log ;;
  *) echo "unknown target $target"; exit 2 ;;
esac
