# refine: our kernel vs the unmodified reference cuda_refine.cu (sm_100a build) on the same GPU — parity test + timing
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_refine.py -q -x > gpurun_out/pytest_refine.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_refine.log
tail -15 gpurun_out/pytest_refine.log
timeout 900 python scripts/bench_refine_vs_reference.py > gpurun_out/refine_vs_reference.jsonl 2> gpurun_out/refine_vs_reference.err; echo "rc=$?"
cat gpurun_out/refine_vs_reference.jsonl; tail -5 gpurun_out/refine_vs_reference.err
echo done
