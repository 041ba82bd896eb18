set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
python bench.py > gpurun_out/bench_r1.json 2> gpurun_out/bench_r1.err; echo "bench rc=$?" >> gpurun_out/bench_r1.err
python bench.py --impl reference --steps 100 --warmup 3 > gpurun_out/bench_ref_r1.json 2> gpurun_out/bench_ref_r1.err
python bench.py --steps 20 --warmup 3 --no-extras --no-cpu > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file gpurun_out/launches.csv python bench.py --steps 20 --warmup 3 --no-extras --no-cpu > gpurun_out/ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:scan_gemv_tma -s 5 -c 2 -o gpurun_out/prof_scan_f16 python bench.py --steps 20 --warmup 3 --no-extras --no-cpu > gpurun_out/ncu2.log 2>&1
echo done
