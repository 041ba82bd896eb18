# full-size parity + properties (K1 / K2 / K3 shapes); refreshed ncu launch list of the bench command
set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_full_size.py -q -x > gpurun_out/pytest_full.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_full.log
tail -30 gpurun_out/pytest_full.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --kernel-name-base demangled -k 'regex:scan_|merge_|rescore|prep_|exchange' -c 60 --csv --log-file gpurun_out/launches_bench.csv python bench.py --steps 20 --warmup 3 --no-extras --no-cpu > gpurun_out/ncu_bench.log 2>&1; echo "ncu bench rc=$?"
echo done
