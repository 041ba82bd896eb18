set -x
cd $GRAFT_REPO_ROOT
B="python bench.py --steps 2 --warmup 3 --no-extras --no-cpu"
timeout 200 $B > gpurun_out/ev_plain.log 2>&1 && timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches_bench_fp16_b1.csv $B > gpurun_out/ev_ncu0.log 2>&1
timeout 100 python scripts/run_rows_once.py i8 8 6 > gpurun_out/ev_plain2.log 2>&1 && timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2_launches_exact_i8_b8.csv python scripts/run_rows_once.py i8 8 6 > gpurun_out/ev_ncu1.log 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:scan_rows -s 3 -c 1 -f -o gpurun_out/r2_rows_i8_b8 python scripts/run_rows_once.py i8 8 > gpurun_out/ev_ncu2.log 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:scan_rows -s 3 -c 1 -f -o gpurun_out/r2_rows_i8_b32 python scripts/run_rows_once.py i8 32 > gpurun_out/ev_ncu3.log 2>&1
timeout 100 python scripts/run_refine_once.py 768 10000 256 > gpurun_out/ev_plain3.log 2>&1 && timeout 300 ncu --set full --clock-control none --import-source on -k regex:refine_l2 -s 2 -c 1 -f -o gpurun_out/r2_refine_k5 python scripts/run_refine_once.py 768 10000 256 > gpurun_out/ev_ncu4.log 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:rescore_exact -s 3 -c 1 -f -o gpurun_out/r2_rescore_i8_b8 python scripts/run_rows_once.py i8 8 > gpurun_out/ev_ncu5.log 2>&1
ls -la gpurun_out/*.ncu-rep
