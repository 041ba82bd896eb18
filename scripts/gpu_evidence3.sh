cd $GRAFT_REPO_ROOT
timeout 100 python scripts/run_rows_once.py i8 64 > gpurun_out/ev3_plain.log 2>&1 || exit 1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:scan_rows -s 3 -c 1 -f -o gpurun_out/r2_rows_i8_b64_final python scripts/run_rows_once.py i8 64 > gpurun_out/ev3_ncu1.log 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:scan_rows -s 3 -c 1 -f -o gpurun_out/r2_rows_i8_b32_final python scripts/run_rows_once.py i8 32 > gpurun_out/ev3_ncu2.log 2>&1
timeout 100 python scripts/run_refine_once.py 384 10000 500 > gpurun_out/ev3_plain2.log 2>&1 && timeout 300 ncu --set full --clock-control none --import-source on -k regex:refine_l2 -s 2 -c 1 -f -o gpurun_out/r2_refine_d384_r500 python scripts/run_refine_once.py 384 10000 500 > gpurun_out/ev3_ncu3.log 2>&1
B="python bench.py --steps 2 --warmup 3 --no-extras --no-cpu"
timeout 200 $B > gpurun_out/ev3_plain3.log 2>&1 && timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches_bench_fp16_b1_v2.csv $B > gpurun_out/ev3_ncu4.log 2>&1
ls -la gpurun_out/*final*.ncu-rep gpurun_out/r2_refine_d384_r500.ncu-rep
