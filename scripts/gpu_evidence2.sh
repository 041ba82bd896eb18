cd $GRAFT_REPO_ROOT
timeout 100 python scripts/run_rows_once.py i8 64 > gpurun_out/ev2_plain.log 2>&1 || exit 1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:scan_rows -s 3 -c 1 -f -o gpurun_out/r2_rows_i8_b64 python scripts/run_rows_once.py i8 64 > gpurun_out/ev2_ncu1.log 2>&1
NVDB_B200_ROWS=2 timeout 300 ncu --set full --clock-control none --import-source on -k regex:scan_rows -s 3 -c 1 -f -o gpurun_out/r2_rows_f16_b8 python scripts/run_rows_once.py f16 8 > gpurun_out/ev2_ncu2.log 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:scan_gemv -s 3 -c 1 -f -o gpurun_out/r2_gemv_f16_b1_4stage python bench.py --steps 2 --warmup 3 --no-extras --no-cpu > gpurun_out/ev2_ncu3.log 2>&1
ls -la gpurun_out/*.ncu-rep
