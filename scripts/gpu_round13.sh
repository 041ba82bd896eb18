set -x
mkdir -p gpurun_out
AB="f16:gemv:1:10;f32:gemv:1:10;f16:gemv:2:10;f16:gemv:8:10;i8:gemv:1:10"
for rep in 1 2; do
for v in packed scalar; do
  L=nano-vectordb_b200/lib/libnvdb_b200.so; [ $v = scalar ] && L=nano-vectordb_b200/lib/libnvdb_b200_scalar.so
  NVDB_B200_LIB=$PWD/$L NVDB_AB="$AB" timeout 600 python scripts/ab_bench.py > gpurun_out/ab_fma_${v}_$rep.jsonl 2> gpurun_out/ab_fma_${v}_$rep.err
  echo "== $v $rep"; cat gpurun_out/ab_fma_${v}_$rep.jsonl | cut -c1-60,100-220
done; done
echo done
