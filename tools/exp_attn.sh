mkdir -p gpurun_out
for v in "" _stg400 _stg800 _stg1200; do
  echo "variant [$v]"
  VITVS_LIB=$PWD/vitvs_b200/libvitvs_b200$v.so timeout 200 python tools/prof_kernels.py attention 20 2>&1 | tail -1
done
