mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_kernels.py -m gpu -q -x -k attention -p no:cacheprovider 2>&1 | tail -2
for v in "" $VARIANTS; do
  echo "variant [$v]"
  VITVS_LIB=$PWD/vitvs_b200/libvitvs_b200$v.so timeout 200 python tools/prof_kernels.py attention 20 2>&1 | tail -1
done
