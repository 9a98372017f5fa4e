# gpurun -- 'bash tools/exp_slabs.sh': interleaved A/B of GEMM epilogue staging variants (experiment builds loaded through VITVS_LIB)
mkdir -p gpurun_out
for v in latewait slabs2; do
  VITVS_LIB=$PWD/vitvs_b200/libvitvs_b200_$v.so timeout 300 python tools/gemm_check.py > gpurun_out/slabs_check_$v.log 2>&1; tail -2 gpurun_out/slabs_check_$v.log
done
for i in 1 2 3; do for v in base latewait slabs2; do
  if [ $v = base ]; then L=$PWD/vitvs_b200/libvitvs_b200.so; else L=$PWD/vitvs_b200/libvitvs_b200_$v.so; fi
  VITVS_LIB=$L timeout 300 python bench.py --fast --steps 20 --warmup 5 > gpurun_out/slabs_${v}_$i.log 2>&1
  python tools/show_bench.py gpurun_out/slabs_${v}_$i.log | head -2
done; done
