# gpurun -- 'bash tools/exp_lnfuse2.sh': where the folded-LayerNorm consumers lose their time (experiment builds, wrong values)
mkdir -p gpurun_out
for i in 1 2; do for v in unfused base nostats nocs noboth; do
  L=$PWD/vitvs_b200/libvitvs_b200_$v.so; F=1
  if [ $v = base ]; then L=$PWD/vitvs_b200/libvitvs_b200.so; fi
  if [ $v = unfused ]; then L=$PWD/vitvs_b200/libvitvs_b200.so; F=0; fi
  VITVS_LN_FUSED=$F VITVS_LIB=$L timeout 300 python bench.py --fast --steps 20 --warmup 5 > gpurun_out/lnf2_${v}_$i.log 2>&1
  echo "== $v $i"; python tools/show_bench.py gpurun_out/lnf2_${v}_$i.log | head -2
done; done
