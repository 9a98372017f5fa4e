# gpurun -- 'bash tools/exp_lnfuse.sh': folded LayerNorm A/B after the TMA-load producer epilogue
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_bench_config.py -q -x -p no:cacheprovider 2>&1 | tail -3
for i in 1 2; do for f in 0 1; do
  VITVS_LN_FUSED=$f timeout 300 python bench.py --fast --steps 20 --warmup 5 > gpurun_out/lnf_${f}_$i.log 2>&1
  echo "== fused=$f $i"; python tools/show_bench.py gpurun_out/lnf_${f}_$i.log | head -2
done; done
