# gpurun -- 'bash tools/exp_lnfuse_latency.sh': folded LayerNorm A/B on latency (single pair, ViT-B and ViT-S) and on the other configs
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_bench_config.py -q -x -p no:cacheprovider 2>&1 | tail -2
for f in 0 1 0 1; do
  echo "== fused=$f latency"; VITVS_LN_FUSED=$f timeout 300 python tools/latency.py 2>&1 | tail -4
done
for f in 0 1; do
  echo "== fused=$f config 3 (ViT-L 256 pairs layer 11)"; VITVS_LN_FUSED=$f timeout 600 python bench.py --config 3 --fast --steps 5 --warmup 3 > gpurun_out/lnf4_c3_$f.log 2>&1; python tools/show_bench.py gpurun_out/lnf4_c3_$f.log | head -2
  echo "== fused=$f config 4 (ViT-g 64 pairs layer 11)"; VITVS_LN_FUSED=$f timeout 600 python bench.py --config 4 --fast --steps 5 --warmup 3 > gpurun_out/lnf4_c4_$f.log 2>&1; python tools/show_bench.py gpurun_out/lnf4_c4_$f.log | head -2
done
