# gpurun -- 'bash tools/exp_lnfuse.sh': folded LayerNorm A/B (VITVS_LN_FUSED=0 runs the LayerNorm kernel), interleaved
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x -s -p no:cacheprovider > gpurun_out/tests_lnf.log 2>&1
echo "exit $?" >> gpurun_out/tests_lnf.log
grep -E "passed|failed|error" gpurun_out/tests_lnf.log | tail -3
for r in 1 2; do
  for f in 0 1; do
    VITVS_LN_FUSED=$f timeout 600 python bench.py --fast > gpurun_out/bench_lnf${f}_r${r}.log 2>&1
    echo "LN_FUSED=$f round $r"; python tools/show_bench.py gpurun_out/bench_lnf${f}_r${r}.log
  done
done
