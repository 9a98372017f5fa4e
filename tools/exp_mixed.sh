# gpurun -- 'bash tools/exp_mixed.sh': GPU suite on the merged library, async host entries, mixed (preferred 4 / fallback 2)
# cluster GEMM parity and interleaved A/B of the step, single-pair latency
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q -x -p no:cacheprovider > gpurun_out/tests.log 2>&1; echo "exit $?" >> gpurun_out/tests.log
tail -3 gpurun_out/tests.log
timeout 600 python tools/async_check.py > gpurun_out/async_check.log 2>&1; echo "exit $?" >> gpurun_out/async_check.log
cat gpurun_out/async_check.log
VITVS_GEMM_CTAS=4 timeout 300 python tools/gemm_check.py > gpurun_out/mixed_check.log 2>&1
echo "exit $?" >> gpurun_out/mixed_check.log
cat gpurun_out/mixed_check.log
for i in 1 2; do for m in 2 4; do
  VITVS_GEMM_CTAS=$m timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_mixed${m}_$i.log 2>&1
done; done
python tools/show_bench.py gpurun_out/bench_mixed*.log
: > gpurun_out/latency2.log
VITVS_LAT_VERBOSE=1 timeout 300 python tools/latency.py >> gpurun_out/latency2.log 2>&1
grep -v "^    " gpurun_out/latency2.log
