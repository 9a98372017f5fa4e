# gpurun -- 'bash tools/exp_latency.sh': single-pair latency under the launch-mode toggles (main library), then with the
# candidate library (small-tile GEMMs + PDL for small batches), plus its parity at the benchmarked configuration
mkdir -p gpurun_out
: > gpurun_out/latency.log
for pdl in 0 1; do for ctas in 2 1; do
  VITVS_LAT_VERBOSE=1 VITVS_PDL=$pdl VITVS_GEMM_CTAS=$ctas timeout 300 python tools/latency.py >> gpurun_out/latency.log 2>&1
done; done
W=$PWD/vitvs_b200/libvitvs_b200_wt.so
echo "--- candidate library" >> gpurun_out/latency.log
VITVS_LIB=$W VITVS_LAT_VERBOSE=1 timeout 300 python tools/latency.py >> gpurun_out/latency.log 2>&1
VITVS_LIB=$W VITVS_PDL=0 timeout 300 python tools/latency.py >> gpurun_out/latency.log 2>&1
VITVS_LIB=$W VITVS_SMALL_TILES=0 timeout 300 python tools/latency.py >> gpurun_out/latency.log 2>&1
VITVS_LIB=$W timeout 900 python -m pytest tests/test_gpu_bench_config.py tests/test_gpu_path.py tests/test_gpu_kernels.py -m gpu -q -x -p no:cacheprovider > gpurun_out/tests_wt.log 2>&1
echo "exit $?" >> gpurun_out/tests_wt.log
timeout 900 python -m pytest tests/test_gpu_bench.py tests/test_gpu_depth.py tests/test_gpu_models.py -m gpu -q -s -p no:cacheprovider > gpurun_out/tests_fix.log 2>&1
echo "exit $?" >> gpurun_out/tests_fix.log
grep -v "^    " gpurun_out/latency.log; tail -3 gpurun_out/tests_wt.log; tail -3 gpurun_out/tests_fix.log
