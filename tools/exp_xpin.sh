# A/B: part of the residual stream kept in L2 by eviction-priority hints (VITVS_XPIN_MB)
mkdir -p gpurun_out
VITVS_XPIN_MB=60 timeout 600 python -m pytest tests/test_gpu_path.py -m gpu -q -x -p no:cacheprovider > gpurun_out/tests_xpin.log 2>&1; echo "exit $?" >> gpurun_out/tests_xpin.log
for r in 1 2 3; do for v in ${XPIN_LIST:-0 32 48 64 80}; do
VITVS_XPIN_MB=$v timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_xpin${v}_$r.log 2>&1
done; done
tail -2 gpurun_out/tests_xpin.log
python tools/show_bench.py gpurun_out/bench_xpin*_?.log | grep -E "^==|layernorm|proj_gemm|ffn2_gemm"
