mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_resize.py tests/test_gpu_path.py -m gpu -q -x -p no:cacheprovider > gpurun_out/tests_resize.log 2>&1
echo "exit $?" >> gpurun_out/tests_resize.log
for ch in 1 2 4; do
VITVS_HOST_CHUNKS=$ch timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_ch$ch.log 2>&1
echo "exit $?" >> gpurun_out/bench_ch$ch.log
done
tail -5 gpurun_out/tests_resize.log
