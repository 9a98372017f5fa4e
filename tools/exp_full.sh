mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x -p no:cacheprovider > gpurun_out/tests2.log 2>&1; echo "exit $?" >> gpurun_out/tests2.log
tail -4 gpurun_out/tests2.log
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_new.log 2>&1; echo "exit $?" >> gpurun_out/bench_new.log
tail -c 200 gpurun_out/bench_new.log
