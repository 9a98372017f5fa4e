# gpurun -- 'bash tools/exp_sel.sh': selection-kernel change: parity tests that touch matching, quick sweep, single-pair latency
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_boundary.py tests/test_gpu_path.py tests/test_gpu_gram.py tests/test_gpu_episode.py tests/test_gpu_bench_config.py -m gpu -q -x -p no:cacheprovider > gpurun_out/sel_tests.log 2>&1; echo "exit $?" >> gpurun_out/sel_tests.log
tail -3 gpurun_out/sel_tests.log
( time timeout 900 python bench.py --config 5 --steps 10 --warmup 3 ) > gpurun_out/r2c_config5_n1.log 2>&1; echo "exit $?" >> gpurun_out/r2c_config5_n1.log
grep '^{' gpurun_out/r2c_config5_n1.log > gpurun_out/r2c_config5_n1.json
python tools/latency.py > gpurun_out/sel_latency.log 2>&1; tail -5 gpurun_out/sel_latency.log
