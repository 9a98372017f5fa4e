# gpurun -- 'bash tools/exp_sim.sh': matching parity tests + config 5 sweep + default bench after a change to the SIM epilogue / normalise kernel
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_boundary.py tests/test_gpu_path.py tests/test_gpu_gram.py -m gpu -q -x -p no:cacheprovider > gpurun_out/sim_tests.log 2>&1; echo "exit $?" >> gpurun_out/sim_tests.log
tail -3 gpurun_out/sim_tests.log
( time timeout 900 python bench.py --config 5 --steps 10 --warmup 3 ) > gpurun_out/r2b_config5_n1.log 2>&1; echo "exit $?" >> gpurun_out/r2b_config5_n1.log
grep '^{' gpurun_out/r2b_config5_n1.log > gpurun_out/r2b_config5_n1.json; tail -5 gpurun_out/r2b_config5_n1.log | cut -c1-300
( time timeout 600 python bench.py --fast ) > gpurun_out/r2b_bench_fast.log 2>&1; python tools/show_bench.py gpurun_out/r2b_bench_fast.log 2>/dev/null | head -30
