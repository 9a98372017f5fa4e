# gpurun -- 'bash tools/profile_round.sh': bench contract test, ncu launch list (first 1500 launches of the bench
# command: set-up, warm-up and the timed device-resident steps) and one --set full capture of the hot kernels.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_bench.py -m gpu -q -x -p no:cacheprovider > gpurun_out/tests_bench.log 2>&1; echo "exit $?" >> gpurun_out/tests_bench.log
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
timeout 600 $CMD > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
timeout 600 $CMD > gpurun_out/plain2.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:"gemm_tc|attention|layernorm" -s 120 -c 8 -f -o gpurun_out/prof $CMD > gpurun_out/ncu_full.log 2>&1
tail -3 gpurun_out/tests_bench.log; tail -2 gpurun_out/ncu_full.log
