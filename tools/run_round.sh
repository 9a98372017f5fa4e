# gpurun -- 'bash tools/run_round.sh [trace]': GPU parity tests, bench, optional attention trace
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x -p no:cacheprovider > gpurun_out/tests.log 2>&1
echo "exit $?" >> gpurun_out/tests.log
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench.log 2>&1
echo "exit $?" >> gpurun_out/bench.log
if [ "$1" = "trace" ]; then bash tools/exp_att_trace.sh; fi
tail -3 gpurun_out/tests.log
