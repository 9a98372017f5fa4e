# gpurun -- 'bash tools/run_r2b.sh': round-2 final evidence on one GPU - parity suite, smoke, both bench arms, ncu launch list +
# full capture of the same command, BASELINE configs 3/4/5, latency and the cached-goal extra
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q -s -p no:cacheprovider --durations=10 > gpurun_out/tests.log 2>&1
echo "exit $?" >> gpurun_out/tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1
echo "exit $?" >> gpurun_out/smoke.log
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_ref.log 2>&1
( time timeout 900 python bench.py ) > gpurun_out/bench.log 2>&1
echo "exit $?" >> gpurun_out/bench.log
grep -E "passed|failed|error" gpurun_out/tests.log | tail -3; tail -2 gpurun_out/smoke.log | cut -c1-200; python tools/show_bench.py gpurun_out/bench.log
bash tools/profile_round.sh
bash tools/run_configs_n1.sh
VITVS_LAT_VERBOSE=1 timeout 300 python tools/latency.py > gpurun_out/latency.log 2>&1; grep "host p50" gpurun_out/latency.log
timeout 300 python tools/cached_goal_bench.py dinov2_vitb14 64 > gpurun_out/cached_goal.log 2>&1; tail -1 gpurun_out/cached_goal.log
