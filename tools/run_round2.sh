# gpurun -- 'bash tools/run_round2.sh': GPU parity suite, smoke, the headline bench and the CPU reference arm, then profiles
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q -s -p no:cacheprovider --durations=10 > gpurun_out/tests.log 2>&1
echo "exit $?" >> gpurun_out/tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1
echo "exit $?" >> gpurun_out/smoke.log
( time timeout 900 python bench.py ) > gpurun_out/bench.log 2>&1
echo "exit $?" >> gpurun_out/bench.log
grep -E "passed|failed|error" gpurun_out/tests.log | tail -5; tail -2 gpurun_out/smoke.log; python tools/show_bench.py gpurun_out/bench.log
bash tools/profile_round.sh
