# gpurun -- 'bash tools/run_final.sh': what the driver runs at round end - GPU tests, smoke, both bench arms (defaults)
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x -p no:cacheprovider > gpurun_out/tests.log 2>&1
echo "exit $?" >> gpurun_out/tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1
echo "exit $?" >> gpurun_out/smoke.log
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_ref.log 2>&1
echo "exit $?" >> gpurun_out/bench_ref.log
( time timeout 900 python bench.py ) > gpurun_out/bench.log 2>&1
echo "exit $?" >> gpurun_out/bench.log
tail -3 gpurun_out/tests.log; tail -2 gpurun_out/smoke.log; tail -c 300 gpurun_out/bench_ref.log; tail -5 gpurun_out/bench.log | cut -c1-300
