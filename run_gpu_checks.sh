# GPU-box check script (gpurun -- 'bash run_gpu_checks.sh [profile]'): parity tests, smoke, bench, optional ncu.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -s -p no:cacheprovider > gpurun_out/tests.log 2>&1
echo "exit $?" >> gpurun_out/tests.log
timeout 300 python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1
echo "exit $?" >> gpurun_out/smoke.log
timeout 900 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench.log 2>&1
echo "exit $?" >> gpurun_out/bench.log
if [ "$1" = "profile" ]; then
  CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
  $CMD > gpurun_out/plain.log 2>&1 &&
  ncu --metrics gpu__time_duration.sum --clock-control none -s 420 -c 330 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
  $CMD > gpurun_out/plain2.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:"gemm_tc|attention" -s 100 -c 5 -f -o gpurun_out/prof $CMD > gpurun_out/ncu_full.log 2>&1
fi
tail -3 gpurun_out/tests.log
