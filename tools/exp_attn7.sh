# attention generation 7 (TS MMA, 2 CTAs/SM) vs generation 6: parity tests + kernel-alone timing
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_kernels.py -m gpu -q -k attention -p no:cacheprovider > gpurun_out/att7_tests.log 2>&1
echo "exit $?" >> gpurun_out/att7_tests.log
( echo "== v7"; timeout 300 python tools/prof_kernels.py attention 20; echo "== v6"; VITVS_ATTN=6 timeout 300 python tools/prof_kernels.py attention 20 ) > gpurun_out/att7_time.log 2>&1
tail -5 gpurun_out/att7_tests.log; cat gpurun_out/att7_time.log
