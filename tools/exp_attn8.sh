mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_kernels.py -m gpu -q -k attention -p no:cacheprovider > gpurun_out/att8_tests.log 2>&1
echo "exit $?" >> gpurun_out/att8_tests.log
tail -2 gpurun_out/att8_tests.log
python - <<'PY'
from vitvs_b200 import build
for st in (600, 1200):
    print(build.build_variant(f"stag{st}", [f"VVS_ATT_STAGGER={st}"]))
print(build.build_variant("trace", ["VVS_ATT_TRACE", "VVS_ATT_STAGGER=600"]))
PY
( echo "== v8"; timeout 300 python tools/prof_kernels.py attention 20;
for st in 600 1200; do echo "== v8 stagger $st";  VITVS_LIB=$PWD/vitvs_b200/libvitvs_b200_stag$st.so timeout 300 python tools/prof_kernels.py attention 20; done ) > gpurun_out/att8_time.log 2>&1
cat gpurun_out/att8_time.log
VITVS_LIB=$PWD/vitvs_b200/libvitvs_b200_trace.so timeout 120 python tools/att_trace.py > gpurun_out/att8_trace.log 2>&1
