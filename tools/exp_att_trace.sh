mkdir -p gpurun_out
python - <<'PY'
from vitvs_b200 import build
print(build.build_variant("trace", ["VVS_ATT_TRACE"]))
print(build.build_variant("trace_nomath_nomma", ["VVS_ATT_TRACE", "VVS_EXP_NOMATH", "VVS_EXP_NOMMA"]))
PY
VITVS_LIB=$PWD/vitvs_b200/libvitvs_b200_trace.so python tools/att_trace.py > gpurun_out/att_trace_full.log 2>&1
VITVS_LIB=$PWD/vitvs_b200/libvitvs_b200_trace_nomath_nomma.so python tools/att_trace.py > gpurun_out/att_trace_skel.log 2>&1
tail -5 gpurun_out/att_trace_full.log
