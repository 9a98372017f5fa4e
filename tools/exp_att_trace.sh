mkdir -p gpurun_out
python - <<'PY'
from vitvs_b200 import build
print(build.build_variant("trace", ["VVS_ATT_TRACE"]))
PY
VITVS_LIB=$PWD/vitvs_b200/libvitvs_b200_trace.so python tools/att_trace.py > gpurun_out/att_trace_full.log 2>&1
tail -3 gpurun_out/att_trace_full.log
