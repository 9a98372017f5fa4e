mkdir -p gpurun_out
python - <<'PY'
from vitvs_b200 import build
for st in (500, 1000, 2000):
    print(build.build_variant(f"stag{st}", [f"VVS_ATT_STAGGER={st}"]))
PY
( echo "== base"; timeout 300 python tools/prof_kernels.py attention 20;
for st in 500 1000 2000; do echo "== stagger $st";  VITVS_LIB=$PWD/vitvs_b200/libvitvs_b200_stag$st.so timeout 300 python tools/prof_kernels.py attention 20; done ) > gpurun_out/att8_time.log 2>&1
cat gpurun_out/att8_time.log
