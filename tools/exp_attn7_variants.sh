# timing experiments on attention generation 7 (variants have wrong numerics; timing only)
mkdir -p gpurun_out
python - <<'PY'
from vitvs_b200 import build
for name, d in [("nomath", ["VVS_EXP_NOMATH"]), ("nomma", ["VVS_EXP_NOMMA"]), ("nomath_nomma", ["VVS_EXP_NOMATH", "VVS_EXP_NOMMA"])]:
    print(build.build_variant(name, d))
PY
( echo "== baseline"; python tools/prof_kernels.py attention 20
for v in nomath nomma nomath_nomma; do echo "== variant $v"; NTOK=${NTOK:-485} VITVS_LIB=$PWD/vitvs_b200/libvitvs_b200_$v.so timeout 120 python tools/prof_kernels.py attention 20; done ) > gpurun_out/att7_variants.log 2>&1
cat gpurun_out/att7_variants.log
( for n in 64 128 256 485 1024; do echo "== NTOK $n"; NTOK=$n python tools/prof_kernels.py attention 20;  NTOK=$n VITVS_LIB=$PWD/vitvs_b200/libvitvs_b200_nomath_nomma.so python tools/prof_kernels.py attention 20; done ) > gpurun_out/att7_ntok.log 2>&1
cat gpurun_out/att7_ntok.log
