for v in "" tilebar2 tilebar3 tilebar4; do echo "== ${v:-nobar(wrong numerics)}"; L=$PWD/vitvs_b200/libvitvs_b200${v:+_$v}.so; VITVS_LIB=$L timeout 100 python tools/prof_kernels.py attention 20 2>&1 | tail -1; done
python - <<'PY'
from vitvs_b200 import build
print(build.build_variant("tilebar1", ["VVS_EXP_TILE_BAR=1"]))
print(build.build_variant("tilebar4_stag", ["VVS_EXP_TILE_BAR=4", "VVS_ATT_STAGGER=800"]))
PY
for v in tilebar1 tilebar4_stag; do echo "== $v"; VITVS_LIB=$PWD/vitvs_b200/libvitvs_b200_$v.so timeout 100 python tools/prof_kernels.py attention 20 2>&1 | tail -1; done
VITVS_LIB=$PWD/vitvs_b200/libvitvs_b200_tilebar4.so timeout 100 python tools/att_check.py 8x485x40 1x1370x12 3x50x2 2x257x3 2>&1 | tail -4
