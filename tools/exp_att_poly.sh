# A/B of the polynomial-exp2 fraction in the attention softmax (kernel alone, interleaved rounds)
mkdir -p gpurun_out
python - <<'PY'
from vitvs_b200 import build
for name, m in (("p00", "0x00"), ("p11", "0x11"), ("p49", "0x49"), ("p55", "0x55")):
    print(build.build_variant(name, [f"VVS_ATT_POLY_MASK={m}"]))
PY
: > gpurun_out/att_poly.log
for round in 1 2 3; do
for v in p00 p11 p49 p55; do
echo "== $v" >> gpurun_out/att_poly.log
VITVS_LIB=$PWD/vitvs_b200/libvitvs_b200_$v.so python tools/prof_kernels.py attention 100 >> gpurun_out/att_poly.log 2>&1
done; done
for v in p11 p55; do
echo "== check $v" >> gpurun_out/att_poly.log
VITVS_LIB=$PWD/vitvs_b200/libvitvs_b200_$v.so timeout 100 python tools/att_check.py 8x485x12 2x257x3 >> gpurun_out/att_poly.log 2>&1
done
cat gpurun_out/att_poly.log
