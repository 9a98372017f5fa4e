# A/B of attention variants (kernel alone, interleaved) + parity sweep of each
bash tools/exp_att_ab.sh "base=" "p01=VVS_ATT_POLY_MASK=0x01" "p11=VVS_ATT_POLY_MASK=0x11" "p49=VVS_ATT_POLY_MASK=0x49" > /dev/null 2>&1
for v in p11; do
echo "== parity $v"; VITVS_LIB=$PWD/vitvs_b200/libvitvs_b200_$v.so python tools/att_check.py 1x64x1 1x485x1 2x485x12 3x50x2 1x1370x12 2x257x3 2>&1 | tail -6
done > gpurun_out/att_check_variants.log 2>&1
cat gpurun_out/att_check_variants.log; grep -E "^==|attention" gpurun_out/att_ab.log
