# A/B of attention variants, kernel alone, interleaved rounds: bash tools/exp_att_ab.sh "NAME=DEFINE ..." 
mkdir -p gpurun_out
python - "$@" <<'PY'
import sys
from vitvs_b200 import build
for a in sys.argv[1:]:
    name, _, d = a.partition("=")
    print(build.build_variant(name, [x for x in d.split(",") if x]))
PY
: > gpurun_out/att_ab.log
for round in 1 2 3; do
for a in "$@"; do v=${a%%=*}
echo "== $v" >> gpurun_out/att_ab.log
VITVS_LIB=$PWD/vitvs_b200/libvitvs_b200_$v.so python tools/prof_kernels.py attention 100 >> gpurun_out/att_ab.log 2>&1
done; done
cat gpurun_out/att_ab.log
