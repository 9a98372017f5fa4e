# timing experiments on the attention kernel (wrong numerics in the variants; timing only)
mkdir -p gpurun_out
python - <<'PY'
from vitvs_b200 import build
build.build_library()
for name, d in [("nomufu", ["VVS_EXPERIMENT_NO_MUFU"]), ("nopstore", ["VVS_EXPERIMENT_NO_PSTORE"]),
                ("nomufu_nopstore", ["VVS_EXPERIMENT_NO_MUFU", "VVS_EXPERIMENT_NO_PSTORE"]), ("notrap", ["VVS_NO_WAIT_TRAP"])]:
    print(build.build_variant(name, d))
PY
echo "baseline" > gpurun_out/att_exp.log; python tools/prof_kernels.py attention 20 >> gpurun_out/att_exp.log 2>&1
for v in nomufu nopstore nomufu_nopstore notrap; do echo "variant $v" >> gpurun_out/att_exp.log; VITVS_LIB=$PWD/vitvs_b200/libvitvs_b200_$v.so python tools/prof_kernels.py attention 20 >> gpurun_out/att_exp.log 2>&1; done
cat gpurun_out/att_exp.log
