# gpurun -- 'bash tools/prof_single_pair.sh': ncu --set full of the GEMMs of a single ViT-B frame pair (eager launches)
mkdir -p gpurun_out
cat > /tmp/sp.py <<'PY'
import sys, torch
sys.path.insert(0, '.')
from vitvs_b200 import Engine, ServoConfig, synth
from vitvs_b200.weights import random_state_dict
cfg = ServoConfig(model_type='dinov2_vitb14')
eng = Engine(cfg, max_images=2)
eng.load_weights(random_state_dict('dinov2_vitb14', seed=0, max_block=cfg.layer))
g, c, d = (torch.from_numpy(a[None].copy()).cuda() for a in synth.frame_pair(0))
for _ in range(4):
    eng.frame_pairs_to_twist(g, c, d)
torch.cuda.synchronize()
PY
export VITVS_GRAPHS=0 VITVS_PAIR64=0
timeout 300 python /tmp/sp.py && timeout 900 ncu --set full --clock-control none --import-source on -k regex:"gemm_tc" -s 160 -c 10 -f -o gpurun_out/prof_sp python /tmp/sp.py > gpurun_out/ncu_sp.log 2>&1
tail -2 gpurun_out/ncu_sp.log
ncu -i gpurun_out/prof_sp.ncu-rep --page raw --csv > gpurun_out/prof_sp_raw.csv 2>/dev/null
ls -la gpurun_out/prof_sp*
