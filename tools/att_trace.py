"""Per-block clock stamps of a few attention CTAs (variant built with -DVVS_ATT_TRACE; timing experiment)."""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from vitvs_b200 import Engine, ServoConfig, _lib
eng = Engine(ServoConfig(model_type="dinov2_vitb14"), max_images=2)
n_img, n_tok, heads, D = 128, int(os.environ.get("NTOK", "485")), 12, 768
g = torch.Generator(device="cuda").manual_seed(0)
qkv = torch.randn(n_img * n_tok, 3 * D, device="cuda", generator=g).to(torch.bfloat16)
qkv[:, :D] *= 0.125
for _ in range(3):
    eng.dbg_attention(qkv, n_img, n_tok, heads)
torch.cuda.synchronize()
lib = ctypes.CDLL(os.environ["VITVS_LIB"])
buf = np.zeros(4 * 3 * 16 * 8, dtype=np.int64)
rc = lib.vvs_dbg_att_trace(buf.ctypes.data_as(ctypes.c_void_p))
tr = buf.reshape(4, 3, 16, 8)
nblk = (n_tok + 63) // 64
names = ["top", "s_full", "Sregs", "math", "arrive", "p_full ok", "PV issued", "S issued"]
for slot in range(4):
    t0 = tr[slot, 0, 0, 0]
    life = tr[slot, 2, 15]
    print(f"--- CTA slot {slot}: entry {int(life[5] - t0)}  warp-0 role end {int(life[6] - t0)}  exit {int(life[7] - t0)}  "
          f"(lifetime {int(life[7] - life[5])} clk)")
    print(f"--- CTA slot {slot} (rc {rc}); clk relative to warp 0 loop start")
    for w, wn in enumerate(["warp0 (tile0)", "warp1 (tile0)", "issuer warp, tile 0: top=p_full seen, s_full=PV+S issued, Sregs=next operands landed"]):
        print(" ", wn)
        for j in range(nblk):
            ev = tr[slot, w, j]
            rel = [int(e - t0) if e else -1 for e in ev]
            print(f"    j={j:2d} " + "  ".join(f"{n}:{r}" for n, r in zip(names, rel) if r >= 0 or True))
