"""Print the headline + per-kernel-class table of bench.py JSON lines (one file per argument)."""
import json, sys
for f in sys.argv[1:]:
    l = [x for x in open(f) if x.startswith('{')]
    if not l:
        print(f, "NO JSON LINE"); print(open(f).read()[-1500:]); continue
    d = json.loads(l[-1])
    print(f"== {f}: {d['value']:.1f} {d['unit']}  {d['ms_per_step']:.3f} ms/step  e2e {d['e2e']['value']:.1f}  "
          f"clocks {d.get('clocks')}")
    for n, v in d.get('kernels', {}).items():
        print(f"   {n:12s} {v['ms_per_step']:8.4f} ms  x{v['launches_per_step']:<3d} share {v['share']:.3f}  "
              f"{v.get('tflops', '')} {v.get('frac_of_sustained_peak', '')}")
