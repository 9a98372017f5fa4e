"""Print the essentials of bench.py JSON lines: python tools/show_bench.py file [file ...]"""
import json
import sys

for path in sys.argv[1:]:
    for ln in open(path):
        if not ln.startswith("{"):
            continue
        d = json.loads(ln)
        k = d.get("kernels", {})
        cls = "  ".join(f"{n} {v['ms_per_step']:.3f}" for n, v in list(k.items())[:8])
        acct = d.get("step_accounting") or {}
        print(f"{path}: {d['value']:.1f} {d['unit']}  {d['ms_per_step']:.3f} ms/step (p50 {d.get('step_ms_p50')}, p95 "
              f"{d.get('step_ms_p95')})  e2e {d['e2e']['value'] if d.get('e2e') else None}  "
              f"sm {d.get('clocks', {}).get('sm_mhz')} MHz\n    kernels: {cls}\n    accounting: kernels "
              f"{acct.get('kernels_ms')} gaps {acct.get('gaps_ms')} memsets {acct.get('copies_memsets_ms')} outside "
              f"{acct.get('outside_graphs_ms')} sum/step {acct.get('sum_over_step')}  lat {d.get('latency', {}) and d['latency'].get('p50_ms')}")
