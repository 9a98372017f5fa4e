"""Top stall lines of one kernel from an ncu source-page CSV (ncu -i rep --page source --csv --kernel-name regex:X)."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
body = [r for r in rows[2:] if len(r) == len(hdr)]
tot = sum(int(r[ix["# Samples"]]) for r in body)
agg = {s: sum(int(r[ix[s]]) for r in body) for s in stalls}
print("total samples", tot, "instr", sum(int(r[ix["Instructions Executed"]]) for r in body))
print({k: v for k, v in sorted(agg.items(), key=lambda kv: -kv[1]) if v})
order = sorted(range(len(body)), key=lambda i: -int(body[i][ix["# Samples"]]))[:top]
for i in sorted(order):
    r = body[i]
    st = {s[6:]: int(r[ix[s]]) for s in stalls if int(r[ix[s]])}
    print(f"{i:5d} {int(r[ix['# Samples']]):6d} {int(r[ix['Instructions Executed']]):9d} {r[ix['Source']].strip()[:70]:70s} {st}")
