"""Summarise `ncu --page source --csv` output: top sampled instructions with their dominant stall reasons.
usage: ncu -i rep --page source --csv --kernel-name regex:NAME --launch-count 1 > src.csv; python tools/ncu_top_stalls.py src.csv [N]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 25
h = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[h]; idx = {k: i for i, k in enumerate(hdr)}
data = [r for r in rows[h + 1:] if len(r) == len(hdr) and r[idx["# Samples"]].isdigit()]
print(rows[0][1][:120] if len(rows[0]) > 1 else "")
S = idx["# Samples"]
tot = sum(int(r[S]) for r in data)
print("total samples", tot, "instructions", len(data))
for r in sorted(data, key=lambda r: -int(r[S]))[:n]:
    st = {k[6:]: int(r[idx[k]]) for k in hdr if k.startswith("stall_") and "Not" not in k and r[idx[k]].isdigit() and int(r[idx[k]]) > 0}
    st = dict(sorted(st.items(), key=lambda kv: -kv[1])[:3])
    print(f"{100*int(r[S])/max(tot,1):5.1f}% {r[idx['Instructions Executed']]:>8s}  {r[idx['Source']].strip()[:64]:64s} {st}")
