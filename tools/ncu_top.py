"""Summarise `ncu --page source --csv` output: top stalled SASS instructions of the first captured kernel."""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 40
h = rows[1]
data = []
for r in rows[2:]:
    if len(r) != len(h) or r[0] == "Address":
        if data:
            break
        continue
    data.append(r)
si, src = h.index("# Samples"), h.index("Source")
ie = h.index("Instructions Executed")
stall_cols = [i for i, c in enumerate(h) if c.startswith("stall_") and "Not Issued" not in c]
tot = sum(int(r[si] or 0) for r in data)
print("total samples", tot, "instructions", len(data))
agg = {}
for r in data:
    for i in stall_cols:
        if r[i] not in ("", "0"):
            agg[h[i][6:]] = agg.get(h[i][6:], 0) + int(r[i])
print("stall totals:", dict(sorted(agg.items(), key=lambda kv: -kv[1])[:8]))
idx = {id(r): k for k, r in enumerate(data)}
for r in sorted(data, key=lambda r: -int(r[si] or 0))[:n]:
    st = {h[i][6:]: int(r[i]) for i in stall_cols if r[i] not in ("", "0")}
    st = dict(sorted(st.items(), key=lambda kv: -kv[1])[:3])
    print(str(idx[id(r)]).rjust(5), r[si].rjust(7), r[ie].rjust(9), r[src].strip()[:60].ljust(60), st)
