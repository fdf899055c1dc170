"""Top SASS instructions by warp-stall samples from an `ncu --page source --csv` dump.
    ncu -i rep.ncu-rep --page source --csv --kernel-name regex:NAME > src.csv ; python tools/ncu_hot.py src.csv [top]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
hdr = rows[1]
col = {h: i for i, h in enumerate(hdr)}
stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
data = []
for n, r in enumerate(rows[2:]):
    if len(r) < len(hdr):
        continue
    try:
        s = int(r[col["# Samples"]])
    except ValueError:
        continue
    st = {h: int(r[col[h]] or 0) for h in stall_cols}
    data.append((s, n, r[col["Source"]].strip(), int(r[col["Instructions Executed"]] or 0), st))
tot = sum(d[0] for d in data)
agg = {}
for d in data:
    for k, v in d[4].items():
        agg[k] = agg.get(k, 0) + v
print("total samples", tot)
print("stall mix:", ", ".join(f"{k[6:]}={100.0 * v / max(tot, 1):.1f}%" for k, v in sorted(agg.items(), key=lambda x: -x[1])[:8]))
for s, n, src, ie, st in sorted(data, reverse=True)[:top]:
    main = sorted(st.items(), key=lambda x: -x[1])[:2]
    print(f"{100.0 * s / max(tot, 1):5.1f}%  #{n:4d} exec={ie:9d}  {src[:70]:70s} " + " ".join(f"{k[6:]}:{v}" for k, v in main if v))
