"""Top SASS lines by stall samples from `ncu -i X.ncu-rep --page source --csv` (run on the GPU box; only this summary travels)."""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]
idx = {h: i for i, h in enumerate(hdr)}
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
data = []
for r in rows[2:]:
    try:
        data.append((int(r[idx["# Samples"]]), int(r[idx["Instructions Executed"]]), r[idx["Source"]][:90], r))
    except Exception:
        pass
tot = sum(d[0] for d in data)
ti = sum(d[1] for d in data)
print("kernel", rows[0][1][:100])
print("samples", tot, "warp instructions", ti, "sass lines", len(data))
byop, byopi = collections.Counter(), collections.Counter()
for s, i, src, r in data:
    op = (src.split()[1] if src.startswith("@") else src.split()[0]).split(".")[0]
    byop[op] += s
    byopi[op] += i
for op, c in byop.most_common(18):
    print(f"  {op:10s} samples {100 * c / tot:5.1f}%  inst {100 * byopi[op] / ti:5.1f}%")
agg = collections.Counter()
for s, i, src, r in data:
    for h in stalls:
        try:
            agg[h] += int(r[idx[h]])
        except Exception:
            pass
print("stall mix:", {h: round(100 * c / tot, 1) for h, c in agg.most_common(9)})
n_top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
for s, i, src, r in sorted(data, key=lambda d: -d[0])[:n_top]:
    top = sorted(((int(r[idx[h]] or 0), h[6:]) for h in stalls), reverse=True)[:2]
    print(f"{100 * s / tot:5.2f}% inst {i:9d} {src[:70]:70s} {top}")
