"""Summarise an `ncu --page source --csv` export: stall mix and hottest SASS lines."""
import csv, sys
path = sys.argv[1]; topn = int(sys.argv[2]) if len(sys.argv) > 2 else 40
rows = list(csv.reader(open(path)))
hdr = rows[1]; data = rows[2:]
ix = {h: i for i, h in enumerate(hdr)}
def f(r, k):
    try: return float(r[ix[k]])
    except Exception: return 0.0
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
tot = {s: sum(f(r, s) for r in data) for s in stalls}
allsamp = sum(f(r, "# Samples") for r in data)
print("total samples", allsamp, "instructions", sum(f(r, "Instructions Executed") for r in data))
for s, v in sorted(tot.items(), key=lambda kv: -kv[1])[:10]:
    print(f"  {s:28s} {v:10.0f} {100*v/max(allsamp,1):5.1f}%")
print("hottest lines by samples:")
order = sorted(range(len(data)), key=lambda i: -f(data[i], "# Samples"))[:topn]
for i in sorted(order):
    r = data[i]
    top = max(stalls, key=lambda s: f(r, s))
    print(f"{i:6d} {f(r,'# Samples'):7.0f} {f(r,'Instructions Executed'):9.0f} {top:22s} {r[ix['Source']].strip()[:90]}")
