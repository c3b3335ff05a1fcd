"""Aggregate an `ncu --page source --csv` export by SASS address block: samples and executed instructions."""
import csv, sys
path = sys.argv[1]; blk = int(sys.argv[2], 0) if len(sys.argv) > 2 else 0x200
rows = list(csv.reader(open(path)))
hdr = rows[1]; data = rows[2:]
ix = {h: i for i, h in enumerate(hdr)}
def f(r, k):
    try: return float(r[ix[k]])
    except Exception: return 0.0
first = None
agg = {}
for n, r in enumerate(data):
    try: addr = int(r[ix["Address"]], 16)
    except Exception: continue
    if first is None or addr < first_addr:
        if first is None: first_addr = addr
        first = True
    off = addr - first_addr
    if off < 0: break  # next kernel instance
    b = off // blk
    a = agg.setdefault(b, [0.0, 0.0, 0, ""])
    a[0] += f(r, "# Samples"); a[1] += f(r, "Instructions Executed"); a[2] += 1
    if not a[3] and ("MUFU" in r[ix["Source"]] or "BAR" in r[ix["Source"]] or "SYNCS" in r[ix["Source"]] or "ATOM" in r[ix["Source"]]): a[3] = r[ix["Source"]].strip()[:50]
tot = sum(a[0] for a in agg.values()); toti = sum(a[1] for a in agg.values())
print("samples", tot, "warp instructions", toti)
for b, a in sorted(agg.items()):
    if a[0] >= 0.01 * tot:
        print(f"  +0x{b*blk:05x} samples {a[0]:6.0f} ({100*a[0]/tot:4.1f}%)  inst {a[1]:10.0f} ({100*a[1]/max(toti,1):4.1f}%)  {a[3]}")
