"""Join an `ncu --page source --csv` (SASS view) export with `nvdisasm -g -c` line info: warp instructions and
stall samples per source line of one kernel.  usage: ncu_lines.py <source.csv> <disasm.sass> <mangled kernel name>"""
import csv
import re
import sys

src_csv, sass, kern = sys.argv[1:4]
# offset -> (file, line) from nvdisasm
line_of, cur, inside = {}, None, False
for ln in open(sass):
    if ln.startswith(".text."):
        inside = ln.strip().rstrip(":") == ".text." + kern
        continue
    if not inside:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/", ln)
    if m:
        line_of[int(m.group(1), 16)] = cur
rows = list(csv.reader(open(src_csv)))
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
base, agg, tot, tots = None, {}, 0.0, 0.0
for r in rows[2:]:
    try:
        addr = int(r[ix["Address"]], 16)
    except Exception:
        continue
    if base is None:
        base = addr
    if addr < base:
        break
    n, s = float(r[ix["Instructions Executed"]] or 0), float(r[ix["# Samples"]] or 0)
    key = line_of.get(addr - base, ("?", 0))
    a = agg.setdefault(key, [0.0, 0.0])
    a[0] += n
    a[1] += s
    tot += n
    tots += s
print(f"warp instructions {tot:.0f}  samples {tots:.0f}")
thr = float(sys.argv[4]) if len(sys.argv) > 4 else 0.01
for key, (n, s) in sorted(agg.items(), key=lambda kv: (kv[0][0], kv[0][1])):
    if n >= thr * tot or s >= thr * tots:
        print(f"{key[0]:>22}:{key[1]:<5} inst {n:10.0f} {100 * n / tot:5.1f}%   samples {s:6.0f} {100 * s / max(tots, 1):5.1f}%")
