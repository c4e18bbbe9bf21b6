"""Per-SASS-instruction executed counts + stall samples from `ncu --page source --csv` (profiles/ helper)."""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
# first kernel only; header row is the one starting with "Address"
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]
ia, isrc, iex, ismp = hdr.index("Address"), hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("# Samples")
body = []
for r in rows[hi + 1:]:
    if not r or r[0] == "Kernel Name" or r[0] == "Address":
        break
    try:
        body.append((r[ia], r[isrc], int(r[iex] or 0), int(r[ismp] or 0)))
    except ValueError:
        pass
tot = sum(b[2] for b in body)
tots = sum(b[3] for b in body)
print(f"total inst {tot}, samples {tots}")
thr = float(sys.argv[2]) if len(sys.argv) > 2 else 0.004
for a, s, e, sm in body:
    if e > thr * tot or sm > thr * tots:
        print(f"{a[-5:]} {e / tot * 100:5.1f}% inst {sm / max(tots, 1) * 100:5.1f}% smp  {s[:70]}")
