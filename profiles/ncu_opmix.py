"""Opcode mix of one kernel from `ncu --page source --csv -k regex:<kernel>`: share of executed warp instructions and of
warp-stall samples per SASS opcode (profiles/ helper).  usage: python profiles/ncu_opmix.py src.csv [kernel-index]"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
heads = [i for i, r in enumerate(rows) if r and r[0] == "Address"]
which = int(sys.argv[2]) if len(sys.argv) > 2 else 0
hi = heads[which]
end = heads[which + 1] if which + 1 < len(heads) else len(rows)
hdr = rows[hi]
idx = {h: i for i, h in enumerate(hdr)}


def num(s):
    try:
        return int(s.replace(",", ""))
    except ValueError:
        return 0


body = [r for r in rows[hi + 1:end] if r and len(r) == len(hdr) and r[0] != "Address"]
tot_inst = sum(num(r[idx["Instructions Executed"]]) for r in body)
tot_smp = sum(num(r[idx["# Samples"]]) for r in body)
print(f"warp instructions {tot_inst}, stall samples {tot_smp}")
byop, smp = collections.Counter(), collections.Counter()
for r in body:
    src = r[idx["Source"]].strip().split()
    if not src:
        continue
    op = src[1] if src[0].startswith("@") and len(src) > 1 else src[0]
    op = op.split(".")[0]
    byop[op] += num(r[idx["Instructions Executed"]])
    smp[op] += num(r[idx["# Samples"]])
print("| opcode | executed | samples |\n|---|---|---|")
for op, c in byop.most_common(24):
    print(f"| {op} | {100 * c / max(tot_inst, 1):.1f}% | {100 * smp[op] / max(tot_smp, 1):.1f}% |")
