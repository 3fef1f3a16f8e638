"""Hot SASS instructions of an `ncu --page source --csv` dump: top lines by stall samples (with the dominant stall reason)
and an opcode histogram weighted by executed warp instructions.  Usage: python tools/ncu_sass_hot.py file.csv [ntop]"""
import csv
import sys
from collections import Counter

rows = list(csv.reader(open(sys.argv[1])))
ntop = int(sys.argv[2]) if len(sys.argv) > 2 else 30
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
data = []
for n, r in enumerate(rows[2:]):
    try:
        data.append((int(r[ix["# Samples"]] or 0), int(r[ix["Instructions Executed"]] or 0), r[ix["Source"]].strip(), n,
                     {c: int(r[ix[c]] or 0) for c in stall_cols}))
    except Exception:
        pass
tot = sum(d[0] for d in data) or 1
toti = sum(d[1] for d in data) or 1
print("total samples %d, executed warp instructions %d, SASS lines %d" % (tot, toti, len(data)))
ops = Counter()
for s, e, src, n, st in data:
    ops[src.split()[0] if not src.startswith("@") else src.split()[1]] += e
print("opcode histogram (share of executed warp instructions):")
for op, e in ops.most_common(18):
    print("  %-28s %5.1f%%" % (op, 100.0 * e / toti))
print("hottest instructions by samples:")
for s, e, src, n, st in sorted(data, reverse=True)[:ntop]:
    top = max(st.items(), key=lambda kv: kv[1])
    print("  %5.2f%%  line %5d  exec %9d  %-18s %s" % (100.0 * s / tot, n, e, top[0][6:], src[:110]))
