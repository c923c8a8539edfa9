"""Segment the SASS of one kernel in an ncu source-page CSV by execution count (hot regions)."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
which = int(sys.argv[2]) if len(sys.argv) > 2 else 0
starts = [i for i, r in enumerate(rows) if r and r[0] == "Address"]
s = starts[which]; e = starts[which + 1] - 1 if which + 1 < len(starts) else len(rows)
hdr = rows[s]; data = [r for r in rows[s + 1:e] if len(r) == len(hdr)]
isrc = hdr.index("Source"); ie = hdr.index("Instructions Executed"); it = hdr.index("Thread Instructions Executed"); iss = hdr.index("# Samples")
out = [(int(r[ie]), int(r[it]), int(r[iss]), r[isrc][:80]) for r in data]
tot = sum(o[0] for o in out); tots = sum(o[2] for o in out)
print("kernel", rows[s - 1][1][:60], "total warp inst", tot, "samples", tots)
def flush(a, b):
    e_ = sum(o[0] for o in out[a:b]); t = sum(o[1] for o in out[a:b]); sm = sum(o[2] for o in out[a:b])
    if e_ > 0.003 * tot:
        print(f"[{a:4d}-{b:4d}) n={b-a:3d} inst%={100*e_/tot:5.1f} samp%={100*sm/tots:5.1f} lanes={t/max(e_,1):5.1f}  first: {out[a][3]}")
a = 0
for i in range(1, len(out) + 1):
    if i == len(out) or abs(out[i][0] - out[a][0]) > 0.12 * max(out[a][0], 1) + 1000:
        flush(a, i); a = i
if len(sys.argv) > 3:
    lo, hi = map(int, sys.argv[3].split(":"))
    for k in range(lo, hi): print(k, out[k])
