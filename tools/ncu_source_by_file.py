"""Aggregate an ncu source-page CSV (ncu -i rep --page source --csv --print-source cuda,sass) by source file,
and optionally list the hottest source lines of one file.  usage: ncu_source_by_file.py src.csv [file-substr [top]]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
secs, cur = [], None
for r in rows:
    if r and r[0] == "File Path": cur = {"file": r[1], "rows": []}; secs.append(cur)
    elif r and r[0] == "Line No": cur["hdr"] = r
    elif cur is not None and "hdr" in cur and len(r) == len(cur["hdr"]): cur["rows"].append(r)
def num(x):
    try: return int(x)
    except ValueError: return 0
out, tot_i, tot_s = [], 0, 0
for s in secs:
    h = s["hdr"]; ie, it, iss, ino = (h.index(k) for k in ("Instructions Executed", "Thread Instructions Executed", "# Samples", "stall_no_inst"))
    I = sum(num(r[ie]) for r in s["rows"]); T = sum(num(r[it]) for r in s["rows"]); S = sum(num(r[iss]) for r in s["rows"]); N = sum(num(r[ino]) for r in s["rows"])
    out.append((s["file"], I, T, S, N)); tot_i += I; tot_s += S
for f, I, T, S, N in sorted(out, key=lambda x: -x[3]):
    print(f"{f[-40:]:40s} inst% {100*I/tot_i:5.1f} samp% {100*S/tot_s:5.1f} lanes {T/max(I,1):5.1f} no_inst share of samples {100*N/max(S,1):5.1f}%")
if len(sys.argv) > 2:
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 25
    for s in secs:
        if sys.argv[2] in s["file"]:
            h = s["hdr"]; ie, it, iss = (h.index(k) for k in ("Instructions Executed", "Thread Instructions Executed", "# Samples"))
            for r in sorted(s["rows"], key=lambda r: -num(r[iss]))[:top]:
                print(f"  L{r[0]:>4} samp% {100*num(r[iss])/tot_s:5.2f} inst% {100*num(r[ie])/tot_i:5.2f} lanes {num(r[it])/max(num(r[ie]),1):5.1f}  {r[1][:110]}")
