"""Hottest CUDA source lines of one kernel (executed warp instructions and stall samples per line) from
   ncu -i X.ncu-rep --page source --csv --print-source sass,cuda > x.csv ; python profiles/ncu_lines.py x.csv [n_top]"""
import csv, sys

csv.field_size_limit(1 << 30)
rows = list(csv.reader(open(sys.argv[1])))
ntop = int(sys.argv[2]) if len(sys.argv) > 2 else 60
secs, cur = [], None
for r in rows:
    if r and r[0] == "File Path":
        cur = {"file": r[1].split("/")[-1], "rows": []}
        secs.append(cur)
    elif r and r[0] == "Line No":
        cur["hdr"] = r
    elif r and r[0] == "Function Name":
        pass
    elif cur is not None and "hdr" in cur:
        cur["rows"].append(r)


def num(x):
    try:
        return int(x)
    except ValueError:
        return 0


lines, tot_i, tot_s = [], 0, 0
for s in secs:
    h = s["hdr"]
    isamp, iinst, isrc = h.index("# Samples"), h.index("Instructions Executed"), h.index("Source")
    for r in s["rows"]:
        if len(r) <= max(isamp, iinst):
            continue
        lines.append((num(r[iinst]), num(r[isamp]), s["file"], r[0], r[isrc].strip()[:110]))
        tot_i += num(r[iinst])
        tot_s += num(r[isamp])
print("files:", [s["file"] for s in secs], "executed warp instructions", tot_i, "samples", tot_s)
for inst, samp, f, ln, src in sorted(lines, reverse=True)[:ntop]:
    print(f"{100.0 * inst / max(tot_i, 1):5.1f}% inst {100.0 * samp / max(tot_s, 1):5.1f}% smp  {f}:{ln}  {src}")
