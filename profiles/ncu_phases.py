"""Stall samples / executed instructions of evolve_small_kernel grouped by phase, from
   ncu -i X.ncu-rep --page source --csv --print-source sass,cuda > x.csv ; python profiles/ncu_phases.py x.csv"""
import collections, csv, sys

rows = list(csv.reader(open(sys.argv[1])))
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


phase, inst, tot = collections.Counter(), collections.Counter(), 0
for s in secs:
    h = s["hdr"]
    isamp, iinst = h.index("# Samples"), h.index("Instructions Executed")
    for r in s["rows"]:
        if not r[0].isdigit():
            continue
        ln, n, ins, src = int(r[0]), num(r[isamp]), num(r[iinst]), r[1]
        f = s["file"]
        k = f
        if f == "bqa_small.cuh":
            if 40 <= ln <= 86:
                k = "rhs_exact (flip loop)"
            elif 88 <= ln <= 118:
                k = "stage_input_exact (combine)"
            elif 120 <= ln <= 172:
                k = "final853_exact"
            elif 315 <= ln <= 360:
                k = "sum3 / reductions"
            else:
                k = "bqa_small other"
        phase[k] += n
        inst[k] += ins
        tot += n
ti = sum(inst.values())
for k, v in phase.most_common():
    print(f"{k:34s} samples {v:9d} {100 * v / tot:5.1f}%   instructions {inst[k]:.3e} {100 * inst[k] / ti:5.1f}%")
