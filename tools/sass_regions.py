"""Condenses an `ncu --page source --csv` dump into contiguous SASS regions with similar execution counts:
where a kernel's instructions and stall samples go. Usage: sass_regions.py <one_source.csv> [min_pct] [dump a b]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr, data = rows[1], rows[2:]
iS, iE, iSm = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("# Samples")
tot = sum(int(r[iE]) for r in data)
tots = sum(int(r[iSm]) for r in data)
if len(sys.argv) > 3 and sys.argv[3] == "dump":
    for k in range(int(sys.argv[4]), int(sys.argv[5])):
        print(k, data[k][iE], data[k][iSm], data[k][iS].strip())
    sys.exit(0)
minpct = float(sys.argv[2]) if len(sys.argv) > 2 else 0.5
print(f"{len(data)} SASS lines, {tot:.4g} warp instructions, {tots} samples")
groups, cur = [], None
for k, r in enumerate(data):
    e, s = int(r[iE]), int(r[iSm])
    if cur and (0.8 * cur["e"] <= e <= 1.25 * cur["e"]):
        cur["n"] += 1; cur["sum"] += e; cur["smp"] += s; cur["end"] = k
    else:
        if cur:
            groups.append(cur)
        cur = {"start": k, "end": k, "e": e, "n": 1, "sum": e, "smp": s}
groups.append(cur)
for g in groups:
    if 100 * g["sum"] / tot > minpct or 100 * g["smp"] / tots > minpct:
        print("%4d-%4d n=%3d exec/inst=%9d  inst%%=%5.1f  samples%%=%5.1f   %s" % (
            g["start"], g["end"], g["n"], g["e"], 100 * g["sum"] / tot, 100 * g["smp"] / tots,
            data[g["start"]][iS].strip()[:50]))
