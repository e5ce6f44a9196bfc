"""Top SASS lines by stall samples from an `ncu --page source --csv` export."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]
ia, isrc, isamp, iexe = hdr.index("Address"), hdr.index("Source"), hdr.index("# Samples"), hdr.index("Instructions Executed")
body = []
for r in rows[2:]:
    if len(r) <= iexe or not r[isamp].isdigit():
        if body:
            break  # first launch only
        continue
    body.append(r)
tot = sum(int(r[isamp] or 0) for r in body)
tote = sum(int(r[iexe] or 0) for r in body)
print("total samples", tot, "warp instr executed", tote)
top = sorted(range(len(body)), key=lambda i: -int(body[i][isamp] or 0))[: int(sys.argv[2]) if len(sys.argv) > 2 else 25]
for i in sorted(top):
    r = body[i]
    print(f"{i:5d} {100*int(r[isamp] or 0)/max(tot,1):5.1f}%  exe={int(r[iexe] or 0):9d}  {r[isrc].strip()[:110]}")
