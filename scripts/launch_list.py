"""Prints the last frame's kernels from an `ncu --metrics gpu__time_duration.sum --csv` log."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
i0 = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
rows = rows[i0 + 1:]
names = [(r[4].split("(")[0].replace("void <unnamed>::", "").replace("<unnamed>::", ""), float(r[-1]) / 1000, r[8]) for r in rows]
idx = [i for i, (n, _, _) in enumerate(names) if "k_set_params" in n]
tot = 0.0
for n, t, g in names[idx[-1]:]:
    tot += t
    print(f"{t:8.1f} us  grid {g:16s} {n}")
print(f"{tot:8.1f} us  total")
