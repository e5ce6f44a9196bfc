"""profiles/ncu_traffic.json from an `ncu --set full` capture exported with `ncu -i X.ncu-rep --page raw --csv`:
DRAM bytes (read + write) per kernel launch of the LAST frame in the capture, and per bench.py stage."""
import csv, json, os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from bench import csrc_hash
rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
ni, ri, wi = hdr.index('Kernel Name'), hdr.index('dram__bytes_read.sum'), hdr.index('dram__bytes_write.sum')
mult = {'byte': 1, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}
frames, cur = [], None
for r in rows[2:]:
    n = r[ni]
    if 'k_set_params' in n:
        cur = []
        frames.append(cur)
    if cur is None:
        continue
    key = n.split('(')[0].replace('void ', '').replace('<unnamed>::', '')
    cur.append((key, int(float(r[ri].replace(',', '')) * mult[units[ri]] + float(r[wi].replace(',', '')) * mult[units[wi]])))
last = frames[-1]
stage = {"preprocess": 0, "sort_depth": 0, "ranges": 0, "duplicate": 0, "sort_tiles": 0, "blend": 0}
seen_dup = False
for k, b in last:
    if k.startswith("k_preprocess"): stage["preprocess"] += b
    elif k.startswith("k_tile_scan"): stage["ranges"] += b
    elif k.startswith("k_duplicate"): stage["duplicate"] += b; seen_dup = True
    elif k.startswith("k_blend"): stage["blend"] += b
    elif k.startswith("k_rs_"): stage["sort_tiles" if seen_dup else "sort_depth"] += b
json.dump({"source": sys.argv[2] if len(sys.argv) > 2 else "", "workload": "3M_sh3_1080p", "csrc_hash": csrc_hash(),
           "per_kernel": [[k, b] for k, b in last], "per_stage": stage}, open(sys.argv[3] if len(sys.argv) > 3 else "profiles/ncu_traffic.json", "w"), indent=1)
print(stage)
