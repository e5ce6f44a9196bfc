"""Prints the metrics we care about from an `ncu --page raw --csv` export, one block per launch."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
want = ['Kernel Name', 'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'dram__throughput.avg.pct_of_peak_sustained_elapsed', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread', 'launch__grid_size',
        'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active',
        'l1tex__throughput.avg.pct_of_peak_sustained_active', 'lts__throughput.avg.pct_of_peak_sustained_elapsed',
        'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'smsp__thread_inst_executed_per_inst_executed.ratio']
stall = [h for h in hdr if h.startswith('smsp__average_warps_issue_stalled') and h.endswith('_per_issue_active.ratio')]
if not stall:
    stall = [h for h in hdr if 'issue_stalled' in h and h.endswith('.pct')]
only = sys.argv[2] if len(sys.argv) > 2 else None
seen = set()
for r in rows[2:]:
    name = r[hdr.index('Kernel Name')]
    if only and only not in name: continue
    key = name[:50]
    if key in seen: continue
    seen.add(key)
    print('---', name[:100])
    for w in want[1:]:
        if w in hdr:
            i = hdr.index(w)
            print(f"  {w:75s} {r[i]:>16s} {units[i]}")
    st = sorted(((float(r[hdr.index(h)].replace(',', '') or 0), h) for h in stall), reverse=True)[:6]
    for v, h in st:
        print(f"  stall {h.replace('smsp__average_warps_issue_stalled_','').replace('_per_issue_active.ratio',''):40s} {v:.2f}")
