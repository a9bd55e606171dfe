"""Development aid: print the headline metrics and a SASS opcode histogram from `ncu --page raw/source --csv` dumps."""
import csv, sys
from collections import Counter
raw, src = sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else None
rows = list(csv.reader(open(raw)))
hdr, units, r = rows[0], rows[1], rows[2]
want = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'launch__registers_per_thread', 'launch__grid_size',
        'launch__occupancy_limit_shared_mem', 'launch__occupancy_limit_registers', 'smsp__inst_executed.sum',
        'sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'launch__shared_mem_per_block_dynamic',
        'sm__warps_active.avg.per_cycle_active', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum',
        'l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed', 'sm__cycles_elapsed.avg']
for w in want:
    if w in hdr:
        i = hdr.index(w); print(w, r[i], units[i])
for i, h in enumerate(hdr):
    if 'warp_issue_stalled' in h and h.endswith('per_warp_active.pct') and float(r[i]) > 3:
        print(h.replace('smsp__average_', ''), r[i])
if src:
    rows = list(csv.reader(open(src)))
    hdr, data = rows[1], rows[2:]
    ie, te, ss = hdr.index('Instructions Executed'), hdr.index('Thread Instructions Executed'), hdr.index('# Samples')
    tot = sum(int(x[ie]) for x in data); print('total warp instr', tot)
    c, cs = Counter(), Counter()
    for x in data:
        t = x[1].split()
        op = (t[1] if t[0].startswith('@') else t[0]).split('.')[0]
        c[op] += int(x[ie]); cs[op] += int(x[ss])
    for op, n in c.most_common(18): print("%-10s %10d (%.3f) samples=%d" % (op, n, n / tot, cs[op]))
