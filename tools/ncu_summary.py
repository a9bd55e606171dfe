"""Summarise an ncu capture for profiles/: headline raw metrics of one kernel + profiles/traffic.json (bench.py reads
the measured DRAM bytes per launch from it). usage: ncu_summary.py <rep.ncu-rep> <kernel key> <samples_per_launch> <out.txt> [algorithmic bytes per sample = 16.125]"""
import csv, json, os, subprocess, sys
rep, key, spl, out = sys.argv[1], sys.argv[2], int(sys.argv[3]), sys.argv[4]
ALG = float(sys.argv[5]) if len(sys.argv) > 5 else 16.125
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], stdout=subprocess.PIPE, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units, r = rows[0], rows[1], rows[2]
want = ["Kernel Name", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
        "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.per_cycle_active",
        "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed", "sm__cycles_elapsed.avg"]
lines = ["# ncu --set full --clock-control none (cold-cache, serialised; compare shares, not absolutes)", "# source: " + os.path.basename(rep)]
vals = {}
for w in want:
    if w in hdr:
        i = hdr.index(w); vals[w] = r[i]
        lines.append("%-70s %s %s" % (w, r[i], units[i]))
for i, h in enumerate(hdr):
    if "warp_issue_stalled" in h and h.endswith("_per_warp_active.pct"):
        try:
            if float(r[i]) > 1.0: lines.append("%-70s %s %%" % (h, r[i]))
        except ValueError:
            pass
def to_bytes(v, u):
    m = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    return float(v.replace(",", "")) * m[u]
ir, iw = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
traffic = to_bytes(r[ir], units[ir]) + to_bytes(r[iw], units[iw])
lines.append("dram bytes per launch (read+write): %.0f ; algorithmic: %.0f ; ratio %.3f" % (traffic, ALG * spl, traffic / (ALG * spl)))
open(out, "w").write("\n".join(lines) + "\n")
tj = os.path.join(ROOT, "profiles", "traffic.json")
t = json.load(open(tj)) if os.path.exists(tj) else {}
t[key] = {"samples_per_launch": spl, "dram_bytes": traffic, "source": os.path.basename(out)}
json.dump(t, open(tj, "w"), indent=1)
print("\n".join(lines))
