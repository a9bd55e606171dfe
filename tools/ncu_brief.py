"""Headline raw metrics of every kernel in one or more .ncu-rep files (ncu --set full), as text for profiles/.
usage: ncu_brief.py <rep.ncu-rep> [...]"""
import csv, os, subprocess, sys
WANT = ["launch__grid_size", "launch__block_size", "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic",
        "launch__shared_mem_per_block_static", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
        "launch__waves_per_multiprocessor", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct", "sm__warps_active.avg.per_cycle_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed", "l1tex__t_sector_hit_rate.pct", "sm__cycles_elapsed.avg"]
for rep in sys.argv[1:]:
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], stdout=subprocess.PIPE, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        print("# source: %s (ncu --set full --clock-control none; cold-cache, serialised)" % os.path.basename(rep))
        print("%-70s %s" % ("Kernel Name", r[hdr.index("Kernel Name")][:110]))
        for w in WANT:
            if w in hdr:
                i = hdr.index(w)
                print("%-70s %s %s" % (w, r[i], units[i]))
        for i, h in enumerate(hdr):
            if "warp_issue_stalled" in h and h.endswith("_per_warp_active.pct"):
                try:
                    if float(r[i]) > 4.0:
                        print("%-70s %s %%" % (h, r[i]))
                except ValueError:
                    pass
        print()
