"""Print the handful of ncu metrics this project steers by, per kernel of a .ncu-rep (argv[1])."""
import csv
import subprocess
import sys

raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr = rows[0]
ix = {h: i for i, h in enumerate(hdr)}
want = ["gpu__time_duration.sum", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "l1tex__data_pipe_lsu_wavefronts.sum.pct_of_peak_sustained_elapsed",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "launch__registers_per_thread", "sm__icc_request_hit_rate.pct", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "smsp__warps_active.avg.per_cycle_active", "l1tex__t_sector_hit_rate.pct", "smsp__inst_executed.sum"]
for r in rows[2:]:
    print(r[ix["Kernel Name"]][:60])
    for w in want:
        if w in ix:
            print("    %-80s %s" % (w, r[ix[w]]))
    st = [(float(r[i].replace(",", "")), h) for h, i in ix.items()
          if h.startswith("smsp__average_warps_issue_stalled") and h.endswith("per_issue_active.ratio") and r[i] not in ("", "n/a")]
    st.sort(reverse=True)
    for v, h in st[:7]:
        print("        stall %-28s %.2f" % (h.split("stalled_")[1].split("_per")[0], v))
