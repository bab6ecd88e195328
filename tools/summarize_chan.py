"""Summary of the channelizer captures of tools/chan_profile.sh:  python tools/summarize_chan.py <tag> [blocks=2048] [channels=64]
gpurun_out/<tag>_chan{2048,8192}_d<D>.raw.csv  ->  profiles/<tag>_chan_kernels.json (same keys as <tag>_kernels.json)."""
import csv
import glob
import json
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1] if len(sys.argv) > 1 else "r02e"
blocks = int(sys.argv[2]) if len(sys.argv) > 2 else 2048
n_ch = int(sys.argv[3]) if len(sys.argv) > 3 else 64
want = {
    "duration_us": "gpu__time_duration.sum",
    "dram_read_bytes": "dram__bytes_read.sum",
    "dram_write_bytes": "dram__bytes_write.sum",
    "dram_throughput_pct": "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "fma_pipe_cycles_active_pct": "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
    "lsu_inst_pct": "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "issue_active_pct": "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "l1_data_pipe_wavefronts_pct": "l1tex__data_pipe_lsu_wavefronts.sum.pct_of_peak_sustained_elapsed",
    "smem_wavefronts_pct": "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
    "smem_wavefronts": "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
    "smem_bank_conflicts": "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "l2_hit_pct": "lts__t_sector_hit_rate.pct",
    "icache_hit_pct": "sm__icc_request_hit_rate.pct",
    "registers_per_thread": "launch__registers_per_thread",
    "grid": "launch__grid_size",
    "block": "launch__block_size",
    "inst_executed": "smsp__inst_executed.sum",
}
scale = {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1.0, "ms": 1e3, "us": 1.0, "ns": 1e-3, "s": 1e6}
kern = []
for path in sorted(glob.glob(os.path.join(ROOT, "gpurun_out", f"{tag}_chan*_d*.raw.csv"))):
    m = re.search(r"chan(\d+)_d(\d)", path)
    fftn, d = int(m.group(1)), int(m.group(2))
    rows = list(csv.reader(l for l in open(path) if not l.startswith("==")))
    hdr, units, data = rows[0], rows[1], rows[2:]
    ix = {h: i for i, h in enumerate(hdr)}
    for r in data:
        name = re.sub(r"\(.*\)$", "", re.sub(r"^void ", "", r[ix["Kernel Name"]])).replace("r2iq::", "").replace("(int)", "")
        e = {"kernel": name, "fftn": fftn, "d": d, "channels": n_ch, "blocks": blocks}
        for k, mname in want.items():
            if mname not in ix:
                e[k] = None
                continue
            v = float(r[ix[mname]].replace(",", ""))
            if k.endswith("_bytes") or k == "duration_us":
                v *= scale.get(units[ix[mname]], 1.0)
            e[k] = round(v, 3)
        stalls = {}
        for h in hdr:
            if "issue_stalled" in h and h.endswith("_per_issue_active.ratio") and "not_issued" not in h:
                stalls[h.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", "")] = round(float(r[ix[h]]), 3)
        e["warp_cycles_per_issue_by_reason"] = dict(sorted(stalls.items(), key=lambda kv: -kv[1])[:7])
        # one shared input stream (2 B/sample) + n_ch complex64 output streams decimated by 2^(d+1)
        alg = blocks * 65536 * (2.0 + n_ch * 8.0 / (1 << (d + 1)))
        e["algorithmic_bytes"] = alg
        e["dram_traffic_bytes"] = e["dram_read_bytes"] + e["dram_write_bytes"]
        e["traffic_over_algorithmic"] = round(e["dram_traffic_bytes"] / alg, 4)
        e["achieved_gbs_under_ncu"] = round(alg / e["duration_us"] / 1e3, 1)
        kern.append(e)
        print(fftn, d, name, e["duration_us"], e["traffic_over_algorithmic"], e["fma_pipe_cycles_active_pct"],
              e["l1_data_pipe_wavefronts_pct"], e["issue_active_pct"], e["registers_per_thread"])
out = os.path.join(ROOT, "profiles", f"{tag}_chan_kernels.json")
json.dump({"note": f"{n_ch} channels x {blocks} blocks, one launch per capture (the third of the command), ncu --set full "
           "--clock-control none; durations are cold-cache and serialised, the CUDA-event timings of the same commands are in "
           f"{tag}_chan_plain.txt", "kernels": kern}, open(out, "w"), indent=1)
print(out)
