"""Turn the ncu artefacts a gpurun call brought back (gpurun_out/) into the small, tracked summaries
under profiles/:  <tag>_launches.csv (kernel, duration per launch), <tag>_kernels.json (key metrics per
fused kernel from the --set full capture) and traffic.json (dram bytes per launch, read by bench.py)."""
import csv
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1] if len(sys.argv) > 1 else "r01"
blocks = int(sys.argv[2]) if len(sys.argv) > 2 else 8192
src = os.path.join(ROOT, "gpurun_out")
dst = os.path.join(ROOT, "profiles")
os.makedirs(dst, exist_ok=True)

import re


def short_name(n):
    """ncu's demangled name -> the form r2iq_cuda_kernel_name() reports ('r2iq_pair_kernel<0, 8, 0>')."""
    n = re.sub(r"^void ", "", n)
    n = re.sub(r"\((?:[\w:]+(?:, )?)*\)$", "", n)
    return n.replace("r2iq::", "").replace("(int)", "").replace("(bool)", "")


# ---- launch list -------------------------------------------------------------------------------
rows = [r for r in csv.reader(l for l in open(os.path.join(src, f"{tag}_launches.csv")) if not l.startswith("=="))]
hdr = rows[0]
ix = {h: i for i, h in enumerate(hdr)}
out = [("id", "kernel", "grid", "block", "duration_us")]
tot = {}
for r in rows[1:]:
    if len(r) < len(hdr) or r[ix["Metric Name"]] != "gpu__time_duration.sum":
        continue
    val = float(r[ix["Metric Value"]].replace(",", ""))
    unit = r[ix["Metric Unit"]]
    us = val / 1000.0 if unit in ("ns", "nsecond") else val * (1000.0 if unit in ("ms", "msecond") else 1.0)
    name = short_name(r[ix["Kernel Name"]])
    out.append((r[ix["ID"]], name, r[ix["Grid Size"]], r[ix["Block Size"]], f"{us:.2f}"))
    tot[name] = tot.get(name, 0.0) + us
with open(os.path.join(dst, f"{tag}_launches.csv"), "w", newline="") as f:
    csv.writer(f).writerows(out)
total = sum(tot.values())
share = sorted(((v / total, k, v) for k, v in tot.items()), reverse=True)
with open(os.path.join(dst, f"{tag}_launch_shares.txt"), "w") as f:
    f.write(f"# share of summed kernel time over the profiled command ({len(out) - 1} launches, {total / 1000:.2f} ms)\n")
    for s, k, v in share[:25]:
        f.write(f"{100 * s:6.2f}%  {v / 1000:9.3f} ms  {k}\n")
    # the timed step of bench.py is the FFTN 2048 sweep alone (the FFTN 8192 sweep is a separate section of the line)
    head = {k: v for k, v in tot.items() if "wide" not in k}
    ht = sum(head.values())
    if head and ht < total:
        f.write(f"# shares within the headline step (FFTN 2048 kernels only, {ht / 1000:.2f} ms)\n")
        for v, k in sorted(((v, k) for k, v in head.items()), reverse=True):
            f.write(f"{100 * v / ht:6.2f}%  {v / 1000:9.3f} ms  {k}\n")

# ---- per-kernel metrics from the full captures (FFTN 2048 kernels; FFTN 8192 kernels when captured) -----------------
captures = []
for rep_name in (f"{tag}_full.ncu-rep", f"{tag}_wide_full.ncu-rep", f"{tag}_chan_full.ncu-rep"):
    rep = os.path.join(src, rep_name)
    csv_path = rep[:-len(".ncu-rep")] + ".raw.csv"          # tools/profile_round.sh exports the raw page on the box
    if os.path.exists(csv_path):
        raw = open(csv_path).read()
    elif os.path.exists(rep):
        raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    else:
        continue
    rows = list(csv.reader(raw.splitlines()))
    captures.append((rows[0], rows[1], rows[2:]))
want = {
    "duration_us": "gpu__time_duration.sum",
    "dram_read_bytes": "dram__bytes_read.sum",
    "dram_write_bytes": "dram__bytes_write.sum",
    "dram_throughput_pct": "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "fma_pipe_cycles_active_pct": "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
    "fma_inst_pct": "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "alu_inst_pct": "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "lsu_inst_pct": "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "issue_active_pct": "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "l1_data_pipe_wavefronts_pct": "l1tex__data_pipe_lsu_wavefronts.sum.pct_of_peak_sustained_elapsed",
    "smem_wavefronts_pct": "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
    "smem_bank_conflicts": "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "smem_wavefronts": "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
    "l1_sector_hit_pct": "l1tex__t_sector_hit_rate.pct",
    "icache_hit_pct": "sm__icc_request_hit_rate.pct",
    "l2_hit_pct": "lts__t_sector_hit_rate.pct",
    "warps_active_pct": "sm__warps_active.avg.pct_of_peak_sustained_active",
    "registers_per_thread": "launch__registers_per_thread",
    "grid": "launch__grid_size",
    "block": "launch__block_size",
    "dyn_smem_per_block": "launch__shared_mem_per_block_dynamic",
    "inst_executed": "smsp__inst_executed.sum",
    "sm_cycles_elapsed_max": "sm__cycles_elapsed.max",
    "sm_clock_ghz": "sm__cycles_elapsed.avg.per_second",
}
scale = {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1.0, "ms": 1e3, "us": 1.0, "ns": 1e-3, "s": 1e6}
kern, traffic = [], {}
for hdr, units, data in captures:
  ix = {h: i for i, h in enumerate(hdr)}
  for r in data:
    e = {"kernel": short_name(r[ix["Kernel Name"]])}
    for k, m in want.items():
        if m not in ix:
            e[k] = None
            continue
        try:
            v = float(r[ix[m]].replace(",", ""))
        except ValueError:
            e[k] = r[ix[m]]
            continue
        u = units[ix[m]]
        if k.endswith("_bytes") or k == "duration_us" or k == "dyn_smem_per_block":
            v *= scale.get(u, 1.0)
        e[k] = round(v, 3)
    stalls = {}
    for h in hdr:
        if "issue_stalled" in h and h.endswith("_per_issue_active.ratio") and "not_issued" not in h:
            name = h.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", "")
            stalls[name] = round(float(r[ix[h]]), 3)
    e["warp_cycles_per_issue_by_reason"] = dict(sorted(stalls.items(), key=lambda kv: -kv[1])[:9])
    name = e["kernel"]
    d = None
    for cand in range(7):
        if f"<{cand}>" in name or f"<{cand}," in name:
            d = cand
    e["d"] = d
    if d is not None:
        alg = blocks * 65536 * (2.0 + 8.0 / (1 << (d + 1)))
        if "channels" in name:            # tools/chan_bench.py: 64 channels share one input stream
            alg = blocks * 65536 * (2.0 + 64 * 8.0 / (1 << (d + 1)))
        e["algorithmic_bytes"] = alg
        e["dram_traffic_bytes"] = e["dram_read_bytes"] + e["dram_write_bytes"]
        e["traffic_over_algorithmic"] = round(e["dram_traffic_bytes"] / alg, 4)
        e["gsamples_per_s_under_ncu"] = round(blocks * 65536 / e["duration_us"] / 1e3, 1)
        traffic[name] = {"blocks": blocks, "dram_bytes": e["dram_traffic_bytes"]}
    kern.append(e)
json.dump({"blocks": blocks, "note": "one launch per decimation, 8192-block capture resident in HBM; "
           "ncu --set full --clock-control none (durations are cold-cache and serialised)", "kernels": kern},
          open(os.path.join(dst, f"{tag}_kernels.json"), "w"), indent=1)
json.dump(traffic, open(os.path.join(dst, "traffic.json"), "w"), indent=1)
for e in kern:
    print(e["d"], e["kernel"][:60], e["duration_us"], e["traffic_over_algorithmic"] if e["d"] is not None else "",
          e["fma_pipe_cycles_active_pct"], e["issue_active_pct"], e["icache_hit_pct"], e["registers_per_thread"])
