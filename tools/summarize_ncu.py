#!/usr/bin/env python
"""Turn gpurun_out/<tag>_*.ncu-rep / <tag>_launches.csv into small text summaries under profiles/.
usage: python tools/summarize_ncu.py <tag>   (run where ncu is installed; no GPU needed)"""
import collections
import csv
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "launch__grid_size",
        "launch__block_size", "smsp__inst_executed.sum", "sm__inst_executed.avg.per_cycle_elapsed",
        "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__thread_inst_executed_per_inst_executed.ratio",
        "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum"]


def to_bytes(v, unit):
    mult = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
    return float(v) * mult.get(unit, 1)


def raw(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    return [(dict(zip(hdr, r)), dict(zip(hdr, units))) for r in rows[2:]]


def main():
    tag = sys.argv[1]
    go = os.path.join(ROOT, "gpurun_out")
    prof = os.path.join(ROOT, "profiles")
    os.makedirs(prof, exist_ok=True)
    traffic = {}
    tpath = os.path.join(prof, "traffic.json")
    if os.path.exists(tpath):
        traffic = json.load(open(tpath))
    names = {"fused_kernel": "k12_fused", "stream_kernel": "k12_fused", "choose_sparse_kernel": "k3_choose", "bytesum_kernel": "k12_check", "choose_kernel": "k3_choose", "count_kernel": "k3_count",
             "emit_kernel": "k3_emit", "classify_kernel": "k3_classify", "block_kernel": "k2_block",
             "warp_kernel": "k2_warp"}
    for f in sorted(os.listdir(go)):
        if not (f.startswith(tag + "_") and f.endswith(".ncu-rep")):
            continue
        lines = []
        for d, u in raw(os.path.join(go, f)):
            kn = d.get("Kernel Name", "")
            lines.append(f"== {kn}")
            for h in WANT:
                if h in d:
                    lines.append(f"  {h} [{u[h]}] = {d[h]}")
            st = []
            for h in d:
                if "issue_stalled" in h and h.endswith("per_issue_active.ratio"):
                    try:
                        st.append((float(d[h]), h.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", "")))
                    except ValueError:
                        pass
            lines.append("  warps stalled per issue: " + ", ".join(f"{n}={v:.2f}" for v, n in sorted(st, reverse=True)[:8]))
            try:
                tb = to_bytes(d["dram__bytes_read.sum"], u["dram__bytes_read.sum"]) + to_bytes(d["dram__bytes_write.sum"], u["dram__bytes_write.sum"])
                for key, nice in names.items():
                    if key in kn:
                        traffic[nice] = {"dram_bytes_per_launch": tb, "source": f"profiles/{f[:-8]}.txt", "kernel": kn,
                                         "capture": f[:-8], "workload": os.environ.get("LIME_NCU_WORKLOAD", "cfg2")}
                lines.append(f"  dram traffic per launch = {tb / 1e9:.3f} GB")
            except (KeyError, ValueError):
                pass
        open(os.path.join(prof, f[:-8] + ".txt"), "w").write("\n".join(lines) + "\n")
        print("wrote", f[:-8] + ".txt")
    lc = os.path.join(go, tag + "_launches.csv")
    if os.path.exists(lc):
        rows = list(csv.DictReader(l for l in open(lc) if not l.startswith("==")))
        agg = collections.OrderedDict()
        for x in rows:
            agg.setdefault(x["Kernel Name"], []).append(float(x["Metric Value"].replace(",", "")))
        tot = sum(sum(v) for v in agg.values())
        with open(os.path.join(prof, tag + "_launches_summary.txt"), "w") as f:
            f.write(f"# ncu --metrics gpu__time_duration.sum --clock-control none ; {len(rows)} launches; shares of summed kernel time\n")
            for k, v in agg.items():
                f.write(f"{k[:70]:70s} launches={len(v):3d} avg_ms={sum(v) / len(v) / 1e6:8.4f} share={sum(v) / tot * 100:5.1f}%\n")
        with open(os.path.join(prof, tag + "_launches.csv"), "w") as f:
            f.write(open(lc).read())
        print("wrote launches")
    json.dump(traffic, open(tpath, "w"), indent=1)


if __name__ == "__main__":
    main()
