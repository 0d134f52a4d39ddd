#!/usr/bin/env python
"""Summarise one `ncu --set full` capture of a scan kernel into profiles/ (tracked) and refresh
profiles/ncu_scan_traffic.json, which bench.py reads for roofline.traffic / dram_frac / lsu_frac.

    python scripts/ncu_summary.py gpurun_out/<tag>_prof_scan_<wl>.ncu-rep <workload> <tag>

Writes profiles/<tag>_scan_<wl>_ncu.json (key metrics), profiles/<tag>_scan_<wl>_ncu_regions.txt (warp-stall samples and
executed instructions per kernel region, split at block barriers, from the source page) and updates the traffic table."""
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KEYS = {
    "gpu__time_duration.sum": "time_ms",
    "dram__bytes_read.sum": "dram_read",
    "dram__bytes_write.sum": "dram_write",
    "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed": "lsu_pipe_pct",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed": "lsu_shared_pct",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum": "smem_bank_conflict_wavefronts",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum": "smem_wavefronts",
    "smsp__issue_active.avg.pct_of_peak_sustained_active": "issue_active_pct",
    "sm__warps_active.avg.pct_of_peak_sustained_active": "occupancy_pct",
    "launch__registers_per_thread": "registers",
    "lts__t_sector_hit_rate.pct": "l2_hit_pct",
    "smsp__inst_executed.sum": "warp_instructions",
    "smsp__average_warp_latency_per_inst_issued.ratio": "cycles_per_issue",
    "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active": "fma_pipe_pct",
    "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active": "alu_pipe_pct",
}
UNIT = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "ms": 1.0, "us": 1e-3, "s": 1e3, "ns": 1e-6}


def page(rep, name):
    out = subprocess.run(["ncu", "-i", rep, "--page", name, "--csv"], capture_output=True, text=True).stdout
    return list(csv.reader(io.StringIO(out)))


def main():
    rep, wl, tag = sys.argv[1:4]
    rows = page(rep, "raw")
    hdr, units, row = rows[0], rows[1], rows[2]
    rec = {"kernel": row[hdr.index("Kernel Name")], "workload": wl, "source": os.path.basename(rep)}
    for h, u, v in zip(hdr, units, row):
        if h in KEYS:
            x = float(v.replace(",", ""))
            rec[KEYS[h]] = x * UNIT.get(u, 1.0) if KEYS[h] in ("time_ms", "dram_read", "dram_write") else x
        if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio"):
            rec.setdefault("stall_per_issue", {})[h[len("smsp__average_warps_issue_stalled_"):-len("_per_issue_active.ratio")]] = round(float(v), 3)
    rec["dram_bytes_per_launch"] = rec["dram_read"] + rec["dram_write"]
    out = os.path.join(ROOT, "profiles", f"{tag}_scan_{wl}_ncu.json")
    json.dump(rec, open(out, "w"), indent=1)
    # regions of the kernel between block barriers: where the warp time goes
    src = page(rep, "source")
    sh = src[1]
    ci = {h: i for i, h in enumerate(sh)}
    data = src[2:]
    tot = sum(int(r[ci["# Samples"]]) for r in data) or 1
    toti = sum(int(r[ci["Instructions Executed"]]) for r in data) or 1
    lines, acc, acci, start = [f"{rec['kernel']}\nregion (SASS line range, ends at)  warp-stall samples %  executed warp instructions %"], 0, 0, 0
    for i, r in enumerate(data):
        acc += int(r[ci["# Samples"]])
        acci += int(r[ci["Instructions Executed"]])
        t = r[1].strip()
        if "BAR." in t or t.startswith("EXIT") or " EXIT" in t or t.startswith("RET") or i == len(data) - 1:
            if acc or acci:
                lines.append(f"{start:5d}-{i:5d}  {t[:44]:44s} {100 * acc / tot:6.2f}  {100 * acci / toti:6.2f}")
            acc, acci, start = 0, 0, i + 1
    top = sorted(((int(r[ci["# Samples"]]), i, r[1].strip()[:70]) for i, r in enumerate(data)), reverse=True)[:12]
    lines.append("\ntop instructions by samples")
    lines += [f"{i:5d} {100 * s / tot:6.2f}%  {t}" for s, i, t in top]
    open(os.path.join(ROOT, "profiles", f"{tag}_scan_{wl}_ncu_regions.txt"), "w").write("\n".join(lines) + "\n")
    tpath = os.path.join(ROOT, "profiles", "ncu_scan_traffic.json")
    tab = [t for t in json.load(open(tpath)) if not (t["workload"] == wl and t["n_gpus"] == 1)] if os.path.exists(tpath) else []
    tab.append({"workload": wl, "scan_variant": 3, "n_gpus": 1, "dram_bytes_per_launch": rec["dram_bytes_per_launch"],
                "lsu_pipe_pct": rec.get("lsu_pipe_pct"), "issue_active_pct": rec.get("issue_active_pct"),
                "kernel_ms_under_ncu": rec.get("time_ms"), "source": f"profiles/{tag}_scan_{wl}_ncu.json ({rec['kernel'][:60]})"})
    json.dump(tab, open(tpath, "w"), indent=1)
    print(json.dumps({k: rec[k] for k in ("kernel", "time_ms", "dram_bytes_per_launch", "lsu_pipe_pct", "issue_active_pct", "occupancy_pct", "registers") if k in rec}))


if __name__ == "__main__":
    main()
