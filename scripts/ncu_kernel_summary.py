#!/usr/bin/env python
"""Summarise one `ncu --set full` capture of any kernel into profiles/<tag>_<name>_ncu.json + _ncu_top.txt
(key metrics, warp-stall reasons per issue, the 25 SASS lines with most warp-stall samples).

    python scripts/ncu_kernel_summary.py gpurun_out/<file>.ncu-rep <name> <tag>"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from ncu_summary import KEYS, ROOT, UNIT, page  # noqa: E402

EXTRA = {"sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active": "tensor_pipe_pct",
         "l1tex__t_sector_pipe_lsu_mem_global_op_ld_hit_rate.pct": "l1_hit_pct",
         "smsp__cycles_active.avg": "smsp_cycles_active"}


def main():
    rep, name, tag = sys.argv[1:4]
    rows = page(rep, "raw")
    hdr, units, row = rows[0], rows[1], rows[2]
    rec = {"kernel": row[hdr.index("Kernel Name")], "source": os.path.basename(rep)}
    keys = dict(KEYS, **EXTRA)
    for h, u, v in zip(hdr, units, row):
        if h in keys:
            x = float(v.replace(",", ""))
            rec[keys[h]] = x * UNIT.get(u, 1.0) if keys[h] in ("time_ms", "dram_read", "dram_write") else x
        if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio") and float(v) >= 0.05:
            rec.setdefault("stall_per_issue", {})[h[len("smsp__average_warps_issue_stalled_"):-len("_per_issue_active.ratio")]] = round(float(v), 3)
    json.dump(rec, open(os.path.join(ROOT, "profiles", f"{tag}_{name}_ncu.json"), "w"), indent=1)
    src = page(rep, "source")
    ci = {h: i for i, h in enumerate(src[1])}
    data = src[2:]
    tot = sum(int(r[ci["# Samples"]]) for r in data) or 1
    toti = sum(int(r[ci["Instructions Executed"]]) for r in data) or 1
    top = sorted(((int(r[ci["# Samples"]]), i, r[1].strip()[:80], int(r[ci["Instructions Executed"]])) for i, r in enumerate(data)), reverse=True)[:25]
    lines = [rec["kernel"], "SASS line  warp-stall samples %  executed warp instructions %  instruction"]
    lines += [f"{i:5d} {100 * s / tot:6.2f}% {100 * e / toti:6.2f}%  {t}" for s, i, t, e in top]
    open(os.path.join(ROOT, "profiles", f"{tag}_{name}_ncu_top.txt"), "w").write("\n".join(lines) + "\n")
    print(json.dumps(rec))


if __name__ == "__main__":
    main()
