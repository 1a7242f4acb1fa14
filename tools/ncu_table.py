#!/usr/bin/env python
"""Per-kernel table from an `ncu --set full` report: duration, DRAM traffic, throughput percentages, registers,
occupancy, issue activity. Also updates profiles/ncu_traffic.json (read by bench.py for roofline.traffic) and profiles/ncu_issue.json (warp
instructions, active lanes per instruction and issue-slot use of each stage's largest kernel: roofline.issue).

    python tools/ncu_table.py gpurun_out/prof_C2_all.ncu-rep C2 profiles/r01_ncu_C2_kernels.md
"""
import csv
import json
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
STAGE_OF = {"preprocess_kernel": "preprocess_fwd", "surfel_preprocess_kernel": "preprocess_fwd",
            "render_forward_kernel": "blend_fwd", "surfel_render_forward_kernel": "blend_fwd",
            "render_forward_warp_kernel": "blend_fwd", "surfel_render_forward_warp_kernel": "blend_fwd",
            "render_backward_kernel": "blend_bwd", "surfel_render_backward_kernel": "blend_bwd",
            "render_backward_warp_kernel": "blend_bwd", "surfel_render_backward_warp_kernel": "blend_bwd",
            "preprocess_backward_kernel": "preprocess_bwd", "surfel_preprocess_backward_kernel": "preprocess_bwd",
            "scatter_kernel": "binning", "tile_sort_kernel": "binning", "tile_scan_kernel": "binning",
            "tile_sort_small_kernel": "binning", "tile_radix_kernel": "binning", "tile_sort_long_kernel": "binning"}
COLS = [("gpu__time_duration.sum", "time"), ("dram__bytes_read.sum", "dram rd"), ("dram__bytes_write.sum", "dram wr"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram %"),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm %"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue %"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ %"),
        ("launch__registers_per_thread", "regs"), ("smsp__inst_executed.sum", "warp inst"),
        ("smsp__thread_inst_executed_per_inst_executed.ratio", "thr/inst"),
        ("lts__t_sector_hit_rate.pct", "L2 hit %")]


def to_bytes(v, unit):
    f = float(v.replace(",", ""))
    return f * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1)


def main():
    rep, cfg, out = sys.argv[1], sys.argv[2], sys.argv[3]
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(txt.splitlines()))
    hdr, units, data = rows[0], rows[1], rows[2:]
    ix = {h: i for i, h in enumerate(hdr)}
    lines = [f"# ncu --set full, config {cfg} ({Path(rep).name}); one launch per kernel from the 3rd resident step\n",
             "| kernel | " + " | ".join(c[1] for c in COLS) + " |", "|---|" + "---|" * len(COLS)]
    traffic = {}
    issue = {}
    seen = set()
    for r in data:
        name = r[ix["Kernel Name"]].split("(")[0].replace("void ", "").replace("gsr::", "")
        if name in seen:            # the capture window may wrap into the next step: first launch of each kernel only
            continue
        seen.add(name)
        base = name.split("<")[0]
        cells = []
        for key, _ in COLS:
            i = ix.get(key)
            cells.append("-" if i is None else f"{r[i]} {units[i]}".strip())
        lines.append(f"| {name} | " + " | ".join(cells) + " |")
        st = STAGE_OF.get(base)
        if st:
            b = to_bytes(r[ix["dram__bytes_read.sum"]], units[ix["dram__bytes_read.sum"]]) + \
                to_bytes(r[ix["dram__bytes_write.sum"]], units[ix["dram__bytes_write.sum"]])
            traffic[st] = traffic.get(st, 0) + b
            num = lambda key: float(r[ix[key]].replace(",", "")) if key in ix and r[ix[key]] not in ("", "n/a") else None
            wi, lanes, iss = num("smsp__inst_executed.sum"), num("smsp__thread_inst_executed_per_inst_executed.ratio"), \
                num("smsp__issue_active.avg.pct_of_peak_sustained_active")
            if wi is not None and wi > issue.get(st, {}).get("warp_inst", 0):        # the stage's largest kernel
                issue[st] = {"kernel": name, "warp_inst": int(wi), "lanes_active": lanes, "issue_active_pct": iss}
    Path(out).write_text("\n".join(lines) + "\n")
    tj = ROOT / "profiles" / "ncu_traffic.json"
    allt = json.loads(tj.read_text()) if tj.exists() else {}
    allt[cfg] = {k: int(v) for k, v in traffic.items()}
    tj.write_text(json.dumps(allt, indent=1, sort_keys=True) + "\n")
    ij = ROOT / "profiles" / "ncu_issue.json"       # read by bench.py for roofline.issue
    alli = json.loads(ij.read_text()) if ij.exists() else {}
    alli[cfg] = issue
    ij.write_text(json.dumps(alli, indent=1, sort_keys=True) + "\n")
    print("\n".join(lines))


if __name__ == "__main__":
    main()
