#!/usr/bin/env python
"""Markdown table of the single-GPU results from profiles/r02_final_{ours,ref}_C*.json (bench.py lines of both arms, same
box type, same harness): medians of the per-step CUDA-event times, the end-to-end modes, and the ratios.
    python tools/results_table.py"""
import json
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
NAMES = {"C1": "post_refine_gs: 50 k, SH0, 512x512, fwd+bwd", "C2": "train_3dgs: 1 M, SH3, 1297x840, fwd+bwd",
         "C3": "large-scene shape, one view on one GPU: 6 M, SH3, 1920x1080, fwd+bwd",
         "C4": "train_2dgs: 1 M surfels, colour + 7-channel map, fwd+bwd",
         "C5": "align_3dgs sweep: 200 k, SH0, 512x512, forward, all 128 poses of rotations_128.npz"}


def load(p):
    return json.loads(Path(p).read_text().strip().splitlines()[-1])


def main():
    print("| Config | Reference CUDA, resident (median) | This repo, resident (median) | Ratio | Reference CUDA, end to end | This repo, end to end: stock calls / one CUDA graph | Ratio (stock / graph) | same inputs_sha, binning_sha, num_rendered of the first view |")
    print("|---|---|---|---|---|---|---|---|")
    for c in ("C1", "C2", "C3", "C4", "C5"):
        o, r = load(ROOT / "profiles" / f"r02_final_ours_{c}.json"), load(ROOT / "profiles" / f"r02_final_ref_{c}.json")
        om, rm = o["ms_per_step_spread"]["median"], r["ms_per_step_spread"]["median"]
        re2e = r["e2e"]["ms_per_step_spread"]["median"]
        modes = o["e2e"]["modes"]
        stock = modes["stock"]["spread"]["median"]
        graph = modes.get("graph", {}).get("spread", {}).get("median")
        same = all(o["config"][k] == r["config"][k] for k in ("inputs_sha", "binning_sha", "num_rendered_first_view"))
        e2e_txt = f"{stock:.3f} ms" + (f" / **{graph:.3f} ms**" if graph else " (one batched call)")
        ratio_txt = f"{re2e / stock:.2f}x" + (f" / {re2e / graph:.2f}x" if graph else "")
        print(f"| {c} {NAMES[c]} | {rm:.3f} ms ({r['value_median'] if 'value_median' in r else r['value']:.0f} Mpix/s) | **{om:.3f} ms** ({o['value_median']:.0f} Mpix/s) | {rm / om:.2f}x | "
              f"{re2e:.3f} ms | {e2e_txt} | {ratio_txt} | {'yes' if same else 'NO'} |")


if __name__ == "__main__":
    main()
