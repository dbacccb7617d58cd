"""Build / extend profiles/traffic.json (read by bench.py: `roofline.traffic`) from ncu CSVs written by tools/ncu_traffic.sh.

    python tools/make_traffic_json.py <key> <csv> <launches-of-the-step-captured> [<source text>]

key = "<nside>/<fofbgtype>/<pp_threshold>" as bench.py forms it.  The CSV holds every launch of `warmup + steps (+ e2e)` runs of
the step; a timer group's traffic per launch (= per step: every group below is bracketed once per step) is the sum over its
kernels' DRAM bytes divided by the number of steps captured."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from ncu_traffic import load  # noqa: E402

# timer group (vrgpu_kernel_times name) -> kernels launched inside its bracket
GROUPS = {
    "ptree.walk": ["k_pt_walk"],
    "fof6d.link": ["k_link6"],
    "fof3d.link_cells": ["k_link_blocks", "k_link_queue", "k_block_roots", "k_link_cells"],
    "ptree.build": ["k_kd_big", "k_kd_small", "k_kd_level_flags", "k_kd_level_children", "k_kd_perm", "k_kd_init_rec"],
    "ptree.moments": ["k_pt_big_sums", "k_pt_sub_sums", "k_pt_leaf_sums", "k_pt_cm", "k_pt_bmax", "k_pt_r2max"],
    "fof6d.dispersion": ["k_group_dispersion_big", "k_group_dispersion", "k_seq_"],
    "potential.direct.sym": ["k_potential_sym"],
    "potential.direct.tiled": ["k_potential_direct"],
    "fof3d.cell_keys": ["k_cell_keys"],
    "fof3d.gather_pos": ["k_gather_pos"],
}


def main():
    key, path, nsteps = sys.argv[1], sys.argv[2], int(sys.argv[3])
    source = sys.argv[4] if len(sys.argv) > 4 else f"ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum --clock-control none ({os.path.basename(path)})"
    per = load(path)
    out = {}
    for group, kernels in GROUPS.items():
        tot = sum(a["bytes"] for k, a in per.items() if any(name in k for name in kernels))
        if tot > 0:
            out[group] = tot / nsteps
    # the sort is used by several stages: count the 3DFOF stage's passes only when the caller says how many sorts a step holds
    dst = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "profiles", "traffic.json")
    tab = json.load(open(dst)) if os.path.exists(dst) else {}
    tab[key] = {"dram_bytes_per_launch": out, "source": source, "steps_captured": nsteps}
    json.dump(tab, open(dst, "w"), indent=1, sort_keys=True)
    for k, v in sorted(out.items(), key=lambda kv: -kv[1]):
        print(f"{k:28s} {v / 1e6:12.1f} MB per step")


if __name__ == "__main__":
    main()
