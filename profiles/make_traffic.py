#!/usr/bin/env python
"""profiles/traffic.json from the `ncu --set full` captures of profiles/capture.sh: per workload and kernel, the
median over the captured launches of dram__bytes_read.sum + dram__bytes_write.sum (bytes per launch).  bench.py
reports it as roofline.traffic.  Usage: python profiles/make_traffic.py <tag> (reads gpurun_out/<tag>_full_<wl>.ncu-rep)."""
import csv
import json
import os
import statistics
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
SHORT = ("k_core", "k_laser", "k_grid_tma", "k_grid_gather", "k_grid", "k_explore", "k_ga3c_tc", "k_ga3c_state", "k_ga3c_fwd",
         "k_rvo4", "k_rvo", "k_flatten")


def main(tag):
    dst = os.path.join(ROOT, "profiles", "traffic.json")
    out = json.load(open(dst)) if os.path.exists(dst) else {}  # a partial pass refreshes only the workloads it captured
    for wl in ("c2", "c3", "full", "c4", "c5", "rvo"):
        path = os.path.join(ROOT, "gpurun_out", "%s_full_%s.ncu-rep" % (tag, wl))
        if not os.path.exists(path):
            continue
        txt = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
        rows = list(csv.reader(txt.splitlines()))
        hdr, units = rows[0], rows[1]
        col = {h: i for i, h in enumerate(hdr)}
        per = {}
        for r in rows[2:]:
            name = r[col["Kernel Name"]]
            short = next((s for s in SHORT if s + "(" in name or s + "<" in name), None)
            if short is None:
                continue
            tot = 0.0
            for m in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
                tot += float(r[col[m]].replace(",", "")) * UNIT[units[col[m]]]
            per.setdefault(short, []).append(tot)
        out[wl] = {k: statistics.median(v) for k, v in per.items()}
    json.dump(out, open(dst, "w"), indent=1, sort_keys=True)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else "r1g")
