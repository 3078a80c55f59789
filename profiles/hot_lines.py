#!/usr/bin/env python
"""Per-source-line share of executed warp instructions and stall samples from an `ncu --set full --import-source on`
report (needs -lineinfo).  Usage: python profiles/hot_lines.py report.ncu-rep [top] [kernel-substring] [samples]
(`samples`: rank the lines by stall samples instead of executed instructions)"""
import csv
import subprocess
import sys


def main(path, top=25, want=None, by=0):
    txt = subprocess.run(["ncu", "-i", path, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True,
                         text=True).stdout
    fpath, func, hdr = None, None, None
    agg = {}        # function -> {(file, line): [inst, samples, text]}
    for r in csv.reader(txt.splitlines()):
        if len(r) >= 2 and r[0] == "File Path":
            fpath = r[1].split("/")[-1]
        elif len(r) >= 2 and r[0] == "Function Name":
            func = r[1]
        elif len(r) > 8 and r[0] == "Line No":
            hdr = r
            ci, cs = hdr.index("Instructions Executed"), hdr.index("# Samples")
        elif hdr is not None and len(r) > 8 and r[0].isdigit():
            try:
                ie, ss = int(r[ci].replace(",", "") or 0), int(r[cs].replace(",", "") or 0)
            except ValueError:
                continue
            a = agg.setdefault(func, {}).setdefault((fpath, int(r[0])), [0, 0, r[1]])
            a[0] += ie
            a[1] += ss
    for f, lines in agg.items():
        if want and want not in f:
            continue
        ti, ts = sum(v[0] for v in lines.values()) or 1, sum(v[1] for v in lines.values()) or 1
        print("== %s | warp instructions (all captured launches): %d | stall samples: %d" % (f[:80], ti, ts))
        for (fp, ln), (ie, ss, text) in sorted(lines.items(), key=lambda kv: -kv[1][by])[:top]:
            print("%-18s %5d  %5.1f%% inst  %5.1f%% samples  %s" % (fp, ln, 100.0 * ie / ti, 100.0 * ss / ts, text.strip()[:90]))


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 25, sys.argv[3] if len(sys.argv) > 3 else None,
         1 if len(sys.argv) > 4 and sys.argv[4] == "samples" else 0)
