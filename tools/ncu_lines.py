#!/usr/bin/env python
"""Top CUDA source lines of an .ncu-rep by warp-stall samples (needs -lineinfo + --import-source on):
    python tools/ncu_lines.py gpurun_out/x.ncu-rep [top_n]"""
import csv
import subprocess
import sys


def main():
    rep = sys.argv[1]
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    fname, hdr, lines = "", None, []
    for r in rows:
        if not r:
            continue
        if r[0] == "File Path":
            fname = r[1].split("/")[-1]
            continue
        if r[0] == "Line No":
            hdr = r
            continue
        if hdr is None or len(r) < len(hdr) or not r[0].strip():
            continue
        idx = {n: i for i, n in enumerate(hdr)}
        try:
            s = int(r[idx["# Samples"]] or 0)
        except ValueError:
            continue
        stalls = sorted(((int(r[i] or 0), n) for n, i in idx.items() if n.startswith("stall_") and "Not Issued" not in n and r[i].isdigit()), reverse=True)[:3]
        lines.append((s, fname, r[0], r[1].strip()[:100], r[idx["Instructions Executed"]], stalls))
    tot = sum(x[0] for x in lines)
    print("total samples", tot)
    for s, f, ln, src, ex, st in sorted(lines, reverse=True)[:top]:
        print(f"{s:6d} {100.0 * s / max(tot, 1):5.1f}%  {f}:{ln}  ex={ex}  {src}   {[(a, b[6:]) for a, b in st if a]}")


if __name__ == "__main__":
    main()
