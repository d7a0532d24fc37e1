#!/usr/bin/env python
"""Per-source-line summary of an ncu report's source page.

  python tools/ncu_lines.py gpurun_out/prof.ncu-rep <kernel-regex> [top_n] [launch_skip]

Runs `ncu -i REP --page source --print-source cuda,sass --csv` and prints, for the
lines with the most warp-stall samples: samples, share, executed warp instructions
and the source text.  Needs -lineinfo at compile time and --import-source on.
"""
import csv
import io
import os
import subprocess
import sys


def main():
    rep, kern = sys.argv[1], sys.argv[2]
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
    skip = sys.argv[4] if len(sys.argv) > 4 else "0"
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv",
                          "--kernel-name", f"regex:{kern}", "--launch-skip", skip, "--launch-count", "1"],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
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
        if hdr is None or len(r) < len(hdr) or r[2] != "-":
            continue                      # SASS rows carry an address in column 2
        try:
            smp = int(r[hdr.index("# Samples")])
            ins = int(r[hdr.index("Instructions Executed")])
        except ValueError:
            continue
        stalls = sorted(((int(r[k]) if r[k].isdigit() else 0, hdr[k].replace("stall_", "")) for k in range(len(hdr))
                         if hdr[k].startswith("stall_") and "Not Issued" not in hdr[k]), reverse=True)[:3]
        lines.append((smp, ins, fname, r[0], r[1].strip(), " ".join(f"{n}:{v}" for v, n in stalls if v)))
    tot_s = sum(x[0] for x in lines) or 1
    tot_i = sum(x[1] for x in lines) or 1
    print(f"total samples {tot_s}, total warp instructions {tot_i}")
    byfile = {}
    for s, i, f, *_ in lines:
        a = byfile.setdefault(f, [0, 0])
        a[0] += s
        a[1] += i
    for f, (s, i) in byfile.items():
        print(f"  {f:22s} samples {s / tot_s:6.1%}  instr {i / tot_i:6.1%}")
    if os.environ.get("BYINSTR"):
        lines.sort(key=lambda x: -x[1])
    else:
        lines.sort(reverse=True)
    for s, i, f, ln, src, st in lines[:top]:
        print(f"{s:7d} {s / tot_s:6.1%} {i:10d} {i / tot_i:6.1%}  {f}:{ln:>4s}  {src[:90]}   [{st}]")


if __name__ == "__main__":
    main()
