#!/usr/bin/env python3
"""tools/ncu_source_summary.py -- per-source-line summary of an ncu report (run where ncu is, no GPU needed).

  python tools/ncu_source_summary.py report.ncu-rep [launch_index] [top_n]

Reads `ncu --page source --csv --print-source cuda,sass` and prints, for the hottest
source lines of the selected launch: warp instructions executed, their share,
average active threads per instruction and stall samples.  Used to decide where
the issue slots of the traversal kernels go (profiles/*.md quote its output).
"""
import csv
import io
import subprocess
import sys


def main():
    rep = sys.argv[1]
    launch = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass",
                          "--launch-skip", str(launch), "--launch-count", "1"],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    fname, func, hdr, lines = "", "", None, []
    for r in rows:
        if not r:
            continue
        if r[0] == "File Path":
            fname = r[1].split("/")[-1]
        elif r[0] == "Function Name":
            func = r[1]
        elif r[0] == "Line No":
            hdr = r
        elif hdr and r[0].isdigit() and len(r) == len(hdr):
            d = dict(zip(hdr, r))
            # hdr has two "Source" columns; index explicitly
            try:
                inst = int(r[hdr.index("Instructions Executed")])
                thr = int(r[hdr.index("Thread Instructions Executed")])
                smp = int(r[hdr.index("# Samples")])
            except ValueError:
                continue
            lines.append((inst, thr, smp, fname, int(r[0]), r[1].strip()))
    tot = sum(l[0] for l in lines) or 1
    tthr = sum(l[1] for l in lines)
    tsmp = sum(l[2] for l in lines) or 1
    print(f"# {func}")
    print(f"# warp instructions {tot:,}; thread instructions {tthr:,}; avg active threads {tthr / tot:.2f}; samples {tsmp:,}")
    print(f"{'inst%':>6} {'cum%':>6} {'thr/inst':>8} {'smp%':>6}  file:line  source")
    cum = 0.0
    for inst, thr, smp, f, ln, src in sorted(lines, reverse=True)[:top]:
        cum += inst / tot
        print(f"{100 * inst / tot:6.2f} {100 * cum:6.1f} {thr / max(inst, 1):8.2f} {100 * smp / tsmp:6.2f}  {f}:{ln}  {src[:100]}")


if __name__ == "__main__":
    main()
