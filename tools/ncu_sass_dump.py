#!/usr/bin/env python3
"""tools/ncu_sass_dump.py report.ncu-rep [launch] -- per-SASS-instruction counters of one launch:
offset, warp-level executions, average active threads, stall samples, instruction."""
import csv
import io
import subprocess
import sys


def load(rep, launch=0):
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--launch-skip", str(launch),
                          "--launch-count", "1"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    hdr = rows[hi]
    ii, it, ism, src = (hdr.index(k) for k in ("Instructions Executed", "Thread Instructions Executed",
                                               "# Samples", "Source"))
    data, base, seen = [], None, set()
    for r in rows[hi + 1:]:
        if len(r) < len(hdr) or not r[0].startswith("0x"):
            continue
        a = int(r[0], 16)
        if a in seen:      # the CSV repeats the listing once per view
            break
        seen.add(a)
        base = a if base is None else base
        data.append((a - base, int(r[ii]), int(r[it]), int(r[ism]), r[src].strip()))
    return data


if __name__ == "__main__":
    data = load(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 0)
    tot = sum(d[1] for d in data) or 1
    print(f"# {tot / 1e6:.1f}M warp instructions, {sum(d[2] for d in data) / tot:.2f} threads/inst, {len(data)} static")
    for d in data:
        print(f"{d[0]:#06x} {d[1] / 1e6:8.2f}M {d[2] / max(d[1], 1):5.1f} {d[3]:6d}  {d[4]}")
