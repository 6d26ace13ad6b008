#!/usr/bin/env python3
"""tools/ncu_sass_phases.py report.ncu-rep launch 'name:lo-hi,name:lo-hi,...'
Sums ncu's per-SASS-instruction counters over address ranges (hex offsets from the kernel start,
as printed by cuobjdump -sass) so the issue slots of a kernel can be attributed to its phases."""
import csv
import io
import subprocess
import sys


def load(rep, launch):
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--launch-skip", str(launch),
                          "--launch-count", "1"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    hdr = rows[hi]
    ii, it, ism = (hdr.index(k) for k in ("Instructions Executed", "Thread Instructions Executed", "# Samples"))
    data, base = [], None
    for r in rows[hi + 1:]:
        if len(r) < len(hdr) or not r[0].startswith("0x"):
            continue
        a = int(r[0], 16)
        base = a if base is None else base
        data.append((a - base, int(r[ii]), int(r[it]), int(r[ism]), r[1].strip()))
    return rows[0][1] if rows and len(rows[0]) > 1 else "", data


def main():
    rep, launch = sys.argv[1], int(sys.argv[2])
    name, data = load(rep, launch)
    tot = sum(d[1] for d in data) or 1
    tsm = sum(d[3] for d in data) or 1
    print(f"# {name}: {tot / 1e6:.1f}M warp instructions, {sum(d[2] for d in data) / tot:.2f} active threads/inst")
    spec = sys.argv[3] if len(sys.argv) > 3 else "all:0-100000"
    for part in spec.split(","):
        nm, rg = part.split(":")
        lo, hi = (int(x, 16) for x in rg.split("-"))
        s = [d for d in data if lo <= d[0] < hi]
        wi = sum(d[1] for d in s)
        print(f"{nm:28s} [{lo:#06x},{hi:#06x}) static={len(s):4d} warp-inst={wi / 1e6:9.1f}M ({100 * wi / tot:5.1f}%) "
              f"threads/inst={sum(d[2] for d in s) / max(wi, 1):5.2f} samples={100 * sum(d[3] for d in s) / tsm:5.1f}%")
    if len(sys.argv) > 4:  # print individual instructions at these offsets
        want = {int(x, 16) for x in sys.argv[4].split(",")}
        for d in data:
            if d[0] in want:
                print(f"  {d[0]:#06x} exec={d[1]:>12,} threads/inst={d[2] / max(d[1], 1):5.2f}  {d[4]}")


if __name__ == "__main__":
    main()
