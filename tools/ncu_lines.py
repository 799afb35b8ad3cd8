#!/usr/bin/env python
"""Per-source-line view of one kernel of an .ncu-rep (no GPU needed).

  python tools/ncu_lines.py REPORT.ncu-rep KERNEL_REGEX [--so ipk_b200/libipkgpu.so] [--top 40] [--launch N]

ncu's CSV source page is SASS-only; this joins it with `nvdisasm --print-line-info` of the cubin inside the shared
library (built with -lineinfo) by instruction order, and sums executed warp instructions, stall samples and shared
bank conflicts per CUDA source line.
"""
from __future__ import annotations

import argparse
import collections
import csv
import io
import os
import re
import subprocess
import sys
import tempfile


def sass_lines(so: str, kernel_re: str):
    """[(address offset, file, line)] for the first kernel of the cubin whose mangled name matches"""
    tmp = tempfile.mkdtemp()
    subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(so)], cwd=tmp, check=True, stdout=subprocess.DEVNULL)
    cubin = [os.path.join(tmp, f) for f in os.listdir(tmp) if f.endswith(".cubin")][0]
    txt = subprocess.run(["nvdisasm", "--print-line-info", "-c", cubin], check=True, stdout=subprocess.PIPE, text=True).stdout
    out, active, cur = [], False, ("?", 0)
    for ln in txt.splitlines():
        m = re.match(r"\s*\.section\s+\.text\.(\S+?),", ln)
        if m:
            if active and out:
                break
            active = re.search(kernel_re, m.group(1)) is not None
            continue
        if not active:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if m:
            cur = (os.path.basename(m.group(1)), int(m.group(2)))
            continue
        m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
        if m:
            out.append((int(m.group(1), 16), cur[0], cur[1], m.group(2).strip()))
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("report")
    ap.add_argument("kernel")
    ap.add_argument("--so", default=os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "ipk_b200", "libipkgpu.so"))
    ap.add_argument("--top", type=int, default=40)
    ap.add_argument("--launch", type=int, default=0, help="which matching launch of the report (0 = first)")
    ap.add_argument("--mangled", default=None, help="regex on the mangled name in the cubin (default: the kernel regex)")
    ap.add_argument("--by", default="samples", choices=["samples", "inst", "thread"])
    ap.add_argument("--sass", action="store_true", help="also list the hottest SASS instructions")
    a = ap.parse_args()
    raw = subprocess.run(["ncu", "-i", a.report, "--page", "source", "--csv", "--kernel-name", f"regex:{a.kernel}"],
                         check=True, stdout=subprocess.PIPE, text=True).stdout
    # the CSV holds one block per launch: a "Kernel Name" row, a header row, then instruction rows
    blocks, cur = [], None
    for row in csv.reader(io.StringIO(raw)):
        if not row:
            continue
        if row[0] == "Kernel Name":
            cur = {"name": row[1], "hdr": None, "rows": []}
            blocks.append(cur)
        elif cur is not None and cur["hdr"] is None:
            cur["hdr"] = row
        elif cur is not None:
            cur["rows"].append(row)
    if not blocks:
        sys.exit("no matching kernel in the report")
    b = blocks[min(a.launch, len(blocks) - 1)]
    hdr = b["hdr"]
    col = {n: i for i, n in enumerate(hdr)}
    lines = sass_lines(a.so, a.mangled or a.kernel)
    if len(lines) != len(b["rows"]):
        print(f"warning: {len(lines)} SASS instructions in the cubin vs {len(b['rows'])} in the report (rebuilt since?)", file=sys.stderr)
    n = min(len(lines), len(b["rows"]))

    def num(row, name):
        try:
            return float(row[col[name]])
        except (KeyError, ValueError):
            return 0.0
    agg = collections.defaultdict(lambda: [0.0, 0.0, 0.0, 0.0, 0])
    tot = [0.0, 0.0, 0.0, 0.0]
    hot = []
    for i in range(n):
        r = b["rows"][i]
        key = (lines[i][1], lines[i][2])
        v = (num(r, "Instructions Executed"), num(r, "# Samples"), num(r, "Thread Instructions Executed"), num(r, "L1 Conflicts Shared N-Way"))
        for j in range(4):
            agg[key][j] += v[j]
            tot[j] += v[j]
        agg[key][4] += 1
        hot.append((v[1], v[0], lines[i][3], key))
    print(f"kernel: {b['name'][:100]}")
    print(f"total warp instructions {tot[0]:.0f}, stall samples {tot[1]:.0f}, avg active threads {tot[2] / max(tot[0], 1):.1f}, shared conflicts {tot[3]:.0f}")
    print(f"{'file:line':28s} {'inst%':>7s} {'samp%':>7s} {'thr/inst':>8s} {'thrinst%':>8s} {'conflicts':>10s}  sass")
    for key, v in sorted(agg.items(), key=lambda kv: -kv[1][{'samples': 1, 'inst': 0, 'thread': 2}[a.by]])[: a.top]:
        print(f"{key[0] + ':' + str(key[1]):28s} {100 * v[0] / max(tot[0], 1):7.2f} {100 * v[1] / max(tot[1], 1):7.2f} {v[2] / max(v[0], 1):8.1f} {100 * v[2] / max(tot[2], 1):8.2f} {v[3]:10.0f}  {v[4]}")
    if a.sass:
        print("\nhottest SASS instructions (stall samples):")
        for s, e, txt, key in sorted(hot, key=lambda t: -t[0])[: a.top]:
            print(f"  {100 * s / max(tot[1], 1):6.2f}%  exec {e:12.0f}  {key[0]}:{key[1]:<5d} {txt[:90]}")


if __name__ == "__main__":
    main()
