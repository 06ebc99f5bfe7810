#!/usr/bin/env python
"""Group the per-line table of tools/ncu_lines.py by the function the line sits in (function
ranges are found in the current sources: a definition starts at a line containing `name(` at brace
depth 0 of the namespace and ends where its braces close).
usage: python tools/ncu_groups.py lines_all.txt"""
import os, re, sys

SRC = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "gsmfretda_b200", "csrc")


def function_ranges(path):
    out = []
    depth = 0
    cur = None
    pending = None
    for no, line in enumerate(open(path), 1):
        code = line.split("//")[0]
        if depth <= 1 and cur is None:
            m = re.search(r"\b([A-Za-z_]\w*)\s*\(", code)
            if m and not code.lstrip().startswith(("#", "}", "if", "for", "while", "return", "static_assert")):
                pending = (m.group(1), no)
        opens, closes = code.count("{"), code.count("}")
        if pending and opens and cur is None:
            cur = (pending[0], pending[1], depth)
            pending = None
        depth += opens - closes
        if cur and depth <= cur[2]:
            out.append((cur[0], cur[1], no))
            cur = None
    return out


ranges = {}
for f in os.listdir(SRC):
    if f.endswith((".cuh", ".cu", ".h")):
        ranges[f] = function_ranges(os.path.join(SRC, f))

tot = {}
for l in open(sys.argv[1]).read().split("\n")[2:]:
    m = re.match(r"(\S.*?)\s+([\d.]+)\s+([\d.]+)\s+([\d.]+)\s+(\d+)", l)
    if not m:
        continue
    key, wi, ss, act = m.group(1), float(m.group(2)), float(m.group(3)), float(m.group(4))
    g = key
    mm = re.match(r"(\S+):(\d+)$", key)
    if mm:
        f, n = mm.group(1), int(mm.group(2))
        g = f
        for name, a, b in ranges.get(f, []):
            if a <= n <= b:
                g = f"{f}:{name}"
    t = tot.setdefault(g, [0, 0, 0])
    t[0] += wi
    t[1] += ss
    t[2] += wi * act
print(open(sys.argv[1]).readline().strip())
for g, t in sorted(tot.items(), key=lambda kv: -kv[1][0]):
    if t[0] < 0.05 and t[1] < 0.05:
        continue
    print(f"{g:50s} warp-inst {t[0]:6.2f}%  samples {t[1]:6.2f}%  act {t[2]/max(t[0],1e-9):5.1f}")
