#!/usr/bin/env python
"""Group the per-line table of tools/ncu_lines.py by phase of the evaluation kernel.
usage: python tools/ncu_groups.py lines_all.txt"""
import re, sys
groups = [
 ("philox", "pda_device.cuh", 25, 55), ("u01", "pda_device.cuh", 56, 60),
 ("det_logf", "pda_device.cuh", 61, 84), ("det_log f64", "pda_device.cuh", 85, 108), ("det_exp f64", "pda_device.cuh", 109, 130),
 ("det_norm_icdf", "pda_device.cuh", 131, 164), ("pow6/root6", "pda_device.cuh", 165, 182),
 ("xor128", "pda_device.cuh", 183, 209), ("binomial_inversion f64", "pda_device.cuh", 210, 235),
 ("btpe", "pda_device.cuh", 236, 316), ("binomial dispatch", "pda_device.cuh", 317, 335),
 ("poisson_itr", "pda_device.cuh", 355, 378),
 ("icdf_fast", "pda_device.cuh", 379, 417), ("binomial_fast", "pda_device.cuh", 418, 453),
 ("poisson_fast", "pda_device.cuh", 454, 470), ("cell_of", "pda_device.cuh", 471, 480), ("hist_index", "pda_device.cuh", 481, 493),
 ("pick", "pda_eval.cuh", 64, 76), ("lane_g/off", "pda_eval.cuh", 100, 112),
 ("scan_linear", "pda_eval.cuh", 128, 158), ("scan_indexed", "pda_eval.cuh", 159, 193),
 ("lane_begin", "pda_eval.cuh", 194, 221), ("lane_step", "pda_eval.cuh", 222, 311),
 ("kernel prologue", "pda_eval.cuh", 355, 403), ("refill", "pda_eval.cuh", 404, 499),
 ("step+finish", "pda_eval.cuh", 500, 528), ("epilogue", "pda_eval.cuh", 529, 556),
]
tot = {}
for l in open(sys.argv[1]).read().split('\n')[2:]:
    m = re.match(r'(\S.*?)\s+([\d.]+)\s+([\d.]+)\s+([\d.]+)\s+(\d+)', l)
    if not m: continue
    key, wi, ss, act = m.group(1), float(m.group(2)), float(m.group(3)), float(m.group(4))
    g = None
    mm = re.match(r'(\S+):(\d+)$', key)
    if mm:
        f, n = mm.group(1), int(mm.group(2))
        for name, ff, a, b in groups:
            if f == ff and a <= n <= b: g = name; break
        if g is None: g = f
    else: g = key
    t = tot.setdefault(g, [0, 0, 0])
    t[0] += wi; t[1] += ss; t[2] += wi * act
print(open(sys.argv[1]).readline().strip())
for g, t in sorted(tot.items(), key=lambda kv: -kv[1][0]):
    if t[0] < 0.005 and t[1] < 0.05: continue
    print(f"{g:58s} warp-inst {t[0]:6.2f}%  samples {t[1]:6.2f}%  act {t[2]/max(t[0],1e-9):5.1f}")
