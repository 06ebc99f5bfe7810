#!/usr/bin/env python
"""Join an ncu SASS source page with nvdisasm line info: executed thread-instructions and
stall samples per CUDA source line (inlined call sites attributed to the innermost line).
usage: python tools/ncu_lines.py prof.ncu-rep libgsmfret.so <mangled-kernel-substring> [top]"""
import csv, io, re, subprocess, sys, tempfile, os, collections

rep, so, kern = sys.argv[1], sys.argv[2], sys.argv[3]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(so)], cwd=tmp, check=True, capture_output=True)
cubin = next(f for f in sorted(os.listdir(tmp)) if f.endswith(".cubin") and kern in subprocess.run(["cuobjdump", "-elf", os.path.join(tmp, f)], capture_output=True, text=True).stdout)
dis = subprocess.run(["nvdisasm", "--print-line-info", os.path.join(tmp, cubin)], capture_output=True, text=True).stdout
lines = dis.split("\n")
# locate function
start = next(i for i, l in enumerate(lines) if l.startswith(".text.") and kern in l)
addr2line = {}
cur = None
for l in lines[start + 1:]:
    if l.startswith("//---------------------") or l.startswith(".text."):
        break
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (os.path.basename(m.group(1)), int(m.group(2)))
        continue
    m = re.match(r'^(\$[\w$.]+):', l)   # helper subroutine placed after the kernel body (no line info)
    if m:
        nm = m.group(1)
        nm = re.sub(r'^\$_ZN.*?\$', '$', nm)
        cur = ("sub " + nm[:60], 0)
        continue
    m = re.match(r'\s*/\*([0-9a-f]{4,})\*/\s+(.*?);', l)
    if m:
        addr2line[int(m.group(1), 16)] = (cur, m.group(2))
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]
c = {h: i for i, h in enumerate(hdr)}
agg = collections.defaultdict(lambda: [0, 0, 0, 0])
base = None
tot = [0, 0, 0]
for r in rows[hi + 1:]:
    if len(r) < len(hdr):
        continue
    a = int(r[c["Address"]], 16)
    if base is None:
        base = a
    key = addr2line.get(a - base, (None, ""))[0]
    ie = int(r[c["Instructions Executed"]] or 0)
    te = int(r[c["Thread Instructions Executed"]] or 0)
    ss = int(r[c["# Samples"]] or 0)
    g = agg[key]
    g[0] += ie; g[1] += te; g[2] += ss; g[3] += 1
    tot[0] += ie; tot[1] += te; tot[2] += ss
print(f"total warp-inst {tot[0]:,}  thread-inst {tot[1]:,}  samples {tot[2]:,}  avg active {tot[1]/max(1,tot[0]):.2f}")
src_cache = {}
def src(key):
    if not key or key[1] == 0: return ""
    f, n = key
    for d in ("gsmfretda_b200/csrc",):
        p = os.path.join(os.path.dirname(os.path.abspath(so)), "csrc", f)
        if os.path.exists(p):
            if p not in src_cache: src_cache[p] = open(p).read().split("\n")
            L = src_cache[p]
            return L[n - 1].strip()[:90] if n - 1 < len(L) else ""
    return ""
print(f"{'line':28s} {'warp-inst%':>10s} {'samples%':>9s} {'act':>5s} {'#sass':>5s}  source")
for key, g in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    nm = (f"{key[0]}:{key[1]}" if key[1] else key[0]) if key else "?"
    print(f"{nm:28s} {100*g[0]/tot[0]:10.2f} {100*g[2]/max(1,tot[2]):9.2f} {g[1]/max(1,g[0]):5.1f} {g[3]:5d}  {src(key)}")
