#!/usr/bin/env python
"""Opcode mix of the executed instructions of the first kernel in an .ncu-rep (source page): per opcode, and
folded into the pipe classes SURVEY.md 8(d) counts by (ALU / FMA / FP64 / XU / LSU / control / uniform / other).
With the number of burst-samples of the launch it also prints thread-instructions per sample per class — the
executed side of the "executed vs algorithmic" comparison in DESIGN.md section 8.
usage: python tools/ncu_opmix.py prof.ncu-rep [top [samples_per_launch [algorithmic_ops_per_sample]]]"""
import csv, io, collections, subprocess, sys

PIPES = {
    "ALU": "LOP3 ISETP SHF SEL IADD3 IADD LEA VIADD PRMT FSETP FSEL VIMNMX IMNMX FMNMX POPC FLO BREV PLOP3 "
           "I2FP F2FP P2R R2P SGXT BMSK IABS ICMP FCHK FSET CS2R MOV VABSDIFF HSETP2 HMNMX2 ISCADD LOP".split(),
    "FMA": "FFMA FMUL FADD IMAD IMUL HFMA2 HADD2 HMUL2 IDP IDP4A FSWZADD".split(),
    "FP64": "DFMA DMUL DADD DSETP DMNMX".split(),
    "XU": "MUFU I2F F2I F2F I2I FRND".split(),
    "LSU": "LDG STG LDS STS LDL STL LD ST ATOM ATOMS ATOMG RED LDSM SHFL MATCH LDC REDUX CCTL MEMBAR ERRBAR "
           "QSPC LDGSTS LDGDEPBAR DEPBAR FENCE".split(),
    "control": "BRA BSSY BSYNC BREAK EXIT RET CALL JMP JMX BRX WARPSYNC BAR NOP VOTE YIELD NANOSLEEP KILL BPT "
               "BMOV S2R RPCMOV ELECT ENDCOLLECTIVE".split(),
    "uniform": "UMOV UIADD3 ULOP3 ULEA USHF UISETP USEL UPRMT UFLO UPOPC UBREV UIMAD UF2FP ULDC LDCU R2UR S2UR "
               "UBMSK UMEMSETS UP2UR UR2UP UPLOP3 UCLEA UFFMA UFADD UFMUL UFSETP UI2FP UVIADD UVIMNMX UFSEL "
               "UFMNMX UI2F UF2I UISCADD UGETNEXTWORKID CREDUX".split(),
}
OP2PIPE = {op: p for p, ops in PIPES.items() for op in ops}

out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv"], capture_output=True, text=True).stdout
top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
samples = float(sys.argv[3]) if len(sys.argv) > 3 else None
algo = float(sys.argv[4]) if len(sys.argv) > 4 else None
rows = list(csv.reader(io.StringIO(out)))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
h = rows[hi]
ie, it, src = h.index("Instructions Executed"), h.index("Thread Instructions Executed"), h.index("Source")
tot, thr, n = collections.Counter(), collections.Counter(), collections.Counter()
for r in rows[hi + 1:]:
    if len(r) <= it or not r[ie].isdigit():
        continue
    t = r[src].split()
    op = (t[1] if t[0].startswith('@') else t[0]).split('.')[0]
    tot[op] += int(r[ie]); thr[op] += int(r[it]); n[op] += 1
T, TT = sum(tot.values()), sum(thr.values())
print(f"total warp-inst {T:,}  thread-inst {TT:,}  lanes/inst {TT / T:.2f}")
pw, pt = collections.Counter(), collections.Counter()
unknown = collections.Counter()
for op in tot:
    p = OP2PIPE.get(op, "other")
    pw[p] += tot[op]; pt[p] += thr[op]
    if p == "other":
        unknown[op] += tot[op]
print("\nby pipe class" + (f" (per burst-sample of {samples:,.0f})" if samples else ""))
print(f"{'class':9s} {'warp-inst':>9s} {'thread-inst':>11s} {'lanes':>6s}" + (f" {'thread-inst/sample':>19s}" if samples else ""))
for p in ("ALU", "FMA", "FP64", "XU", "LSU", "control", "uniform", "other"):
    if not pw[p]:
        continue
    line = f"{p:9s} {pw[p] / T * 100:8.2f}% {pt[p] / TT * 100:10.2f}% {pt[p] / pw[p]:6.1f}"
    if samples:
        line += f" {pt[p] / samples:19.1f}"
    print(line)
if samples:
    print(f"{'all':9s} {'':9s} {'':11s} {TT / T:6.1f} {TT / samples:19.1f}" +
          (f"   = {TT / samples / algo:.2f} x the {algo:.0f} algorithmic" if algo else ""))
if unknown:
    print("other =", ", ".join(f"{k} {v / T * 100:.2f}%" for k, v in unknown.most_common(12)))
print("\nby opcode")
for k, v in tot.most_common(top):
    print(f"{k:12s} {v / T * 100:6.2f}%  act {thr[k] / max(v, 1):5.1f}  {OP2PIPE.get(k, 'other'):8s} ({n[k]} sass)")
