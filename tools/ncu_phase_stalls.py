"""Per address range (phase) stall-reason breakdown of one kernel of an ncu report.
usage: ncu_phase_stalls.py <report.ncu-rep> name:lo-hi [name:lo-hi ...]   (hex byte offsets from the kernel start)"""
import csv, io, subprocess, sys, collections
rep = sys.argv[1]
phases = []
for a in sys.argv[2:]:
    nm, r = a.split(":"); lo, hi = r.split("-"); phases.append((nm, int(lo, 16), int(hi, 16)))
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hi_ = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi_]
cols = [c for c in hdr if c.startswith("stall_") and "Not Issued" not in c]
idx = {c: hdr.index(c) for c in cols}
ia, ismp, iex = hdr.index("Address"), hdr.index("# Samples"), hdr.index("Instructions Executed")
data = [r for r in rows[hi_ + 1:] if len(r) > iex and r[ia].startswith("0x")]
base = int(data[0][ia], 16)
tot = sum(int(r[ismp] or 0) for r in data)
for nm, lo, hi in phases:
    agg = collections.Counter(); smp = ex = 0
    for r in data:
        a = int(r[ia], 16) - base
        if lo <= a < hi:
            smp += int(r[ismp] or 0); ex += int(r[iex] or 0)
            for c in cols: agg[c] += int(r[idx[c]] or 0)
    s = sum(agg.values()) or 1
    print(f"{nm:8s} samples {100*smp/tot:5.1f}%  warp-instr {ex/1e6:7.1f}M  " + " ".join(f"{c[6:]}:{100*v/s:.0f}" for c, v in agg.most_common(8)))
