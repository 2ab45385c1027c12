"""Stall samples / executed warp-instructions of one kernel of an ncu report along the SASS address axis, in buckets,
with the share of fp64-pipe opcodes and the mean issue rate proxy (exec per sample).  usage: ncu_buckets.py <report.ncu-rep> [bucket_bytes]"""
import csv, io, subprocess, sys, collections
rep = sys.argv[1]
bucket = int(sys.argv[2], 0) if len(sys.argv) > 2 else 0x800
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]
ia, isrc, ismp, iex = hdr.index("Address"), hdr.index("Source"), hdr.index("# Samples"), hdr.index("Instructions Executed")
insts = [(int(r[ia], 16), r[isrc].strip(), int(r[ismp] or 0), int(r[iex] or 0)) for r in rows[hi + 1:]
         if len(r) > iex and r[ia].startswith("0x")]
base = insts[0][0]
tot = sum(i[2] for i in insts); totx = sum(i[3] for i in insts)
B = collections.OrderedDict()
for a, s, smp, ex in insts:
    k = (a - base) // bucket
    b = B.setdefault(k, [0, 0, 0, 0, collections.Counter()])
    op = s.split()[0] if not s.startswith("@") else s.split()[1]
    op = op.split(".")[0]
    b[0] += smp; b[1] += ex; b[3] += 1
    if op in ("DFMA", "DMUL", "DADD", "DSETP"): b[2] += ex
    b[4][op] += ex
print(f"total samples {tot}, warp-instr {totx}")
print(" addr      samples%  exec%  fp64%ofexec  exec/static   top opcodes")
for k, (smp, ex, f64, n, ops) in B.items():
    if smp < 0.002 * tot and ex < 0.002 * totx: continue
    top = " ".join(f"{o}:{100*v/max(ex,1):.0f}" for o, v in ops.most_common(4))
    print(f"{k*bucket:#08x} {100*smp/tot:7.2f} {100*ex/totx:7.2f} {100*f64/max(ex,1):7.1f} {ex/max(n,1)/1e3:10.0f}k   {top}")
