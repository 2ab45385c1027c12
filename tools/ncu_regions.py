"""Time (stall samples) and executed instructions of an ncu report aggregated per source FILE and per coarse
source region of hmc_kernel.cuh.  usage: ncu_regions.py <report.ncu-rep> <object.o> <kernel mangled-name substring>"""
import collections, csv, io, os, re, subprocess, sys, tempfile
rep, obj, kname = sys.argv[1:4]
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]
ia, isrc, ismp, iex = hdr.index("Address"), hdr.index("Source"), hdr.index("# Samples"), hdr.index("Instructions Executed")
insts = [(int(r[ia], 16), r[isrc].strip(), int(r[ismp] or 0), int(r[iex] or 0)) for r in rows[hi + 1:]
         if len(r) > iex and r[ia].startswith("0x")]
base = insts[0][0]
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=tmp, capture_output=True)
line_of = {}
for cub in os.listdir(tmp):
    txt = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, cub)], capture_output=True, text=True).stdout
    inside, cur = False, ("?", 0)
    for ln in txt.splitlines():
        if ln.startswith("//---") and ".text." in ln:
            inside = kname in ln
        if not inside:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if m:
            cur = (os.path.basename(m.group(1)), int(m.group(2)))
            continue
        m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(\S.*);", ln)
        if m:
            line_of[int(m.group(1), 16)] = cur
tot = sum(i[2] for i in insts)
totx = sum(i[3] for i in insts)
S, X = collections.Counter(), collections.Counter()
# contiguous address ranges: print a run-length summary of (file) along the address axis with samples
runs = []
for addr, s, smp, ex in insts:
    f, l = line_of.get(addr - base, ("?", 0))
    S[f] += smp
    X[f] += ex
    if runs and runs[-1][0] == f:
        runs[-1][2] = addr - base
        runs[-1][3] += smp
        runs[-1][4] += ex
        runs[-1][5] += 1
        runs[-1][6].add(l)
    else:
        runs.append([f, addr - base, addr - base, smp, ex, 1, {l}])
print("file                     samples%  exec%")
for f, v in S.most_common():
    print(f"{f:24s} {100*v/tot:7.2f} {100*X[f]/totx:7.2f}")
print("\naddress-ordered regions with >=0.8% of samples:")
for f, a0, a1, smp, ex, n, ls in runs:
    if smp >= 0.008 * tot:
        print(f"{a0:#07x}-{a1:#07x} {f:22s} lines {min(ls)}-{max(ls)}  n={n:4d} samples {100*smp/tot:5.2f}% exec {100*ex/totx:5.2f}%")
