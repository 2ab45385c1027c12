"""Per source line: warp-instructions executed, average active threads, stall samples (divergence view of an ncu report).
usage: ncu_threads.py <report.ncu-rep> <object.o> <kernel mangled-name substring> [min_pct]"""
import collections, csv, io, os, re, subprocess, sys, tempfile
rep, obj, kname = sys.argv[1:4]
minpct = float(sys.argv[4]) if len(sys.argv) > 4 else 0.7
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]
ia, ismp, iex, ith = hdr.index("Address"), hdr.index("# Samples"), hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed")
insts = [(int(r[ia], 16), int(r[ismp] or 0), int(r[iex] or 0), int(r[ith] or 0)) for r in rows[hi + 1:]
         if len(r) > ith and r[ia].startswith("0x")]
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
agg = collections.defaultdict(lambda: [0, 0, 0])
for a, s, e, t in insts:
    k = line_of.get(a - base, ("?", 0))
    agg[k][0] += s; agg[k][1] += e; agg[k][2] += t
ts, te, tt = (sum(v[i] for v in agg.values()) for i in range(3))
print(f"total: samples {ts}, warp-instr {te}, avg active threads {tt / max(te, 1):.2f}")
print("samples%  exec%  avg_threads  line")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    if 100.0 * v[1] / te >= minpct or 100.0 * v[0] / ts >= minpct:
        print(f"{100.0 * v[0] / ts:7.2f} {100.0 * v[1] / te:6.2f}  {v[2] / max(v[1], 1):8.2f}    {k[0]}:{k[1]}")
