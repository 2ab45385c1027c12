"""Joins an ncu report's per-instruction stall samples with source lines.

usage: ncu_lines.py <report.ncu-rep> <object.o or .so with the kernel> <kernel mangled-name substring>
Prints samples / instructions aggregated per (file, line) and per SASS opcode class.
"""
import collections
import csv
import io
import re
import subprocess
import sys
import tempfile
import os

rep, obj, kname = sys.argv[1:4]
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]
ia, isrc, ismp, iex = hdr.index("Address"), hdr.index("Source"), hdr.index("# Samples"), hdr.index("Instructions Executed")
insts = []
for r in rows[hi + 1:]:
    if len(r) <= iex or not r[ia].startswith("0x"):
        continue
    insts.append((int(r[ia], 16), r[isrc].strip(), int(r[ismp] or 0), int(r[iex] or 0)))
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
by_line = collections.Counter()
ex_line = collections.Counter()
by_op = collections.Counter()
ex_op = collections.Counter()
tot = sum(i[2] for i in insts)
totx = sum(i[3] for i in insts)
for addr, s, smp, ex in insts:
    key = line_of.get(addr - base, ("?", 0))
    by_line[key] += smp
    ex_line[key] += ex
    op = s.split()[1] if s.startswith("@") else s.split()[0]
    op = op.split(".")[0]
    by_op[op] += smp
    ex_op[op] += ex
print(f"total samples {tot}, warp-instructions executed {totx}")
print("--- by source line (samples%, exec%)")
for k, v in by_line.most_common(45):
    print(f"{100*v/tot:6.2f}% {100*ex_line[k]/totx:6.2f}%  {k[0]}:{k[1]}")
print("--- by opcode (samples%, exec%)")
for k, v in by_op.most_common(25):
    print(f"{100*v/tot:6.2f}% {100*ex_op[k]/totx:6.2f}%  {k}")
