#!/usr/bin/env python
"""Opcode histogram of a cubin / shared library: `python tools/sass_hist.py <file> [function-substring]`.
Prints, per function (or for the whole file with --total), the count of every SASS mnemonic (modifiers stripped)."""
import collections
import re
import subprocess
import sys


def hist(path, match=None):
    out = subprocess.run(["cuobjdump", "-sass", path], capture_output=True, text=True).stdout
    fn = None
    per = collections.OrderedDict()
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            fn = m.group(1)
            per[fn] = collections.Counter()
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if m and fn:
            per[fn][m.group(1).split(".")[0]] += 1
    if match:
        per = {k: v for k, v in per.items() if match in k}
    return per


if __name__ == "__main__":
    total = "--total" in sys.argv
    args = [a for a in sys.argv[1:] if a != "--total"]
    per = hist(args[0], args[1] if len(args) > 1 else None)
    if total:
        c = collections.Counter()
        for v in per.values():
            c.update(v)
        per = {"<all %d functions>" % len(per): c}
    for fn, c in per.items():
        print(f"{fn}: {sum(c.values())} instructions")
        print("  " + "  ".join(f"{k} {v}" for k, v in c.most_common()))
