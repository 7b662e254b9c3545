#!/usr/bin/env python
"""Static SASS statistics of one kernel of the product library: instruction count, opcode
histogram, local-memory (spill) instructions.  With --dump, writes the listing.
   python tools/sass_stats.py [--so lib.so] [--kernel _ZN3grb7k_solveILi3ELi1ELb0EEEv8GrParams] [--dump out.sass]"""
import argparse, collections, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ap = argparse.ArgumentParser()
ap.add_argument("--so", default=os.path.join(ROOT, "grackle_b200", "lib", "libgrackle_b200.so"))
ap.add_argument("--kernel", default="_ZN3grb7k_solveILi3ELi1ELb0EEEv8GrParams")
ap.add_argument("--dump")
a = ap.parse_args()
out = subprocess.run(["cuobjdump", "-sass", "-fun", a.kernel, a.so], capture_output=True, text=True).stdout
ops = collections.Counter(); n = 0; lines = []
for ln in out.splitlines():
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
    if not m: continue
    txt = m.group(2).strip(); lines.append("%s  %s" % (m.group(1), txt))
    t = txt.split()
    op = t[1] if t[0].startswith("@") else t[0]
    ops[op.split(".")[0]] += 1; n += 1
    if op.startswith(("LDL", "STL")): ops["(local) " + op.split(".")[0]] += 1
print("kernel %s: %d instructions (%.1f KB)" % (a.kernel, n, n * 16 / 1024))
for k, v in ops.most_common(40): print("  %6d  %s" % (v, k))
if a.dump:
    open(a.dump, "w").write("\n".join(lines) + "\n")
