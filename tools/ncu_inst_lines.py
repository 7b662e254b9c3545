#!/usr/bin/env python
"""Executed warp-instructions of one kernel by source line and by opcode.
   python tools/ncu_inst_lines.py prof.ncu-rep libgrackle_b200.so <mangled kernel> [ntop]"""
import csv, re, collections, os, subprocess, sys, io, tempfile
rep, so, func = sys.argv[1:4]
ntop = int(sys.argv[4]) if len(sys.argv) > 4 else 40
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(so)], cwd=tmp, check=True, capture_output=True)
dis = ""
for cub in sorted(os.listdir(tmp)):          # the cubin that holds the kernel (one per translation unit)
    d = subprocess.run(["nvdisasm", "-g", os.path.join(tmp, cub)], capture_output=True, text=True).stdout
    if ".text." + func in d:
        dis = d
        break
addr2 = {}; cur = None; inside = False
for ln in dis.splitlines():
    if ln.startswith(".text."):
        inside = (ln.strip().rstrip(":") == ".text." + func); continue
    if not inside: continue
    m = re.search(r'//## File "([^"]+)", line (\d+)(.*)', ln)
    if m:
        cur = (os.path.basename(m.group(1)), int(m.group(2))); continue
    m = re.match(r'\s+/\*([0-9a-f]+)\*/\s+(.*);', ln)
    if m: addr2[int(m.group(1), 16)] = (cur, m.group(2).strip())
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out))); hdr = rows[1]
ia = hdr.index("Address"); ii = hdr.index("Instructions Executed")
base = None
line = collections.Counter(); op = collections.Counter(); tot = 0
for r in rows[2:]:
    if len(r) <= ii: continue
    if r[ia] == "Address": break          # a second kernel's table follows: only the first is analysed
    a = int(r[ia], 16) if r[ia].startswith("0x") else int(r[ia])
    if base is None: base = a
    key, txt = addr2.get(a - base, (("?", 0), "?"))
    o = txt.split()[0] if not txt.startswith("@") else txt.split()[1]
    n = int(float(r[ii] or 0)); line[key] += n; op[o.split(".")[0]] += n; tot += n
print("total warp instructions", tot)
print("== by opcode")
for o, n in op.most_common(25): print("  %5.1f%%  %s" % (100 * n / tot, o))
print("== by line")
for k, n in line.most_common(ntop): print("  %5.1f%%  %s:%d" % (100 * n / tot, k[0], k[1]))
