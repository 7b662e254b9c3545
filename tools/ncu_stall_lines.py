#!/usr/bin/env python
"""Attribute stall samples of one kernel to source lines / opcodes.
   python tools/ncu_stall_lines.py prof.ncu-rep libgrackle_b200.so _ZN3grb7k_solveILi3EEEv8GrParams [stall ...]"""
import csv, re, collections, os, subprocess, sys, io, tempfile
rep, so, func = sys.argv[1:4]
stalls = sys.argv[4:] or ["stall_long_sb", "stall_barrier", "stall_wait"]
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
    m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
    if m: cur = (os.path.basename(m.group(1)), int(m.group(2))); continue
    m = re.match(r'\s+/\*([0-9a-f]+)\*/\s+(.*);', ln)
    if m: addr2[int(m.group(1), 16)] = (cur, m.group(2).strip())
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out))); hdr = rows[1]
ia = hdr.index("Address"); ii = hdr.index("Instructions Executed")
cols = {k: hdr.index(k) for k in stalls}
base = None
by = collections.defaultdict(collections.Counter); byop = collections.defaultdict(collections.Counter)
inst_line = collections.Counter(); tot = collections.Counter()
for r in rows[2:]:
    if len(r) <= ii: continue
    if r[ia] == "Address": break          # a second kernel's table follows: only the first is analysed
    a = int(r[ia], 16) if r[ia].startswith("0x") else int(r[ia])
    if base is None: base = a
    key, txt = addr2.get(a - base, (("?", 0), "?"))
    op = txt.split()[0] if not txt.startswith("@") else txt.split()[1]
    n = int(float(r[ii] or 0)); inst_line[key] += n; tot["inst"] += n
    for k, c in cols.items():
        v = int(float(r[c] or 0)); by[k][key] += v; byop[k][op] += v; tot[k] += v
print(dict(tot))
for k in stalls:
    print("==", k)
    for key, v in by[k].most_common(12): print("  %5.1f%%  %s:%d" % (100 * v / max(tot[k], 1), key[0], key[1]))
    print("  by op:", [(o, round(100 * v / max(tot[k], 1), 1)) for o, v in byop[k].most_common(6)])
print("== instructions by file")
f = collections.Counter()
for (fn, ln), n in inst_line.items(): f[fn] += n
for fn, n in f.most_common(8): print("  %5.1f%%  %s" % (100 * n / tot["inst"], fn))
