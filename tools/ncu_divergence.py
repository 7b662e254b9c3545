#!/usr/bin/env python
"""Lane utilisation of one kernel by source line: warp instructions executed, lanes that did
useful work (predicated-on thread instructions), and the lane-slots lost to divergence.
   python tools/ncu_divergence.py prof.ncu-rep libgrackle_b200.so <mangled kernel> [ntop]"""
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
it = hdr.index("Predicated-On Thread Instructions Executed")
base = None
inst = collections.Counter(); thr = collections.Counter(); tot = tott = 0
hist = collections.Counter()
for r in rows[2:]:
    if len(r) <= it: continue
    if r[ia] == "Address": break
    a = int(r[ia], 16) if r[ia].startswith("0x") else int(r[ia])
    if base is None: base = a
    key, txt = addr2.get(a - base, (("?", 0), "?"))
    n = int(float(r[ii] or 0)); t = int(float(r[it] or 0))
    inst[key] += n; thr[key] += t; tot += n; tott += t
    if n: hist[min(32, int(round(t / n)))] += n
print("warp instructions %d, useful lanes per instruction %.2f" % (tot, tott / tot))
lost = {k: 32 * inst[k] - thr[k] for k in inst}
L = sum(lost.values())
print("== instructions by average active lanes (share of warp instructions)")
acc = 0
for k in sorted(hist):
    print("  %2d lanes  %5.1f%%" % (k, 100 * hist[k] / tot))
print("== lost lane-slots by line (share of all lost, warp-inst share, avg lanes)")
for k, v in sorted(lost.items(), key=lambda kv: -kv[1])[:ntop]:
    print("  %5.1f%%  %5.1f%%  %5.1f  %s:%d" % (100 * v / L, 100 * inst[k] / tot, thr[k] / max(1, inst[k]), k[0], k[1]))
