#!/usr/bin/env python
"""Join an ncu source-page CSV (per-SASS-instruction counts) with nvdisasm -g line info:
prints executed warp-instructions and stall samples per source file:line and per file.

  ncu -i prof.ncu-rep --page source --csv > src.csv
  cuobjdump -xelf all libgrackle_b200.so ; nvdisasm -g kernels.sm_100a.cubin > kern.dis
  python tools/ncu_hot_lines.py src.csv kern.dis _ZN3grb7k_solveILi3EEEv8GrParams
"""
import csv, re, sys, collections, os
src_csv, dis, func = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
addr2line = {}
cur = None; inside = False
for ln in open(dis):
    if ln.startswith(".text."):
        inside = (ln.strip().rstrip(":") == ".text." + func)
        continue
    if not inside: continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
    if m: cur = (os.path.basename(m.group(1)), int(m.group(2))); continue
    m = re.match(r'\s+/\*([0-9a-f]+)\*/\s+(.*);', ln)
    if m: addr2line[int(m.group(1), 16)] = (cur, m.group(2).strip())
rows = list(csv.reader(open(src_csv)))
hdr = rows[1]
ia, ii, isamp = hdr.index("Address"), hdr.index("Instructions Executed"), hdr.index("# Samples")
ino = hdr.index("stall_no_inst")
base = None
per_line = collections.Counter(); per_file = collections.Counter(); samp_line = collections.Counter()
per_op = collections.Counter(); noinst = collections.Counter()
tot = 0; tots = 0
for r in rows[2:]:
    if len(r) <= ii: continue
    a = int(r[ia], 16) if r[ia].startswith("0x") else int(r[ia])
    if base is None: base = a
    n = int(float(r[ii] or 0)); s = int(float(r[isamp] or 0))
    key, txt = addr2line.get(a - base, (("?", 0), "?"))
    per_line[key] += n; per_file[key[0]] += n; samp_line[key] += s
    noinst[key] += int(float(r[ino] or 0))
    op = txt.split()[0] if not txt.startswith("@") else txt.split()[1]
    per_op[op.split(".")[0]] += n
    tot += n; tots += s
print("total warp-instructions %d, samples %d" % (tot, tots))
print("--- by file"); 
for f, n in per_file.most_common(12): print("%6.2f%%  %s" % (100.0 * n / tot, f))
print("--- by opcode")
for f, n in per_op.most_common(25): print("%6.2f%%  %s" % (100.0 * n / tot, f))
print("--- by line (inst%%, sample%%, no_inst samples)")
for k, n in per_line.most_common(top): print("%6.2f%% %6.2f%% %7d  %s:%d" % (100.0 * n / tot, 100.0 * samp_line[k] / max(tots, 1), noinst[k], k[0], k[1]))
