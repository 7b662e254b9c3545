#!/usr/bin/env python
"""Static estimate of exposed global-load round trips in a kernel's SASS: walk the listing in
program order, keep the set of registers with an LDG in flight, and count a 'wait event' whenever
an instruction reads (or overwrites) such a register - every load issued before that point is
then assumed to have returned.  Few events per fetch phase = loads are batched; one event per
load = the compiler serialised them (register pressure).
   python tools/sass_loadwaits.py [--so lib.so] [--kernel NAME] [--from ADDR --to ADDR] [-v]"""
import argparse, os, re, subprocess
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ap = argparse.ArgumentParser()
ap.add_argument("--so", default=os.path.join(ROOT, "grackle_b200", "lib", "libgrackle_b200.so"))
ap.add_argument("--kernel", default="_ZN3grb7k_solveILi3ELi1ELb0EEEv8GrParams")
ap.add_argument("--from", dest="lo", default=None); ap.add_argument("--to", dest="hi", default=None)
ap.add_argument("-v", action="store_true")
a = ap.parse_args()
out = subprocess.run(["cuobjdump", "-sass", "-fun", a.kernel, a.so], capture_output=True, text=True).stdout
ins = []
for ln in out.splitlines():
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
    if m: ins.append((int(m.group(1), 16), m.group(2).strip()))
lo = int(a.lo, 16) if a.lo else 0
hi = int(a.hi, 16) if a.hi else 1 << 40
def regs(tok):
    out = []
    for m in re.finditer(r"\bR(\d+)(\.64)?\b", tok):
        r = int(m.group(1)); out.append(r)
        if m.group(2): out.append(r + 1)
    return out
inflight = {}     # reg -> address of the load
events = []; nload = 0; batch = 0
for addr, txt in ins:
    if addr < lo or addr > hi: continue
    t = txt.split()
    body = txt[len(t[0]):].strip() if t[0].startswith("@") else txt
    op = body.split()[0]
    operands = body[len(op):]
    parts = [p.strip() for p in operands.split(",")]
    if op.startswith("LDG"):
        width = 4 if ".128" in op else (2 if ".64" in op else 1)
        dst = regs(parts[0])[0]
        # address registers are read
        used = [r for p in parts[1:] for r in regs(p)]
        hit = [r for r in used if r in inflight]
        if hit:
            events.append((addr, txt, len(inflight) // 1)); inflight.clear()
        for k in range(width): inflight[dst + k] = addr
        nload += 1
        continue
    # destination = first operand for most ops; double-precision ops touch pairs
    wide = op.startswith(("DFMA", "DMUL", "DADD", "DSETP", "STS.64", "STS.128"))
    used = []
    for p in parts:
        for r in regs(p):
            used.append(r)
            if wide: used.append(r + 1)
    hit = [r for r in used if r in inflight]
    if hit:
        events.append((addr, txt, len(inflight) // 4)); inflight.clear()
print("%d LDG, %d wait events in range" % (nload, len(events)))
if a.v:
    for addr, txt, n in events: print("  %06x  (%2d loads outstanding)  %s" % (addr, n, txt))
