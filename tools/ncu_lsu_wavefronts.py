#!/usr/bin/env python
"""LSU-pipe accounting of one kernel from an ncu report's source page (SASS view): shared-memory
wavefronts, global tag requests and L2 sectors per opcode and per instruction, normalised to one
warp-subcycle (= executions of the lockstep barrier, or --per N).
   python tools/ncu_lsu_wavefronts.py prof.ncu-rep [--per N] [--top 30]"""
import csv, io, subprocess, sys, collections, argparse
ap = argparse.ArgumentParser(); ap.add_argument("rep"); ap.add_argument("--per", type=float, default=0)
ap.add_argument("--top", type=int, default=30); a = ap.parse_args()
out = subprocess.run(["ncu", "-i", a.rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out))); hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
def num(r, k):
    try: return float(r[ix[k]] or 0)
    except Exception: return 0.0
recs = []
for r in rows[2:]:
    if len(r) < len(hdr) or r[0] == "Address": break
    src = r[ix["Source"]].strip()
    toks = src.split()
    op = toks[1] if toks and toks[0].startswith("@") else (toks[0] if toks else "?")
    recs.append(dict(src=src, op=op, n=num(r, "Instructions Executed"), thr=num(r, "Thread Instructions Executed"),
                     shw=num(r, "L1 Wavefronts Shared"), shi=num(r, "L1 Wavefronts Shared Ideal"),
                     tag=num(r, "L1 Tag Requests Global"), l2=num(r, "L2 Theoretical Sectors Global"),
                     samples=num(r, "# Samples")))
per = a.per
if not per:
    # executions of the loop body = the most frequent execution count among the kernel's instructions
    cnt = collections.Counter(x["n"] for x in recs if x["n"] > 0)
    per = cnt.most_common(1)[0][0] if cnt else 1.0
tot = sum(x["n"] for x in recs)
print("instructions per unit: %.1f   (unit count %.0f)" % (tot / per, per))
print("shared wavefronts per unit: %.1f (ideal %.1f)   global tag requests per unit: %.1f   L2 sectors per unit: %.1f"
      % (sum(x["shw"] for x in recs) / per, sum(x["shi"] for x in recs) / per, sum(x["tag"] for x in recs) / per,
         sum(x["l2"] for x in recs) / per))
by = collections.defaultdict(lambda: [0, 0, 0, 0, 0])
for x in recs:
    k = x["op"].split(".")[0] + ("." + ".".join(x["op"].split(".")[1:3]) if x["op"].startswith(("LD", "ST", "ATOM", "RED")) else "")
    b = by[k]; b[0] += x["n"]; b[1] += x["shw"]; b[2] += x["tag"]; b[3] += x["l2"]; b[4] += x["shi"]
print("== by opcode: inst/unit, shared wavefronts/unit (ideal), global tags/unit, L2 sectors/unit")
for k, b in sorted(by.items(), key=lambda kv: -kv[1][0])[:40]:
    print("  %-22s %8.1f %8.1f (%6.1f) %8.1f %8.1f" % (k, b[0] / per, b[1] / per, b[4] / per, b[2] / per, b[3] / per))
print("== top instructions by shared wavefronts")
for x in sorted(recs, key=lambda x: -x["shw"])[:a.top]:
    print("  %8.2f (ideal %6.2f) x%6.2f  %s" % (x["shw"] / per, x["shi"] / per, x["n"] / per, x["src"]))
print("== top instructions by global tag requests")
for x in sorted(recs, key=lambda x: -x["tag"])[:a.top]:
    print("  %8.2f tags %8.2f sectors x%6.2f  %s" % (x["tag"] / per, x["l2"] / per, x["n"] / per, x["src"]))
