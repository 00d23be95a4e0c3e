"""Hottest SASS instructions (by stall samples) of one kernel in an .ncu-rep, in program order
around each hot spot.  usage: python profiles/ncu_sass.py file.ncu-rep "<kernel substring>" [n_top] [context]"""
import csv, io, subprocess, sys
rep, kf = sys.argv[1], sys.argv[2]
ntop = int(sys.argv[3]) if len(sys.argv) > 3 else 12
ctx = int(sys.argv[4]) if len(sys.argv) > 4 else 3
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
fn, h, ins, seen = "", None, [], False
for r in rows:
    if not r: continue
    if r[0] in ("Function Name", "Kernel Name"):
        if seen and kf in fn: break  # first matching launch only
        fn = r[1]; continue
    if "Address" in r and "Source" in r: h = r; continue
    if h is None or len(r) != len(h) or kf not in fn: continue
    seen = True
    ins.append(r)
ia, isrc, ie, iss = h.index("Address"), h.index("Source"), h.index("Instructions Executed"), h.index("# Samples")
tot = sum(int(r[iss] or 0) for r in ins)
print(fn[:120]); print("instructions", len(ins), "samples", tot)
hot = sorted(range(len(ins)), key=lambda i: -int(ins[i][iss] or 0))[:ntop]
for i in sorted(hot):
    print("--")
    for j in range(max(0, i - ctx), min(len(ins), i + ctx + 1)):
        r = ins[j]
        print(f"{'>>' if j == i else '  '} {j:5d} {int(r[iss] or 0):6d} {100*int(r[iss] or 0)/max(tot,1):5.1f}%  {r[isrc][:110]}")
