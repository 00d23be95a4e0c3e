"""Per-source-line cost of a kernel from an .ncu-rep captured with --import-source on and built
with -lineinfo: warp instructions executed, average active threads and stall samples per CUDA
source line.  usage: python profiles/ncu_lines.py file.ncu-rep [n_top] [kernel-name substring]"""
import collections, csv, io, subprocess, sys
rep = sys.argv[1]; ntop = int(sys.argv[2]) if len(sys.argv) > 2 else 40
kfilter = sys.argv[3] if len(sys.argv) > 3 else None  # substring of the kernel (function) name
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
agg = collections.defaultdict(lambda: [0, 0, 0, ""])  # instr, thread instr, samples, text
cur_file, cur, cur_fn = "", None, ""
h = None
for r in rows:
    if not r: continue
    if r[0] == "File Path": cur_file = r[1].split("/")[-1]; continue
    if r[0] == "Function Name": cur_fn = r[1]; continue
    if r[0] == "Line No": h = r; ie = h.index("Instructions Executed"); it = h.index("Thread Instructions Executed"); isamp = h.index("# Samples"); continue
    if h is None or len(r) < len(h): continue
    if kfilter and kfilter not in cur_fn: continue
    if r[0] != "":
        cur = (cur_file, int(r[0])); agg[cur][3] = r[1].strip()[:90]
    elif cur is not None and r[ie].isdigit():
        a = agg[cur]; a[0] += int(r[ie]); a[1] += int(r[it]); a[2] += int(r[isamp] or 0)
tot_i = sum(a[0] for a in agg.values()); tot_s = sum(a[2] for a in agg.values())
print(f"total warp instr {tot_i}  samples {tot_s}")
print("  %instr %samples  lanes  file:line  source")
for k, a in sorted(agg.items(), key=lambda kv: -kv[1][2])[:ntop]:
    if a[0] == 0: continue
    print(f"  {100*a[0]/tot_i:5.1f}  {100*a[2]/max(tot_s,1):5.1f}  {a[1]/a[0]:5.1f}  {k[0]}:{k[1]}  {a[3]}")
