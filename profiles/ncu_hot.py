"""Summarise an .ncu-rep: key raw metrics + hottest SASS lines (needs ncu; no GPU).
usage: python profiles/ncu_hot.py file.ncu-rep [n_top]"""
import collections, csv, subprocess, sys, io
rep = sys.argv[1]; ntop = int(sys.argv[2]) if len(sys.argv) > 2 else 40
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
h = rows[0]
keys = ["gpu__time_duration.sum", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "sm__inst_executed_pipe_xu_realtime.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__cycles_active.avg", "sm__cycles_elapsed.max"]
for v in rows[2:]:
    print("== kernel:", v[h.index("Kernel Name")][:100])
    for k in keys:
        for i, n in enumerate(h):
            if n == k or n.endswith("." + k):
                print(f"  {k} = {v[i]} {rows[1][i]}")
                break
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hi = [i for i, r in enumerate(rows) if "Address" in r and "Source" in r][0]
h = rows[hi]; data = [r for r in rows[hi + 1:] if len(r) == len(h)]
ia, isrc, ie, iss, iav = h.index("Address"), h.index("Source"), h.index("Instructions Executed"), h.index("# Samples"), h.index("Avg. Threads Executed")
tot = sum(int(r[ie]) for r in data if r[ie].isdigit()); tots = sum(int(r[iss]) for r in data if r[iss].isdigit())
print("total warp instr", tot, "samples", tots)
hist = collections.Counter(); hs = collections.Counter()
for r in data:
    if not r[ie].isdigit(): continue
    t = r[isrc].split()
    op = (t[1] if t[0].startswith("@") else t[0]).split(".")[0]
    hist[op] += int(r[ie]); hs[op] += int(r[iss] or 0)
for op, c in hist.most_common(18):
    print(f"  {op:10s} {100*c/tot:5.1f}% instr  {100*hs[op]/max(tots,1):5.1f}% samples")
top = sorted([r for r in data if r[iss].isdigit()], key=lambda r: -int(r[iss]))[:ntop]
for r in top:
    print(r[ia][-5:], r[iss].rjust(6), r[ie].rjust(9), r[iav][:5].rjust(6), r[isrc][:100])
