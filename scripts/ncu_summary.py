"""Summarise an ncu report: key raw metrics + instruction/stall share by source line.
usage: python scripts/ncu_summary.py gpurun_out/prof.ncu-rep [top_n]"""
import collections, csv, io, subprocess, sys
rep = sys.argv[1]; topn = int(sys.argv[2]) if len(sys.argv) > 2 else 30
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
want = ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "smsp__inst_executed.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "launch__occupancy_limit_registers",
        "launch__occupancy_limit_shared_mem", "launch__grid_size", "launch__block_size",
        "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "lts__t_bytes.sum", "l1tex__t_bytes.sum"]
for r in rows[2:]:
    print("== launch")
    for i, h in enumerate(hdr):
        if h in want:
            print(f"  {h} [{units[i]}] = {r[i]}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
agg = collections.Counter(); stall = collections.Counter(); text = {}; cur = None
for r in csv.reader(io.StringIO(src)):
    if not r: continue
    if r[0] == "File Path": cur = r[1].split("/")[-1]; continue
    if r[0] in ("Function Name", "Line No"): continue
    if r[0] != "" and len(r) > 7 and r[2] == "-":
        try: ln = int(r[0]); ie = int(r[7]); s = int(r[6])
        except ValueError: continue
        agg[(cur, ln)] += ie; stall[(cur, ln)] += s; text[(cur, ln)] = r[1].strip()[:80]
tot = sum(agg.values()) or 1; ts = sum(stall.values()) or 1
byfile = collections.Counter(); sf = collections.Counter()
for k, v in agg.items(): byfile[k[0]] += v; sf[k[0]] += stall[k]
print("== warp instructions (all profiled launches):", tot)
for f in byfile: print(f"  {f}: {100*byfile[f]/tot:.1f}% inst, {100*sf[f]/ts:.1f}% stall samples")
for k, v in agg.most_common(topn):
    print(f"  {k[0][:16]:16s} {k[1]:4d} {100*v/tot:5.1f}% inst {100*stall[k]/ts:5.1f}% stall | {text[k]}")
print("== source lines by stall samples")
for k, v in stall.most_common(topn):
    print(f"  {k[0][:16]:16s} {k[1]:4d} {100*agg[k]/tot:5.1f}% inst {100*v/ts:5.1f}% stall | {text[k]}")
