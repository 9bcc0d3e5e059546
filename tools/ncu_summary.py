"""Summarise an .ncu-rep (read on the CPU box): key metrics, stall totals, opcode mix per warp-step."""
import csv, re, collections, subprocess, sys, io
rep = sys.argv[1]; steps = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
r = list(csv.reader(io.StringIO(raw))); hdr, units, vals = r[0], r[1], r[2]
keys = ['gpu__time_duration.sum', 'launch__grid_size', 'launch__block_size', 'launch__registers_per_thread', 'smsp__inst_executed.sum',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_fmaheavy.avg.pct_of_peak_sustained_active',
        'dram__bytes_read.sum', 'dram__bytes_write.sum', 'lts__t_bytes.sum', 'l1tex__t_sector_hit_rate.pct', 'sm__cycles_elapsed.max',
        'smsp__inst_executed_op_local_ld.sum', 'smsp__inst_executed_op_local_st.sum', 'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active']
print(f"{'profiled_launch_steps':70s} {steps:g} step")   # integrator steps of the profiled launch (bench.py scales per-launch figures by it)
for h, u, v in zip(hdr, units, vals):
    if h in keys: print(f"{h:70s} {v} {u}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
for n, row in enumerate(rows):
    if 'Source' in row and '# Samples' in row: hdr = row; start = n + 1; break
iS, iE = hdr.index('Source'), hdr.index('Instructions Executed')
cols = [(i, h) for i, h in enumerate(hdr) if h.startswith('stall_') and 'Not Issued' not in h]
st = collections.Counter(); ops = collections.Counter(); tot = 0
for row in rows[start:]:
    try: e = int(row[iE])
    except Exception: continue
    for i, h in cols:
        try: st[h] += int(row[i])
        except Exception: pass
    m = re.match(r'\s*(@!?U?P\d+\s+)?([A-Z0-9_.]+)', row[iS])
    if m: ops[m.group(2).split('.')[0]] += e; tot += e
ssum = sum(st.values())
print("stalls:", ", ".join(f"{h[6:]}={c/ssum:.3f}" for h, c in st.most_common(9)))
nw = float(vals[hdr.index('x')] if 'x' in hdr else 0) if False else None
print(f"warp-instr total {tot}; per step {tot/steps:.0f}")
print("ops per step:", ", ".join(f"{o}={c/steps:.0f}" for o, c in ops.most_common(22)))
