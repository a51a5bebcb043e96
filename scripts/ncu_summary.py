"""Summarise one `ncu --set full` report (first kernel): duration, DRAM traffic, issue/pipe utilisation, stalls,
and the instruction-mix evidence for tcgen05 / TMA.  usage: python scripts/ncu_summary.py <report.ncu-rep>"""
import csv, subprocess, sys, io, re
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
d = dict(zip(rows[0], rows[2]))
units = dict(zip(rows[0], rows[1]))
keys = ["Kernel Name", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "sm__inst_executed_pipe_tensor.sum", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_uniform.sum", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "smsp__inst_executed_op_shared_ld.sum", "smsp__inst_executed_op_shared_st.sum"]
for k in keys:
    if k in d: print(f"{k:<72} {d[k]} {units.get(k, '')}")
print("stall reasons (warps per issue-active cycle):")
st = [(k, float(v)) for k, v in d.items() if "issue_stalled" in k and k.endswith("per_issue_active.ratio")]
for k, v in sorted(st, key=lambda kv: -kv[1]):
    if v > 0.05: print(f"  {k.split('issue_stalled_')[1].split('_per_issue')[0]:<24} {v:.2f}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
srows = list(csv.reader(io.StringIO(src)))
h = srows[1]; si, ei = h.index("Source"), h.index("Instructions Executed")
mix = {}
for r in srows[2:]:
    if len(r) <= ei: continue
    op = r[si].split()[0] if r[si].split() else ""
    if op.startswith("@"): op = r[si].split()[1]
    op = op.split(".")[0]
    try: mix[op] = mix.get(op, 0) + float(r[ei])
    except ValueError: pass
tot = sum(mix.values())
print(f"SASS instruction mix (warp instructions executed, total {tot:.0f}):")
for op, n in sorted(mix.items(), key=lambda kv: -kv[1])[:22]:
    print(f"  {op:<12} {n:12.0f} {100 * n / tot:5.1f} %")
for op in ("UTCHMMA", "UTCBAR", "UTMALDG", "UTMASTG", "UBLKCP", "LDTM", "SYNCS"):
    if op in mix: print(f"  [{op}] {mix[op]:.0f}")
