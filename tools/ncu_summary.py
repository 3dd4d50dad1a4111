"""Summarise an .ncu-rep (first kernel): key details, stall breakdown, SASS opcode histogram.
usage: python tools/ncu_summary.py rep.ncu-rep warps_per_gate gates [steps]"""
import collections, csv, io, re, subprocess, sys
rep, wpg, gates = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
steps = int(sys.argv[4]) if len(sys.argv) > 4 else 636
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, vals = rows[0], rows[1], rows[2]
want = r"^(gpu__time_duration.sum|sm__cycles_elapsed.max|launch__registers_per_thread|launch__occupancy_limit.*|sm__warps_active.avg.pct_of_peak_sustained_active|smsp__issue_active.avg.per_cycle_active|sm__inst_executed_pipe_(alu|fma|lsu|adu|uniform).avg.pct_of_peak_sustained_active|sm__pipe_(alu|fma|fmaheavy|fmalite|fp64|shared)_cycles_active.avg.pct_of_peak_sustained_active|sm__inst_executed_pipe_(fmaheavy|fmalite|fp64|xu|cbu|tex).*pct_of_peak_sustained_active|smsp__inst_executed_pipe_(fmaheavy|fmalite|alu|fma).*|sm__inst_executed.avg.per_cycle_elapsed|smsp__warps_eligible.avg.per_cycle_active|smsp__issue_inst0.avg.pct_of_peak_sustained_active|smsp__average_warps_issue_stalled_.*_per_issue_active.ratio|dram__bytes_(read|write).sum|l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed|l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum|l1tex__data_pipe_lsu_wavefronts_mem_shared.sum|smsp__inst_executed.sum|lts__t_sector_hit_rate.pct|l1tex__t_sector_hit_rate.pct|sm__throughput.avg.pct_of_peak_sustained_elapsed|smsp__sass_inst_executed_op_local_(ld|st).sum|l1tex__lsu_writeback_active.*pct.*|l1tex__data_pipe.*pct.*)$"
for h, u, v in zip(hdr, units, vals):
    if re.match(want, h):
        print(f"{h:95s} {u:12s} {v}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
h = rows[1]
ix = {n: i for i, n in enumerate(h)}
body = [r for r in rows[2:] if len(r) > 10]
ops, samp, tot = collections.Counter(), collections.Counter(), 0
for r in body:
    t = r[ix["Source"]].split()
    if t[0].startswith("@"):
        t = t[1:]
    n = int(r[ix["Instructions Executed"]])
    ops[t[0]] += n
    tot += n
    samp[t[0]] += int(r[ix["# Samples"]])
warps = gates * wpg
print(f"\nSASS opcode histogram per warp per CMUX step ({gates} gates x {wpg} warps, {steps} steps)")
for op, n in ops.most_common(28):
    print(f"{op:24s} {n / warps / steps:9.1f}  {100 * n / tot:5.1f}% instr  {100 * samp[op] / max(1, sum(samp.values())):5.1f}% samples")
print(f"total {tot / warps / steps:.1f} instr per warp-step")
print("\nstall samples (all):")
for st in ["stall_wait", "stall_math", "stall_selected", "stall_not_selected", "stall_long_sb", "stall_dispatch", "stall_short_sb",
           "stall_no_inst", "stall_barrier", "stall_mio", "stall_lg", "stall_branch_resolving"]:
    if st in ix:
        print(f"{st:24s} {sum(int(r[ix[st]]) for r in body)}")
