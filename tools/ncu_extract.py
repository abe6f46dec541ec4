#!/usr/bin/env python3
"""Export an ncu report for profiles/: the FULL details page (untruncated) and a raw-counter extract (FP32 op counts
incl. packed, local memory, registers, DRAM bytes, issue).  usage: ncu_extract.py rep.ncu-rep out_prefix n_clips"""
import csv, json, subprocess, sys
rep, out, n = sys.argv[1], sys.argv[2], float(sys.argv[3])
txt = subprocess.run(["ncu", "-i", rep, "--page", "details"], capture_output=True, text=True).stdout
open(out + ".txt", "w").write(txt)
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
h, units, v = rows[0], rows[1], rows[-1]
m = dict(zip(h, v))
unit = dict(zip(h, units))
SCALE = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
def f(k):
    try:
        return float(m[k].replace(",", ""))
    except Exception:
        return None
keys = [k for k in h if k.startswith(("launch__registers", "launch__block_size", "launch__grid_size", "launch__shared_mem_per_block",
                                      "launch__occupancy_limit", "launch__waves")) or k in (
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "smsp__inst_executed.sum", "sm__cycles_elapsed.max",
    "smsp__issue_active.avg.pct", "smsp__warps_eligible.avg.per_cycle_active", "smsp__warps_active.avg.per_cycle_active",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
    "l1tex__t_requests_pipe_lsu_mem_local_op_ld.sum", "l1tex__t_requests_pipe_lsu_mem_local_op_st.sum",
    "smsp__sass_thread_inst_executed_op_ffma_pred_on.sum.per_cycle_elapsed", "smsp__sass_thread_inst_executed_op_fadd_pred_on.sum.per_cycle_elapsed",
    "smsp__sass_thread_inst_executed_op_fmul_pred_on.sum.per_cycle_elapsed", "smsp__sass_thread_inst_executed_op_dfma_pred_on.sum.per_cycle_elapsed",
    "smsp__sass_thread_inst_executed_op_dadd_pred_on.sum.per_cycle_elapsed", "smsp__sass_thread_inst_executed_op_dmul_pred_on.sum.per_cycle_elapsed",
    "lts__t_sectors_srcunit_tex_aperture_peer.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed")]
got = {k: "%s %s" % (m[k], unit.get(k, "")) for k in keys}
d = {"kernel": m.get("Kernel Name"), "clips": n, "counters": got}
rd, wr = f("dram__bytes_read.sum"), f("dram__bytes_write.sum")
if rd is not None and wr is not None:
    tot = rd * SCALE.get(unit["dram__bytes_read.sum"], 1) + wr * SCALE.get(unit["dram__bytes_write.sum"], 1)
    d["dram_bytes_per_launch"] = tot
    d["dram_bytes_per_clip"] = tot / n
ie = f("smsp__inst_executed.sum")
if ie is not None:
    d["warp_instructions_per_clip"] = ie / n
ld, st = f("l1tex__t_requests_pipe_lsu_mem_local_op_ld.sum"), f("l1tex__t_requests_pipe_lsu_mem_local_op_st.sum")
if ld is not None:
    d["local_memory_requests_per_clip"] = {"ld": ld / n, "st": (st or 0) / n}
fl = [f("smsp__sass_thread_inst_executed_op_%s_pred_on.sum.per_cycle_elapsed" % o) for o in ("ffma", "fadd", "fmul")]
if None not in fl:
    d["scalar_fp32_flop_per_cycle_counted_by_ncu_roofline"] = 2 * fl[0] + fl[1] + fl[2]
    d["note"] = ("ncu's FP32 roofline / op_ffma, op_fadd, op_fmul counters see only the SCALAR instructions; the packed "
                 "FFMA2 / FADD2 / FMUL2 (two fp32 operations per lane each; see the opcode mix files) are not counted")
json.dump(d, open(out + "_raw.json", "w"), indent=1)
print(out, "details %d lines," % txt.count("\n"), "dram bytes/clip %.0f, warp-instr/clip %.0f" % (d.get("dram_bytes_per_clip", 0), d.get("warp_instructions_per_clip", 0)))
