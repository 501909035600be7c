"""Summarise an .ncu-rep (first kernel): key throughput metrics, stall mix, hottest SASS lines, instruction mix.
python tools/ncu_summary.py gpurun_out/prof_x.ncu-rep [ntop]"""
import collections
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
ntop = int(sys.argv[2]) if len(sys.argv) > 2 else 25
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
r = list(csv.reader(io.StringIO(raw)))
h, u, v = r[0], r[1], r[2]
KEYS = ["Kernel Name", "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "launch__shared_mem_per_block_static", "launch__waves_per_multiprocessor",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_warps",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed.sum", "smsp__inst_executed.sum", "sm__cycles_elapsed.max", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "sm__pipe_tc_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct", "lts__t_bytes.sum", "lts__t_sectors.sum",
        "l1tex__m_xbar2l1tex_read_bytes.sum", "l1tex__m_l1tex2xbar_write_bytes.sum",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smsp__cycles_active.avg",
        "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum",
        "l1tex__t_requests_pipe_lsu_mem_global_op_st.sum", "l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum",
        "l1tex__t_requests_pipe_lsu_mem_global_op_red.sum", "l1tex__t_sectors_pipe_lsu_mem_global_op_red.sum",
        "lts__t_sectors_op_red.sum", "lts__t_sectors_op_atom.sum", "lts__t_sectors_op_read.sum", "lts__t_sectors_op_write.sum"]
for k in KEYS:
    if k in h:
        i = h.index(k)
        print("%-70s %-10s %s" % (k, u[i], v[i]))
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hh = rows[1]
rows = rows[2:]
iS, iSrc, iI = hh.index("# Samples"), hh.index("Source"), hh.index("Instructions Executed")
stalls = [i for i, k in enumerate(hh) if k.startswith("stall_") and "Not Issued" not in k]
tot = sum(int(x[iS] or 0) for x in rows)
agg = collections.Counter()
for x in rows:
    for i in stalls:
        agg[hh[i]] += int(x[i] or 0)
print("samples", tot, "stall mix:", [(k, n) for k, n in agg.most_common(8)])
print("warp instructions", sum(int(x[iI] or 0) for x in rows))
for x in sorted(rows, key=lambda x: -int(x[iS] or 0))[:ntop]:
    s = {hh[i]: int(x[i] or 0) for i in stalls}
    ss = sorted([(k, n) for k, n in s.items() if n], key=lambda kv: -kv[1])[:2]
    print("%6s %8s  %-64s %s" % (x[iS], x[iI], x[iSrc][:64], ss))
op = collections.Counter()
for x in rows:
    m = x[iSrc].split()
    if m:
        o = m[1] if m[0].startswith("@") and len(m) > 1 else m[0]
        op[o.split(".")[0]] += int(x[iI] or 0)
print("instruction mix:", op.most_common(14))
