"""Summarises an `ncu --page raw --csv` export (one row per captured launch) into a markdown table for profiles/.
usage: python scripts/ncu_summary.py raw.csv > summary.md"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[0]
col = {h: i for i, h in enumerate(hdr)}


def num(r, name, default=0.0):
    try:
        return float(r[col[name]].replace(",", ""))
    except (KeyError, ValueError):
        return default


ROLE = [("k_tc_gemm4<2", "gathered depth step"), ("k_tc_gemm4<0, 0", "input projection (3xTF32, plain operand)"),
        ("k_tc_gemm4<0, 1", "atom update (3xTF32, two K segments, bias+act)"), ("k_tc_gemm4", "forward contraction"),
        ("k_tc_gemm3<1", "atom update (two K segments, bias+act)"),
        ("k_tc_gemm3", "plain GEMM (dX; input projection in tf32 mode)"), ("k_tc_dw5", "weight gradient"),
        ("k_message_grad", "message gradient"), ("k_aggregate", "aggregate / transpose-CSR scatter"),
        ("k_readout_grad", "readout backward"), ("k_readout", "readout"), ("k_expand_rows", "feature expansion (compact upload)"),
        ("k_reduce_partials", "split-K reduction"), ("k_colsum_partial", "bias-gradient column sums"), ("k_tc_pack_b", "weight packing")]
print("| kernel | role | time us | DRAM read MB | DRAM write MB | DRAM total MB | GB/s | L2->SM read MB | LTS % | tensor pipe active % | regs | achieved occupancy % |")
print("|---|---|---|---|---|---|---|---|---|---|---|---|")
for r in rows[2:]:
    name = r[col["Kernel Name"]].replace("void ", "")
    short = name.split("(")[0]
    role = next((v for k, v in ROLE if short.startswith(k)), "")
    t = num(r, "gpu__time_duration.sum")
    def mbytes(name):
        unit = rows[1][col[name]] if name in col else "byte"
        return num(r, name) * {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}.get(unit, 1e-6)
    rd, wr = mbytes("dram__bytes_read.sum"), mbytes("dram__bytes_write.sum")
    tunit = rows[1][col["gpu__time_duration.sum"]]
    t_us = t * {"ns": 1e-3, "us": 1.0, "ms": 1e3, "usecond": 1.0, "nsecond": 1e-3, "msecond": 1e3}.get(tunit, 1.0)
    l2 = num(r, "lts__t_sectors_srcunit_tex_op_read.sum") * 32e-6
    lts = num(r, "lts__throughput.avg.pct_of_peak_sustained_elapsed")
    tens = num(r, "sm__inst_executed_pipe_tensor.avg.pct_of_peak_sustained_active", num(r, "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"))
    regs = num(r, "launch__registers_per_thread")
    occ = num(r, "sm__warps_active.avg.pct_of_peak_sustained_active")
    gbs = (rd + wr) / t_us * 1e3 if t_us else 0.0      # MB / us = TB/s
    print(f"| `{short}` | {role} | {t_us:.1f} | {rd:.0f} | {wr:.0f} | {rd + wr:.0f} | {gbs:.0f} | {l2:.0f} | {lts:.0f} | {tens:.0f} | {regs:.0f} | {occ:.0f} |")
