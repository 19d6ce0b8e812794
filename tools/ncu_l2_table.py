"""L2 -> SM traffic table of the GEMM launches in a `ncu --set full` report (raw page): what the implicit-GEMM convolutions move
through the crossbar next to their DRAM bytes and tensor-pipe activity.   python tools/ncu_l2_table.py report.ncu-rep [kernel-substring]"""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
needle = sys.argv[2] if len(sys.argv) > 2 else "gemm_tc_kernel"
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = rows[0]
ki = hdr.index("Kernel Name")
cols = {"time_us": "gpu__time_duration.sum", "grid": "launch__grid_size", "l2_to_sm_MB": "l1tex__m_xbar2l1tex_read_bytes.sum",
        "l2_to_sm_TBps": "l1tex__m_xbar2l1tex_read_bytes.sum.per_second", "lts_%": "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "l2_hit_%": "lts__t_sector_hit_rate.pct", "dram_rd_MB": "dram__bytes_read.sum",
        "tensor_%": "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "warps_%": "sm__warps_active.avg.pct_of_peak_sustained_active", "regs": "launch__registers_per_thread",
        "smem_KB": "launch__shared_mem_per_block_dynamic"}
idx = {k: hdr.index(v) for k, v in cols.items()}
print("units as reported by ncu: " + ", ".join(f"{k} {rows[1][i]}" for k, i in idx.items() if rows[1][i]))
print()
print("| kernel | " + " | ".join(cols) + " |")
print("|---|" + "---|" * len(cols))
for r in rows[2:]:
    n = r[ki]
    if needle not in n:
        continue
    i = n.index(needle)
    j = n.find(">", i)
    short = n[i:j + 1] if j > 0 else needle
    vals = []
    for k in cols:
        v = r[idx[k]]
        try:
            v = f"{float(v):.4g}"
        except ValueError:
            pass
        vals.append(v)
    print(f"| `{short}` | " + " | ".join(vals) + " |")
