"""Per-kernel summary of an `ncu --set full` report: python tools/ncu_summarize.py report.ncu-rep > profiles/x.txt"""
import csv
import subprocess
import sys

rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units, data = rows[0], rows[1], rows[2:]
col = {h: i for i, h in enumerate(hdr)}
want = [("gpu__time_duration.sum", "time"), ("dram__bytes_read.sum", "dram rd"), ("dram__bytes_write.sum", "dram wr"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram %pk"),
        ("l1tex__m_xbar2l1tex_read_bytes.sum", "L2->SM rd"),
        ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor %"),
        ("sm__inst_executed_pipe_tensor.sum", "tensor inst"),
        ("launch__registers_per_thread", "regs"), ("launch__grid_size", "grid"), ("launch__block_size", "block"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ %"),
        ("smsp__inst_executed.sum", "warp inst")]
want = [(k, n) for k, n in want if k in col]
print("# " + " | ".join(f"{n} [{units[col[k]]}]" for k, n in want))
for r in data:
    name = r[col["Kernel Name"]]
    name = name.split("(")[0].replace("b200::", "")[:60]
    vals = []
    for k, _ in want:
        v = r[col[k]]
        try:
            v = f"{float(v.replace(',', '')):.4g}"
        except ValueError:
            pass
        vals.append(v)
    print(f"{r[col['ID']]:>3} {name:60s} " + " | ".join(vals))
