"""Summarise an .ncu-rep (ncu --set full) into a small text table for profiles/.
   python tools/ncu_summary.py gpurun_out/x.ncu-rep > profiles/x.txt"""
import csv
import io
import subprocess
import sys

WANT = [("Kernel Name", "kernel"), ("gpu__time_duration.sum", "us"), ("dram__bytes_read.sum", "dram_rd_MB"),
        ("dram__bytes_write.sum", "dram_wr_MB"), ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram_%"),
        ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor_%"),
        ("sm__inst_executed.avg.per_cycle_elapsed", "ipc"), ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ_%"),
        ("launch__registers_per_thread", "regs"), ("launch__grid_size", "grid")]

out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr, units = rows[0], rows[1]
idx = [(hdr.index(k), n) for k, n in WANT if k in hdr]
print("# " + sys.argv[1] + "   (ncu --set full --clock-control none; per launch, cold cache, serialised)")
print(" | ".join(n for _, n in idx))
for r in rows[2:]:
    cells = []
    for i, n in idx:
        v = r[i]
        if n == "kernel":
            v = v.replace("void ", "").replace("sfno::", "")[:60]
        else:
            try:
                v = f"{float(v):.2f}"
            except ValueError:
                pass
        cells.append(v)
    print(" | ".join(cells))
