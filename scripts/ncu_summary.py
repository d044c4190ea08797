"""Summarise an .ncu-rep into a small table (per launch): duration, DRAM bytes, DRAM %, tensor %, occupancy, regs.

    python scripts/ncu_summary.py gpurun_out/prof.ncu-rep [more.ncu-rep ...]
"""
import csv
import io
import subprocess
import sys

WANT = [("gpu__time_duration.sum", "us"), ("dram__bytes_read.sum", "rd"), ("dram__bytes_write.sum", "wr"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram%"),
        ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor%"),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm%"),
        ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "l2%"),
        ("l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "l1%"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%"),
        ("launch__registers_per_thread", "regs"), ("launch__grid_size", "grid"), ("launch__block_size", "block")]

for path in sys.argv[1:]:
    raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    cols = [(hdr.index(k), short) for k, short in WANT if k in hdr]
    name_i = hdr.index("Kernel Name")
    print(f"# {path}")
    print("kernel".ljust(44), " ".join(f"{s}[{units[i]}]".rjust(14) for i, s in cols))
    for r in rows[2:]:
        print(r[name_i][:44].ljust(44), " ".join(r[i][:12].rjust(14) for i, _ in cols))
