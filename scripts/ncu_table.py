"""One line per kernel of an .ncu-rep (raw page): time, grid, DRAM traffic and rate, L2 / tensor / SM utilisation.

    python scripts/ncu_table.py gpurun_out/evidence/step.ncu-rep > profiles/round2/ncu_step_table.md
"""
import csv
import re
import subprocess
import sys

rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units, data = rows[0], rows[1], rows[2:]


def col(d, name, default=""):
    return d[hdr.index(name)] if name in hdr else default


def num(d, name):
    try:
        return float(col(d, name, "0").replace(",", ""))
    except ValueError:
        return 0.0


def unit(name):
    return units[hdr.index(name)] if name in hdr else ""


def to_bytes(v, u):
    return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)


def to_us(v, u):
    return v * {"ns": 1e-3, "us": 1, "ms": 1e3, "s": 1e6, "nsecond": 1e-3, "usecond": 1, "msecond": 1e3}.get(u, 1)


print("| # | kernel | us | grid x block | regs | smem KB | DRAM rd+wr MB | DRAM TB/s | DRAM % | L2 % | tensor % | SM % | warps % |")
print("|---:|---|---:|---|---:|---:|---:|---:|---:|---:|---:|---:|---:|")
for i, d in enumerate(data):
    name = col(d, "Kernel Name")
    name = re.sub(r"\(.*$", "", name)
    name = name.replace("void ", "").replace("unnamed>::", "").replace("b200dp::", "")
    t = to_us(num(d, "gpu__time_duration.sum"), unit("gpu__time_duration.sum"))
    rd = to_bytes(num(d, "dram__bytes_read.sum"), unit("dram__bytes_read.sum"))
    wr = to_bytes(num(d, "dram__bytes_write.sum"), unit("dram__bytes_write.sum"))
    smem = to_bytes(num(d, "launch__shared_mem_per_block_dynamic"), unit("launch__shared_mem_per_block_dynamic")) / 1e3
    print(f"| {i} | `{name[:70]}` | {t:.1f} | {col(d, 'launch__grid_size')} x {col(d, 'launch__block_size')} | "
          f"{col(d, 'launch__registers_per_thread')} | {smem:.0f} | {(rd + wr) / 1e6:.1f} | {(rd + wr) / (t * 1e-6) / 1e12:.2f} | "
          f"{num(d, 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed'):.0f} | "
          f"{num(d, 'lts__throughput.avg.pct_of_peak_sustained_elapsed'):.0f} | "
          f"{num(d, 'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active'):.0f} | "
          f"{num(d, 'sm__throughput.avg.pct_of_peak_sustained_elapsed'):.0f} | "
          f"{num(d, 'sm__warps_active.avg.pct_of_peak_sustained_active'):.0f} |")
