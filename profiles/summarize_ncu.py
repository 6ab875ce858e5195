"""ncu report -> the per-kernel summary rows committed under profiles/ (and, with --traffic, the DRAM bytes per launch
bench.py reads for roofline.traffic).  usage: python profiles/summarize_ncu.py REPORT.ncu-rep OUT.csv"""
import csv
import io
import subprocess
import sys

COLS = ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size",
        "lts__t_sector_hit_rate.pct", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "l1tex__t_sector_hit_rate.pct"]


def rows_of(report):
    raw = subprocess.run(["ncu", "-i", report, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    idx = [hdr.index(c) for c in COLS if c in hdr]
    out = [[hdr[i] for i in idx], [units[i] for i in idx]]
    for r in rows[2:]:
        out.append([r[i] for i in idx])
    return out


def to_bytes(value, unit):
    scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[unit]
    return float(value.replace(",", "")) * scale


if __name__ == "__main__":
    rows = rows_of(sys.argv[1])
    with open(sys.argv[2], "w", newline="") as f:
        csv.writer(f).writerows(rows)
    hdr, units = rows[0], rows[1]
    ri, wi, ti = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum"), hdr.index("gpu__time_duration.sum")
    for r in rows[2:]:
        name = r[0].split("(")[0].split("::")[-1]
        print("%-34s %9.1f us  dram %8.1f MB" % (name[:34], float(r[ti].replace(",", "")),
                                                 (to_bytes(r[ri], units[ri]) + to_bytes(r[wi], units[wi])) / 1e6))
