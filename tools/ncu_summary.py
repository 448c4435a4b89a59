"""Summarise `ncu -i X.ncu-rep --page raw --csv` (one row per profiled launch) into the table kept under profiles/.

usage: ncu -i gpurun_out/X.ncu-rep --page raw --csv > /tmp/raw.csv; python tools/ncu_summary.py /tmp/raw.csv
Columns: duration, DRAM traffic (dram__bytes_read.sum + dram__bytes_write.sum), DRAM / L1 / L2 / SM throughput as % of
peak, issue-slot utilisation, the busiest execution pipes (% of peak, instruction-issue based), achieved occupancy and
registers per thread."""
import csv
import sys


def fnum(s):
    try:
        return float(str(s).replace(",", ""))
    except ValueError:
        return float("nan")


def main(path):
    with open(path) as f:
        rows = list(csv.reader(l for l in f if not l.startswith("==")))
    hdr, units, data = rows[0], rows[1], rows[2:]
    col = {h: i for i, h in enumerate(hdr)}

    def get(r, name):
        return fnum(r[col[name]]) if name in col else float("nan")

    def unit(name):
        return units[col[name]] if name in col else ""

    pipes = [h for h in hdr if h.startswith("sm__inst_executed_pipe_") and h.endswith(".avg.pct_of_peak_sustained_active")]
    if not pipes:
        pipes = [h for h in hdr if "inst_executed_pipe_" in h and "pct_of_peak" in h]
    print("%-16s %9s %10s %6s %6s %6s %6s %6s %5s %4s  %s" % ("kernel", "us", "dramMB", "dram%", "l1%", "l2%", "sm%", "issue%", "occ%", "regs", "top pipes (% of peak)"))
    for r in data:
        name = r[col["Kernel Name"]].split("(")[0][-16:]
        t = get(r, "gpu__time_duration.sum")
        tu = unit("gpu__time_duration.sum")
        t_us = t / 1e3 if tu in ("ns", "nsecond") else t * 1e3 if tu in ("ms", "msecond") else t
        rb, wb = get(r, "dram__bytes_read.sum"), get(r, "dram__bytes_write.sum")
        sc = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
        mb = (rb * sc.get(unit("dram__bytes_read.sum"), 1) + wb * sc.get(unit("dram__bytes_write.sum"), 1)) / 1e6
        pp = sorted(((get(r, p), p.split("pipe_")[1].split(".")[0]) for p in pipes if get(r, p) == get(r, p)), reverse=True)[:4]
        print("%-16s %9.1f %10.2f %6.1f %6.1f %6.1f %6.1f %6.1f %5.1f %4d  %s" % (
            name, t_us, mb, get(r, "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
            get(r, "l1tex__throughput.avg.pct_of_peak_sustained_active"), get(r, "lts__throughput.avg.pct_of_peak_sustained_elapsed"),
            get(r, "sm__throughput.avg.pct_of_peak_sustained_elapsed"), get(r, "sm__inst_issued.avg.pct_of_peak_sustained_active"),
            get(r, "sm__warps_active.avg.pct_of_peak_sustained_active"), int(get(r, "launch__registers_per_thread")),
            " ".join("%s=%.0f" % (n, v) for v, n in pp)))


if __name__ == "__main__":
    main(sys.argv[1])
