"""Summarise ncu outputs brought back in gpurun_out/ into the text files kept under profiles/.

    python profiles/summarize.py launches gpurun_out/launches_r1f.csv        > profiles/r1_launches_device_pass.txt
    python profiles/summarize.py kernel   gpurun_out/prof_scan_r1f.ncu-rep   > profiles/r1_scan_full.txt
"""
import collections
import csv
import subprocess
import sys


def launches(path):
    lines = [l for l in open(path) if not l.startswith("==")]
    agg = collections.OrderedDict()
    tot = 0.0
    n = 0
    for row in csv.DictReader(lines):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(row["Metric Value"].replace(",", ""))
        u = row["Metric Unit"]
        v = v / 1e3 if u == "ns" else v * 1e3 if u == "ms" else v * 1e6 if u == "s" else v
        k = row["Kernel Name"].split("(")[0][:80]
        a = agg.setdefault(k, [0, 0.0])
        a[0] += 1
        a[1] += v
        tot += v
        n += 1
    print("# %d launches, %.1f us of kernel time (ncu per-launch times are cold-cache and serialised: compare shares)" % (n, tot))
    print("%-82s %6s %12s %7s" % ("kernel", "n", "total_us", "share"))
    for k, (c, v) in sorted(agg.items(), key=lambda x: -x[1][1]):
        print("%-82s %6d %12.1f %6.1f%%" % (k, c, v, 100 * v / tot))


KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "launch__grid_size", "launch__block_size", "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic",
        "launch__waves_per_multiprocessor", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_tma.avg.pct_of_peak_sustained_active",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "lts__t_sector_hit_rate.pct"]


def kernel(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    h, units, data = rows[0], rows[1], rows[2:]
    for r in data:
        print("# kernel:", r[h.index("Kernel Name")][:100])
        for i, c in enumerate(h):
            if c in KEYS or c.startswith("smsp__average_warps_issue_stalled"):
                print("%-90s %-12s %s" % (c, units[i], r[i]))


if __name__ == "__main__":
    {"launches": launches, "kernel": kernel}[sys.argv[1]](sys.argv[2])
