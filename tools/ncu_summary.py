"""Text summaries of ncu output for profiles/: a launch list (csv of gpu__time_duration) and/or a full report.

    python tools/ncu_summary.py launches <launches.csv>
    python tools/ncu_summary.py report <file.ncu-rep>
"""
import collections
import csv
import io
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed.sum", "lts__t_sector_hit_rate.pct", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__shared_mem_per_block_dynamic"]


def launches(path):
    rows = [r for r in csv.reader(open(path)) if len(r) > 5]
    hdr = rows[0]
    ki, vi, gi, bi = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Grid Size"), hdr.index("Block Size")
    d = collections.OrderedDict()
    for r in rows[1:]:
        try:
            d.setdefault((r[ki], r[gi], r[bi]), []).append(float(r[vi].replace(",", "")))
        except ValueError:
            pass
    tot = sum(sum(v) for v in d.values())
    print("%-78s %7s %10s %8s" % ("kernel  grid x block", "count", "mean us", "share"))
    for (k, g, b), v in sorted(d.items(), key=lambda kv: -sum(kv[1])):
        name = k.replace("ovm::<unnamed>::", "").replace("void ", "")[:56]
        print("%-78s %7d %10.2f %7.1f%%" % ("%s  %s x %s" % (name, g, b), len(v), sum(v) / len(v) / 1e3, 100 * sum(v) / tot))


def report(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    ki = hdr.index("Kernel Name")
    for r in rows[2:]:
        print(r[ki].replace("ovm::<unnamed>::", "")[:110])
        for k in KEYS:
            if k in hdr:
                print("    %-72s %14s %s" % (k, r[hdr.index(k)], units[hdr.index(k)]))


if __name__ == "__main__":
    (launches if sys.argv[1] == "launches" else report)(sys.argv[2])
