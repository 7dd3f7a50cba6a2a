"""Turns an .ncu-rep (--set full) and a launch-list csv (--metrics gpu__time_duration.sum) into the
markdown summary kept under profiles/.  Usage:
    python tools/ncu_summary.py <report.ncu-rep> <launches.csv> "<command line that produced them>" > profiles/NAME.md
"""
import collections
import csv
import subprocess
import sys

METRICS = [
    "Block Size", "Grid Size", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "launch__registers_per_thread", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__t_sector_hit_rate.pct", "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "smsp__thread_inst_executed_per_inst_executed.ratio",
    "smsp__pcsamp_warps_issue_stalled_long_scoreboard", "smsp__pcsamp_warps_issue_stalled_wait",
    "smsp__pcsamp_warps_issue_stalled_barrier", "smsp__pcsamp_warps_issue_stalled_branch_resolving",
    "smsp__pcsamp_warps_issue_stalled_lg_throttle", "smsp__pcsamp_warps_issue_stalled_short_scoreboard",
]


def main():
    rep, launches, cmd = sys.argv[1], sys.argv[2], sys.argv[3]
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units, data = rows[0], rows[1], rows[2:]
    ik = hdr.index("Kernel Name")
    names = [r[ik].split("(")[0].replace("void ", "") for r in data]
    print("# ncu summaries (`--set full`, one launch of each kernel)\n")
    print("Command (run on a B200 under gpurun, after the same command had exited 0 without ncu):\n")
    print("    " + cmd + "\n")
    print("ncu replays each kernel cold-cache and serialised: use SHARES, not absolutes; the bench's CUDA-event "
          "timings are the numbers of record.\n")
    print("| metric | unit | " + " | ".join(names) + " |")
    print("|---|---|" + "---|" * len(names))
    for m in METRICS:
        if m not in hdr:
            continue
        j = hdr.index(m)
        print("| %s | %s | %s |" % (m, units[j], " | ".join(r[j] for r in data)))
    if launches == "-":
        return
    print("\n## Launch list of the timed steps (`ncu --metrics gpu__time_duration.sum`), microseconds\n")
    rows = list(csv.reader(open(launches)))
    for i, r in enumerate(rows):
        if "Kernel Name" in r:
            h, d = r, rows[i + 1:]
            break
    ik, iv, iu = h.index("Kernel Name"), h.index("Metric Value"), h.index("Metric Unit")
    agg = collections.OrderedDict()
    for r in d:
        if len(r) > iv:
            v = float(r[iv].replace(",", ""))
            v = v / 1000.0 if r[iu] in ("ns", "nsecond") else v
            agg.setdefault(r[ik].split("(")[0].replace("void ", "")[:70], []).append(v)
    print("| kernel | launches seen | mean us |\n|---|---|---|")
    for k, v in agg.items():
        print("| %s | %d | %.1f |" % (k, len(v), sum(v) / len(v)))


if __name__ == "__main__":
    main()
