"""Builds profiles/traffic.json (what bench.py's roofline.traffic is read from) out of `ncu --set full` reports.

    python tools/ncu_traffic.py <commit> <bench-name>@<kernel-substring>=<units>:<report.ncu-rep> ... > profiles/traffic.json

<bench-name> is the kernel name bench.py asks for.  For each argument the FIRST launch in the report whose ncu
kernel name contains the substring is taken:
dram_bytes = dram__bytes_read.sum + dram__bytes_write.sum of that launch, `units` = what that launch processed
(records / rows), so that bench.py only quotes the figure for a run of the same size.
"""
import csv
import json
import subprocess
import sys

SCALE = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def raw(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    return rows[0], rows[1], rows[2:]


def main():
    commit = sys.argv[1]
    entries = []
    for arg in sys.argv[2:]:
        key, rest = arg.split("=", 1)
        name, key = key.split("@", 1)
        units, rep = rest.split(":", 1)
        hdr, unit, data = raw(rep)
        ik = hdr.index("Kernel Name")
        ir, iw = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
        it = hdr.index("gpu__time_duration.sum")
        for r in data:
            if key in r[ik]:
                rd = float(r[ir]) * SCALE[unit[ir]]
                wr = float(r[iw]) * SCALE[unit[iw]]
                entries.append({"kernel": name, "kernel_full_name": r[ik], "units": int(units), "dram_bytes": rd + wr,
                                "dram_bytes_read": rd, "dram_bytes_write": wr,
                                "ncu_duration_us": float(r[it]), "ncu_duration_unit": unit[it],
                                "source": "profiles/ <- " + rep.split("/")[-1] + " (ncu --set full --clock-control none)"})
                break
        else:
            print("no launch of", key, "in", rep, file=sys.stderr)
    json.dump({"commit": commit, "note": "per-launch DRAM traffic from ncu; cold-cache, serialised replays",
               "entries": entries}, sys.stdout, indent=1)
    print()


if __name__ == "__main__":
    main()
