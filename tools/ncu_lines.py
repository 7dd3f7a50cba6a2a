"""Per-source-line instruction / stall-sample totals of one kernel from an ncu report.
   python tools/ncu_lines.py <report.ncu-rep> <cubin> <mangled kernel substring> [top N]
The cubin comes from `cuobjdump -xelf all libswgpu.so`; line info needs -lineinfo (build.py passes it)."""
import collections, csv, re, subprocess, sys

rep, cubin, kname = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
sass = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout.splitlines()
line_of = {}
infn, cur = False, None
for l in sass:
    if l.startswith("//---") and ".text." in l:
        infn = kname in l
        cur = None
    if not infn:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/", l)
    if m:
        line_of[int(m.group(1), 16)] = cur
import os
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"] + os.environ.get("NCU_ARGS", "").split(), capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
# a report with several kernels prints one section per kernel: keep the first
for j in range(2, len(rows)):
    if rows[j] and rows[j][0] == "Kernel Name":
        rows = rows[:j]
        break
hdr = rows[1]
ia, iw, it, iss = hdr.index("Address"), hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed"), hdr.index("Warp Stall Sampling (All Samples)")
data = rows[2:]
base = int(data[0][ia], 16)
agg = collections.defaultdict(lambda: [0, 0, 0])
for r in data:
    k = line_of.get(int(r[ia], 16) - base)
    a = agg[k]
    a[0] += int(r[iw]); a[1] += int(r[it]); a[2] += int(r[iss])
tw = sum(a[0] for a in agg.values()); ts = sum(a[2] for a in agg.values())
print("total warp inst %.1fM, samples %d" % (tw / 1e6, ts))
src = {}
for k, a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    text = ""
    if k:
        if k[0] not in src:
            try:
                src[k[0]] = open("sitewhere_b200/csrc/" + k[0]).read().splitlines()
            except OSError:
                src[k[0]] = []
        if 0 < k[1] <= len(src[k[0]]):
            text = src[k[0]][k[1] - 1].strip()[:70]
    print("%-22s %6.2fM %4.1f%%  thr %4.1f  smp %4.1f%% | %s" % ("%s:%d" % k if k else "?", a[0] / 1e6, 100 * a[0] / tw, a[1] / max(a[0], 1), 100 * a[2] / max(ts, 1), text))
