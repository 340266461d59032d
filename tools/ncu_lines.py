"""Per-CUDA-source-line warp-instruction and stall-sample totals of one launch in an ncu report (needs -lineinfo
and --import-source on).  usage: python tools/ncu_lines.py rep.ncu-rep <launch 1..n> <pixels> [min_share_pct]"""
import csv
import subprocess
import sys

rep, kid, px = sys.argv[1], sys.argv[2], float(sys.argv[3])
min_pct = float(sys.argv[4]) if len(sys.argv) > 4 else 0.5
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass,cuda", "--kernel-id", ":::" + kid],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
fname, hdr, per = None, None, {}
for r in rows:
    if len(r) == 2 and r[0] == "File Path":
        fname = r[1].split("/")[-1]
        continue
    if r and r[0] == "Line No":
        hdr = r
        iI, iS, iW, iWi = hdr.index("Instructions Executed"), hdr.index("# Samples"), hdr.index("L1 Wavefronts Shared"), hdr.index("L1 Wavefronts Shared Ideal")
        continue
    if hdr and len(r) > iWi and r[0].isdigit() and r[2] == "-":      # a CUDA source line (aggregated over its SASS)
        k = (fname, int(r[0]))
        a = per.setdefault(k, [0, 0, 0, 0, r[1].strip()])
        a[0] += int(r[iI] or 0); a[1] += int(r[iS] or 0); a[2] += int(r[iW] or 0); a[3] += int(r[iWi] or 0)
tot = sum(v[0] for v in per.values()); ts = sum(v[1] for v in per.values())
print("total warp inst %d  winst/px %.3f  samples %d" % (tot, tot / px, ts))
for (f, ln), v in sorted(per.items()):
    if v[0] * 100.0 / tot >= min_pct or v[1] * 100.0 / max(ts, 1) >= min_pct:
        print("%-16s %4d  inst %5.2f%% (%.3f/px)  samples %5.2f%%  smem wavefronts %d (ideal %d)  | %s"
              % (f, ln, 100.0 * v[0] / tot, v[0] / px, 100.0 * v[1] / max(ts, 1), v[2], v[3], v[4][:90]))
