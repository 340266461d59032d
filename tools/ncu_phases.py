"""Per-phase (between BAR.SYNC) instruction / stall-sample shares of one kernel launch in an ncu report."""
import csv, subprocess, sys
rep, kid, px = sys.argv[1], sys.argv[2], float(sys.argv[3])
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-id", ":::" + kid], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = rows[1]; data = rows[2:]
iS, iE, iSamp = hdr.index('Source'), hdr.index('Instructions Executed'), hdr.index('# Samples')
data = [r for r in data if len(r) > max(iE, iSamp) and r[iE].isdigit()]
tot = sum(int(r[iE]) for r in data); ts = sum(int(r[iSamp]) for r in data)
print('total warp inst', tot, 'samples', ts, 'winst/px %.3f' % (tot / px))
seg = acc = sacc = start = 0
for k, r in enumerate(data):
    acc += int(r[iE]); sacc += int(r[iSamp])
    if 'BAR.SYNC' in r[iS] or k == len(data) - 1:
        print('seg', seg, 'sass', start, '-', k, 'inst %.1f%%' % (100 * acc / tot), 'samples %.1f%%' % (100 * sacc / ts), 'winst/px %.3f' % (acc / px))
        seg += 1; acc = sacc = 0; start = k + 1
if len(sys.argv) > 4:
    lo, hi = map(int, sys.argv[4].split('-'))
    for k in range(lo, hi):
        r = data[k]; print(k, r[iE], r[iSamp], r[iS].strip()[:100])
