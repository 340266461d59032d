"""Executed-instruction histogram by opcode (optionally within a SASS index range) of one launch."""
import csv, subprocess, sys, collections
rep, kid, px = sys.argv[1], sys.argv[2], float(sys.argv[3])
lo, hi = (map(int, sys.argv[4].split('-')) if len(sys.argv) > 4 else (0, 10**9))
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-id", ":::" + kid], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = rows[1]; data = rows[2:]
iS, iE = hdr.index('Source'), hdr.index('Instructions Executed')
c = collections.Counter()
for k, r in enumerate(data):
    if len(r) <= iE or not r[iE].isdigit(): continue
    if not (lo <= k < hi): continue
    t = r[iS].strip().split()
    op = t[1] if t[0].startswith('@') else t[0]
    c[op.rstrip(';')] += int(r[iE])
tot = sum(c.values())
print('total winst/px %.3f' % (tot / px))
for op, n in c.most_common(40):
    print('%-22s %.4f winst/px  %.1f%%' % (op, n / px, 100 * n / tot))
