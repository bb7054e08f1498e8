"""Group the SASS of an ncu report into runs with equal execution count (loops / branches).
python profiles/sass_groups.py rep.ncu-rep [min_total_millions] [dump_from dump_to]"""
import csv, subprocess, sys
rep = sys.argv[1]; thr = float(sys.argv[2]) if len(sys.argv) > 2 else 3.0
out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
his = [i for i, r in enumerate(rows) if r and r[0] == 'Address'] + [len(rows)]
import os
kidx = int(os.environ.get('NCU_KERNEL', '0'))  # which kernel of a multi-kernel report
hi = his[kidx]; rows = rows[:his[kidx + 1]]
h = rows[hi]; iS = h.index('Source'); iE = h.index('Instructions Executed')
body = [r for r in rows[hi + 1:] if len(r) > iE and r[iE] != '']
grp = []
for n, r in enumerate(body):
    c = int(r[iE])
    if grp and grp[-1][2] == c: grp[-1][1] = n
    else: grp.append([n, n, c])
for a, b, c in grp:
    tot = (b - a + 1) * c
    if tot > thr * 1e6: print('%5d-%5d n=%4d exec=%9d total=%7.1fM  %s' % (a, b, b - a + 1, c, tot / 1e6, body[a][iS][:50].strip()))
if len(sys.argv) > 4:
    for n in range(int(sys.argv[3]), int(sys.argv[4]) + 1): print(n, body[n][iE], body[n][iS][:90])
