"""Summarise the SASS page of an ncu report: executed instructions and stall samples by
region (between barriers / calls) and by opcode.   python profiles/sass_hot.py rep.ncu-rep"""
import csv, subprocess, sys, collections
rep = sys.argv[1]
out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
his = [i for i, r in enumerate(rows) if r and r[0] == 'Address'] + [len(rows)]
import os
kidx = int(os.environ.get('NCU_KERNEL', '0'))  # which kernel of a multi-kernel report
hi = his[kidx]; rows = rows[:his[kidx + 1]]
h = rows[hi]
iS, iE, iW = h.index('Source'), h.index('Instructions Executed'), h.index('Warp Stall Sampling (All Samples)')
stalls = [c for c in h if c.startswith('stall_') and 'Not Issued' not in c]
body = [r for r in rows[hi + 1:] if len(r) > max(iW, iE) and r[iE] != '']
tot = sum(int(r[iE]) for r in body); tots = sum(int(r[iW]) for r in body)
print('instructions executed', tot, 'stall samples', tots)
ops = collections.Counter(); 
for r in body:
    op = r[iS].split()[0] if not r[iS].startswith('@') else r[iS].split()[1]
    ops[op.split('.')[0]] += int(r[iE])
print('by opcode:', ', '.join('%s %.1f%%' % (k, v * 100.0 / tot) for k, v in ops.most_common(22)))
st = collections.Counter()
for r in body:
    for c in stalls:
        st[c] += int(r[h.index(c)] or 0)
print('stalls:', ', '.join('%s %.1f%%' % (k, v * 100.0 / max(1, tots)) for k, v in st.most_common(10)))
# regions split at BAR / CALL / RET
reg = []; cur = [0, 0, 0, None]
for n, r in enumerate(body):
    s = r[iS]
    cur[0] += int(r[iE]); cur[1] += int(r[iW]); cur[2] += 1
    if cur[3] is None: cur[3] = n
    if 'BAR.' in s or s.startswith('CALL') or 'RET' in s or 'EXIT' in s:
        reg.append((cur[3], n, cur[0], cur[1], s[:40])); cur = [0, 0, 0, None]
for a, b, e, w, s in reg:
    if e > tot * 0.01 or w > tots * 0.01:
        print('sass %5d-%5d  inst %5.1f%%  stall %5.1f%%  ends %s' % (a, b, e * 100.0 / tot, w * 100.0 / max(1, tots), s))
