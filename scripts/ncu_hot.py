"""Stall hot spots of one kernel launch: python scripts/ncu_hot.py report.ncu-rep kernel-regex [launch-skip] [min-percent]"""
import csv, subprocess, sys
rep, rx = sys.argv[1], sys.argv[2]
skip = sys.argv[3] if len(sys.argv) > 3 else '0'
thr = float(sys.argv[4]) if len(sys.argv) > 4 else 0.7
out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--kernel-name', 'regex:' + rx, '--launch-skip', skip, '--launch-count', '1',
                      '--print-source', 'sass'], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
idx = [i for i, r in enumerate(rows) if r and r[0] == 'Address']
start = idx[0]; end = idx[1] if len(idx) > 1 else len(rows)
hdr = rows[start]
iS, iSm = hdr.index('Source'), hdr.index('# Samples')
body = [r for r in rows[start + 1:end] if len(r) >= len(hdr)]
print(rows[start - 1][1][:120])
tot = sum(int(r[iSm] or 0) for r in body)
print('total samples', tot, 'instructions', len(body))
names = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
for k, r in enumerate(body):
    sm = int(r[iSm] or 0)
    if sm >= tot * thr / 100:
        st = ' '.join(f"{h[6:]}={r[hdr.index(h)]}" for h in names if r[hdr.index(h)] not in ('', '0') and int(r[hdr.index(h)]) * 10 >= sm)
        print(f'{k:4d} {r[iS][:64]:64s} {100 * sm / tot:5.1f}%  {st}')
