"""Summarise `ncu --page source --csv` output: instructions executed and stall samples per SASS line."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
h = next(i for i, r in enumerate(rows) if r and r[0] == 'Address')
hdr = rows[h]
ia, isrc, ist, iex = hdr.index('Address'), hdr.index('Source'), hdr.index('# Samples'), hdr.index('Instructions Executed')
data = [r for r in rows[h + 1:] if len(r) > iex]
tot_ex = sum(float(r[iex] or 0) for r in data); tot_s = sum(float(r[ist] or 0) for r in data)
print('total inst executed %.3e, samples %d, sass lines %d' % (tot_ex, tot_s, len(data)))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
if len(sys.argv) > 3 and sys.argv[3] == 'all':
    for n, r in enumerate(data):
        print('%4d %5.1f%% ex %5.1f%% st  %s' % (n, 100 * float(r[iex] or 0) / tot_ex, 100 * float(r[ist] or 0) / max(tot_s, 1), r[isrc][:110]))
else:
    for r in sorted(data, key=lambda r: -float(r[ist] or 0))[:top]:
        print('%5.1f%% ex %5.1f%% st  %s' % (100 * float(r[iex] or 0) / tot_ex, 100 * float(r[ist] or 0) / max(tot_s, 1), r[isrc][:110]))
