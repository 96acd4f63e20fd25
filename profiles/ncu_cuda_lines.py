"""Summarise `ncu --page source --print-source cuda,sass --csv`: instructions executed and stall samples per CUDA source line."""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
thr = float(sys.argv[2]) if len(sys.argv) > 2 else 0.5
agg = collections.OrderedDict()
cur = hdr = None
for r in rows:
    if not r:
        continue
    if r[0] == 'File Path':
        cur = r[1].split('/')[-1]; continue
    if r[0] == 'Function Name':
        continue
    if r[0] == 'Line No':
        hdr = r; iex = hdr.index('Instructions Executed'); ist = hdr.index('# Samples'); continue
    if hdr is None or not r[0]:
        continue
    off = len(r) - len(hdr)                      # commas inside the source text shift the columns
    try:
        ex = float(r[iex + off] or 0); st = float(r[ist + off] or 0)
    except ValueError:
        continue
    key = (cur, int(r[0]), ','.join(r[1:2 + off]).strip()[:96])
    if key in agg:
        agg[key][0] += ex; agg[key][1] += st
    else:
        agg[key] = [ex, st]
tex = sum(v[0] for v in agg.values()); tst = sum(v[1] for v in agg.values())
print('total inst executed %.3e, stall samples %d' % (tex, tst))
for k, v in agg.items():
    if 100 * v[0] / tex > thr or 100 * v[1] / tst > thr:
        print('%-15s %4d %5.1f%% ex %5.1f%% st  %s' % (k[0], k[1], 100 * v[0] / tex, 100 * v[1] / tst, k[2]))
