"""Per-SASS-instruction view of `ncu --page source --print-source cuda,sass --csv`: warp-level instructions
executed and stall samples per address, grouped into runs of equal execution count (= basic-block regions:
stages, loops, phase B).  usage: ncu_sass_regions.py file.csv [min_share_percent]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
thr = float(sys.argv[2]) if len(sys.argv) > 2 else 0.4
seen, hdr = {}, None
for r in rows:
    if not r:
        continue
    if r[0] == 'Line No':
        hdr = r
        iex, ist = hdr.index('Instructions Executed'), hdr.index('# Samples')
        continue
    if hdr is None or r[0] != '' or len(r) < len(hdr) or not r[2].startswith('0x'):
        continue
    off = len(r) - len(hdr)
    try:
        seen[int(r[2], 16)] = (','.join(r[3:4 + off]).strip(), float(r[iex + off] or 0), float(r[ist + off] or 0))
    except ValueError:
        pass
base = min(seen)
segs = []
for a in sorted(seen):
    s, ex, st = seen[a]
    if segs and segs[-1][1] == ex:
        segs[-1][2] += 1; segs[-1][3] += st; segs[-1][5] = s
    else:
        segs.append([a - base, ex, 1, st, s, s])
tot = sum(s[1] * s[2] for s in segs)
tst = sum(s[3] for s in segs)
print('instructions %d, executed %.4e, stall samples %d' % (len(seen), tot, tst))
cum = 0
for s in segs:
    w = s[1] * s[2]
    cum += w
    if 100 * w / tot > thr:
        print('%6x n=%4d ex=%9d %5.2f%% cum %5.1f%% st %4.1f%% | %s ... %s' %
              (s[0], s[2], s[1], 100 * w / tot, 100 * cum / tot, 100 * s[3] / max(tst, 1), s[4][:36], s[5][:36]))
