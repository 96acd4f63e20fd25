#!/usr/bin/env python
"""Text summary of ncu reports for profiles/: per kernel the headline metrics of the raw page, the executed
instructions and stall samples per CUDA source line, and the SASS regions (runs of equal execution count).
usage: python scripts/ncu_summary.py title report.ncu-rep[:kernel-regex] ... > profiles/<round>_ncu_summary.txt"""
import csv
import io
import re
import subprocess
import sys

METRICS = ['dram__bytes_read.sum', 'dram__bytes_write.sum', 'gpu__time_duration.sum',
           'l1tex__throughput.avg.pct_of_peak_sustained_elapsed', 'launch__block_size', 'launch__grid_size',
           'launch__occupancy_limit_registers', 'launch__occupancy_limit_shared_mem', 'launch__registers_per_thread',
           'lts__t_sector_hit_rate.pct', 'lts__throughput.avg.pct_of_peak_sustained_elapsed',
           'sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active', 'sm__warps_active.avg.pct_of_peak_sustained_active',
           'smsp__inst_executed.sum', 'smsp__issue_active.avg.pct_of_peak_sustained_active']


def page(rep, which, kernel=None):
    cmd = ['ncu', '-i', rep, '--page', which, '--csv']
    if kernel:
        cmd += ['-k', 'regex:' + kernel]
    return list(csv.reader(io.StringIO(subprocess.run(cmd, capture_output=True, text=True).stdout)))


def raw_section(rep, kernel):
    rows = page(rep, 'raw', kernel)
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    seen = set()
    for r in rows[2:]:
        name = r[idx['Kernel Name']]
        if name in seen:
            continue
        seen.add(name)
        print('== %s' % name[:120])
        for m in METRICS:
            if m in idx:
                print('%s %s %s' % (m, units[idx[m]], r[idx[m]]))
        for h in hdr:
            if 'issue_stalled' in h and h.endswith('per_issue_active.ratio'):
                try:
                    if float(r[idx[h]]) >= 0.3:
                        print('%s %s %s' % (h, units[idx[h]], r[idx[h]]))
                except ValueError:
                    pass


def source_section(rep, kernel):
    rows = page(rep, 'source', kernel)
    heads = [i for i, r in enumerate(rows) if r and r[0] == 'Address']
    if not heads:
        return
    h = rows[heads[0]]
    ia, isrc, iex, ist = h.index('Address'), h.index('Source'), h.index('Instructions Executed'), h.index('Warp Stall Sampling (All Samples)')
    end = heads[1] - 1 if len(heads) > 1 else len(rows)
    body = [r for r in rows[heads[0] + 1:end] if len(r) > iex and r[iex] != '']
    tot = sum(float(r[iex]) for r in body)
    tst = sum(float(r[ist] or 0) for r in body) or 1.0
    print('-- %s: SASS regions (runs of equal execution count) with >= 1 %% of the executed instructions or >= 2 %% of the stall samples' % rows[0][1][:80])
    print('instructions %d, executed %.4e, stall samples %d' % (len(body), tot, tst))
    cur, regs = None, []
    for r in body:
        ex = float(r[iex])
        if cur and cur['ex'] == ex:
            cur['n'] += 1; cur['st'] += float(r[ist] or 0); cur['last'] = r[isrc]
        else:
            cur = {'a': r[ia], 'ex': ex, 'n': 1, 'st': float(r[ist] or 0), 'first': r[isrc], 'last': r[isrc]}
            regs.append(cur)
    for c in regs:
        share = c['ex'] * c['n'] / tot * 100
        if share >= 1.0 or c['st'] / tst * 100 >= 2.0:
            print('%6s n=%4d ex=%11d %5.2f%% ex %5.2f%% st | %s ... %s' % (c['a'][-5:], c['n'], c['ex'], share, c['st'] / tst * 100,
                                                                       c['first'].strip()[:44], c['last'].strip()[:44]))
    print('-- the ten instructions with the most stall samples')
    cols = {c: h.index(c) for c in h if c.startswith('stall_') and 'Not Issued' not in c}
    for r in sorted(body, key=lambda r: -float(r[ist] or 0))[:10]:
        big = sorted(((float(r[i] or 0), c) for c, i in cols.items()), reverse=True)[:2]
        print('%6s %5.2f%% st  ex %10s  %-58s %s' % (r[ia][-5:], float(r[ist] or 0) / tst * 100, r[iex], r[isrc].strip()[:58],
                                                   ' '.join('%s=%d' % (c[6:], v) for v, c in big if v > 0)))


def main():
    print('# ' + sys.argv[1])
    for spec in sys.argv[2:]:
        rep, _, kernel = spec.partition(':')
        raw_section(rep, kernel or None)
        if kernel:
            source_section(rep, kernel)


if __name__ == '__main__':
    main()
