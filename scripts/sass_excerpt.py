#!/usr/bin/env python
"""SASS / ptxas evidence for the hot kernels (runs without a GPU): per kernel the register / spill line of
`nvcc -Xptxas -v` and the counts of the mnemonics that identify how it moves data (UBLKCP = 1-D bulk copy
on the TMA engine, SYNCS = mbarrier, CREDUX = warp float min/max, LDG.E.NA.128 = streaming 128-bit loads,
CCTL/prefetch, REDG/ATOMG = global reductions, DFMA = fp64 chain, VOTE/POPC = warp compaction, STL/LDL =
local-memory traffic).  usage: python scripts/sass_excerpt.py > profiles/r02_sass_excerpt.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, 'lidar_object_detection_b200', 'csrc')
HOT = [('fuse.cu', r'lf_fuse_kernelILi8ELi5ELb0ELi4ELb1', 'lf_fuse_kernel<8 warps, 5 loads, no car_bits, row-map build> (point pass)'),
       ('fuse.cu', r'lf_car_stats_kernel', 'lf_car_stats_kernel (per-car decision)'),
       ('erode.cu', r'lf_erode_fast_kernelINS_6SeSpecILi3ELi0ELi1ELi0ELi0EEELi0EE', 'lf_erode_fast_kernel<SeCross3, uint8> (erode + pack, dominant)'),
       ('erode.cu', r'lf_erode_packed_kernelINS_6SeSpecILi3ELi0ELi1ELi0ELi0EEEEE', 'lf_erode_packed_kernel<SeCross3> (masks arrive bit-packed)'),
       ('compact.cu', r'lf_car_compact_kernelIfE', 'lf_car_compact_kernel<float> (order-preserving per-car compaction)'),
       ('project.cu', r'lf_zscatter_batch_kernelILi1E', 'lf_zscatter_batch_kernel<last> (depth maps, scatter)'),
       ('resize.cu', r'lf_resize_linear_sep_kernel', 'lf_resize_linear_sep_kernel (separable bilinear resize: PRMT + IDP.2A per pixel)'),
       ('points.cu', r'lf_count_in_boxes_kernelIfE', 'lf_count_in_boxes_kernel<float> (128 points per warp, box cull by lane)'),
       ('points.cu', r'lf_mask_gather_lean_kernel', 'lf_mask_gather_lean_kernel (funnel-shift bit collection)')]
KEYS = ['UBLKCP', 'SYNCS', 'CREDUX', 'LDG.E.NA.128', 'LDG.E.128', 'LDG.E', 'CCTL', 'STG', 'REDG', 'ATOMG', 'ATOM.E', 'ATOMS', 'RED',
        'DFMA', 'DMUL', 'DADD', 'FFMA', 'MUFU', 'VOTE', 'POPC', 'SHFL', 'MATCH', 'IDP', 'PRMT', 'IMAD.HI', 'SHF', 'LDS', 'STS', 'LDL', 'STL',
        'BAR', 'UTMALDG', 'UTCHMMA']


def main():
    objs = {}
    for src in sorted({h[0] for h in HOT}):
        obj = '/tmp/sass_%s.o' % src[:-3]
        cmd = ['nvcc', '-O3', '-std=c++17', '-lineinfo', '-gencode', 'arch=compute_100a,code=sm_100a', '-Xcompiler',
               '-fPIC,-O3', '-I', os.path.join(ROOT, 'include'), '-I', CSRC, '-DLF_BUILDING=1', '-Xptxas', '-v', '-c',
               os.path.join(CSRC, src), '-o', obj]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode:
            sys.exit(r.stderr)
        objs[src] = (obj, r.stderr)
    print('# nvcc %s, -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo' %
          subprocess.run(['nvcc', '--version'], capture_output=True, text=True).stdout.strip().splitlines()[-2])
    for src, pat, title in HOT:
        obj, log = objs[src]
        lines = log.splitlines()
        print('\n== %s   [%s]' % (title, src))
        for i, l in enumerate(lines):
            if 'Compiling entry function' in l and re.search(pat, l):
                print('   ' + lines[i + 2].strip())
                print('   ' + lines[i + 3].strip())
                break
        sass = subprocess.run(['cuobjdump', '-sass', obj], capture_output=True, text=True).stdout
        body = None
        for f in re.split(r'\n\s*Function : ', sass)[1:]:
            if re.search(pat, f.split('\n')[0]):
                body = f
                break
        if body is None:
            print('   (not found in SASS)')
            continue
        ins = [m.group(1) for m in re.finditer(r'/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)', body)]
        cnt = collections.Counter()
        for op in ins:
            for k in KEYS:
                if op == k or op.startswith(k + '.') or op.startswith(k):
                    cnt[k] += 1
                    break
        print('   %d SASS instructions; ' % len(ins) + ', '.join('%s %d' % (k, cnt[k]) for k in KEYS if cnt[k]))
        shown = 0
        for m in re.finditer(r'/\*([0-9a-f]{4})\*/\s+(.*?);', body):
            if re.search(r'UBLKCP|SYNCS|CREDUX|LDG\.E\.NA\.128|CCTL|REDG|ATOM|IDP|MATCH', m.group(2)) and shown < 6:
                print('      /*%s*/ %s' % (m.group(1), m.group(2).strip()[:100]))
                shown += 1


if __name__ == '__main__':
    main()
