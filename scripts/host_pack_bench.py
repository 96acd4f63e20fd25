"""Host mask packer throughput on this machine: pageable vs pinned source, every SIMD level, thread counts.
Usage: python scripts/host_pack_bench.py [n_masks]"""
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lidar_object_detection_b200 import _abi, io as IO

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
H, W = 376, 1408
lib = _abi.load()
top = lib.lf_host_simd(2)
srcs = {'pageable': torch.zeros((n, H, W), dtype=torch.uint8)}
if torch.cuda.is_available():
    srcs['pinned'] = torch.zeros((n, H, W), dtype=torch.uint8, pin_memory=True)
for t in srcs.values():
    t[:, 100:200, 300:600] = 1
out = {}
for kind, src in srcs.items():
    dst = torch.empty((n, H, (W + 31) // 32), dtype=torch.int32, pin_memory=(kind == 'pinned'))
    for cap in range(top, 0, -1):
        lib.lf_host_simd(cap)
        for threads in (8, 12, 16, os.cpu_count()):
            best = 1e9
            for _ in range(4):
                t0 = time.perf_counter()
                IO.pack_masks(src.numpy(), threads=threads, out=dst.numpy())
                best = min(best, time.perf_counter() - t0)
            out['%s simd%d t%d' % (kind, cap, threads)] = round(src.numel() / best / 1e9, 1)
lib.lf_host_simd(2)
print(json.dumps(out, indent=1))
