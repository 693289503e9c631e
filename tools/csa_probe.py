#!/usr/bin/env python3
"""csa_access vs sa_access through the C ABI on the c2 index (4.6 Mbp, 9.2 M rows): wall time per call of 10 M random
rows (H2D of the rows and D2H of the positions included) and the check that both return the same positions.
GPU box only."""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import longreadmapper_b200 as lm

g = lm.sim_genome(20240, 4_600_000)
m = lm.GpuMapper.build([g], 0)
N = 2 * g.size + 1
rows = np.random.default_rng(3).integers(0, N, 10_000_000).astype(np.uint64)


def t(fn):
    fn()
    ts = []
    for _ in range(5):
        t0 = time.perf_counter()
        r = fn()
        ts.append((time.perf_counter() - t0) * 1e3)
    return min(ts), r


ms, sa = t(lambda: m.sa_access_batch(rows))
print(f"sa_access_batch            {ms:8.2f} ms  {rows.size / ms / 1e3:8.1f} M rows/s")
for ratio in (4, 32, 64):
    ms, pos = t(lambda: m.csa_access_batch(rows, ratio))
    print(f"csa_access_batch ratio {ratio:3d} {ms:8.2f} ms  {rows.size / ms / 1e3:8.1f} M rows/s  equal to SA: {np.array_equal(pos, sa)}")
