#!/usr/bin/env python3
"""Where does the end-to-end time of as_map_batch go?  Times the call on the c2 batch with and without the CIGAR-column
copy and over chunk sizes (AS_CHUNK_READS).  GPU box only."""
import ctypes as C
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import longreadmapper_b200 as lm

n = 1_000_000
g = lm.sim_genome(20240, 4_600_000)
m = lm.GpuMapper.build([g], 0)
reads, lens, _, _ = lm.sim_reads(777, g, n, 150, 150, 0.02, 0.02, 0.8)
stride = reads.shape[1]
L = m.L
p = m.params()
ss = 300
hp = [L.as_host_alloc(x) for x in (4 * n, 56 * n, ss * n)] + [L.as_host_alloc(n * stride) for _ in range(6)]
np.ctypeslib.as_array(C.cast(hp[0], C.POINTER(C.c_uint32)), shape=(n,))[:] = lens


def run(store, label):
    ts = []
    for k in range(6):
        np.ctypeslib.as_array(C.cast(hp[3 + k], C.POINTER(C.c_uint8)), shape=(n, stride))[:] = reads
    for k in range(6):
        t0 = time.perf_counter()
        rc = L.as_map_batch(m.h, C.byref(p), hp[3 + k], stride, hp[0], n, hp[1], hp[2] if store else None, ss)
        assert rc == 0
        ts.append((time.perf_counter() - t0) * 1e3)
    tm = m.last_timing()
    print(f"{label:40s} wall ms min {min(ts[1:]):7.3f} med {sorted(ts[1:])[2]:7.3f}  lib total {tm['total']:7.3f} h2d {tm['h2d']:6.3f} d2h {tm['d2h']:6.3f} "
          f"kern {tm['part1'] + tm['part2'] + tm['gact']:6.3f}  d2h bytes {m.last_transfer()[1] / 1e6:.0f} MB")


for cr in ("", "31250", "62500", "250000", "500000", "1000000"):
    if cr:
        os.environ["AS_CHUNK_READS"] = cr
    else:
        os.environ.pop("AS_CHUNK_READS", None)
    run(True, f"chunk={cr or 'auto'} with ops")
os.environ.pop("AS_CHUNK_READS", None)
run(False, "chunk=auto no ops")
