#!/usr/bin/env python
"""GPU box only: as_map_batch_packed on the c2 batch (or --workload) under several chunk plans; prints the wall time per call, the
device-side stage times and the host-side split (enqueue / wait).  A/B tool, not a bench."""
import argparse
import ctypes as C
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import longreadmapper_b200 as lm      # noqa: E402
from bench import WORKLOADS, GENOME_SEED, READS_SEED      # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--workload", default="c2")
ap.add_argument("--reads", type=int, default=0)
ap.add_argument("--chunks", default="0,125000,250000,333334,500000,1000000")
ap.add_argument("--reps", type=int, default=8)
ap.add_argument("--mids", default="")
ap.add_argument("--knobs", default="", help="as_params knob sets: a=1,b=2;a=3 ...")
a = ap.parse_args()
wl = WORKLOADS[a.workload]
n = a.reads or wl["reads"]
g = lm.sim_genome(GENOME_SEED, wl["genome"])
m = lm.GpuMapper.build([g], 0)
reads, lens, _, _ = lm.sim_reads(READS_SEED, g, n, wl["len"][0], wl["len"][1], wl["err"][0], wl["err"][1], wl["sub"])
wst = (int(lens.max()) + 15) // 16
packed = lm.pack_reads(reads, lens, wst)
L = m.L
cap = n * 64 if int(lens.max()) <= 512 else int(lens.astype(np.int64).sum())
bufs = [L.as_host_alloc(x) for x in (n * wst * 4, n * 4, n * 16, cap)]
np.ctypeslib.as_array(C.cast(bufs[0], C.POINTER(C.c_uint32)), shape=(n, wst))[:] = packed
np.ctypeslib.as_array(C.cast(bufs[1], C.POINTER(C.c_uint32)), shape=(n,))[:] = lens
got = C.c_uint64(0)
plans = [dict(chunk_reads=int(x)) for x in a.chunks.split(",") if x] + [dict(chunk_tail=int(x)) for x in a.mids.split(",") if x]
plans += [dict((kv.split("=")[0], int(kv.split("=")[1])) for kv in ks.split(",")) for ks in a.knobs.split(";") if ks]
for kn in plans:
    p = m.params(**kn)
    cr = kn.get("chunk_reads", -kn.get("chunk_tail", 0))
    if any(k not in ("chunk_reads", "chunk_tail") for k in kn):
        print("knobs", kn, flush=True)
    ts = []
    for k in range(a.reps):
        t0 = time.perf_counter()
        rc = L.as_map_batch_packed(m.h, C.byref(p), bufs[0], wst, bufs[1], n, bufs[2], bufs[3], cap, C.byref(got))
        ts.append((time.perf_counter() - t0) * 1e3)
        assert rc == 0, L.as_last_error()
    tm = m.last_timing()
    print(f"chunk_reads {cr:8d}: wall ms min {min(ts[2:]):.3f} med {float(np.median(ts[2:])):.3f} | device "
          + " ".join(f"{k} {v:.2f}" for k, v in tm.items() if k != "unused")
          + " | host " + " ".join(f"{k} {v:.0f}" for k, v in m.last_host_timing().items()), flush=True)
    ms = (C.c_float * 8)()
    nc = L.as_gpu_last_chunk_timeline(m.h, 0, C.byref(ms))
    for c in range(nc):
        L.as_gpu_last_chunk_timeline(m.h, c, C.byref(ms))
        print("      chunk", c, "h2d %.2f | begin %.2f vote %.2f locus %.2f gact %.2f myers %.2f all %.2f | d2h-fixed %.2f"
              % (ms[5], ms[0], ms[1], ms[2], ms[3], ms[4], ms[6], ms[7]), flush=True)
