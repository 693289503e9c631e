#!/usr/bin/env python
"""CPU only: throughput of the ingest (ingest.c) and emit (emit.c) stages of accaln_b200 on a c2-shaped FASTQ, by thread count.
The results handed to the emitter are synthetic (every read 'mapped' forward with a three-run CIGAR)."""
import ctypes as C
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import longreadmapper_b200 as lm      # noqa: E402
from bench import write_fastq_fast      # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
threads = [int(x) for x in (sys.argv[2] if len(sys.argv) > 2 else "1,4,8").split(",")]
fq = "/tmp/stage_bench.fq"
if not os.path.exists(fq) or os.path.getsize(fq) != n * 315:
    g = lm.sim_genome(1, 4_600_000)
    reads, lens, _, _ = lm.sim_reads(2, g, min(n, 1_000_000), 150, 150, 0.02, 0.02, 0.8)
    write_fastq_fast(fq, reads, lens, max(1, n // reads.shape[0]))
H = lm.host_lib()


class Batch(C.Structure):
    _fields_ = [("n", C.c_size_t), ("max_len", C.c_uint32), ("wstride", C.c_size_t), ("packed", C.c_void_p), ("lens", C.c_void_p),
                ("name", C.c_void_p), ("seq", C.c_void_p), ("qual", C.c_void_p), ("name_len", C.c_void_p), ("text", C.c_void_p),
                ("text_len", C.c_size_t), ("text_owned", C.c_int), ("owner", C.c_void_p)]


H.as_ingest_open.restype = C.c_void_p
H.as_ingest_open.argtypes = [C.c_char_p, C.c_int, C.c_size_t]
H.as_ingest_next.argtypes = [C.c_void_p, C.POINTER(Batch), C.c_void_p, C.c_void_p]
H.as_packed_batch_free.argtypes = [C.POINTER(Batch), C.c_void_p]
H.as_ingest_close.argtypes = [C.c_void_p]
H.as_emitter_new.restype = C.c_void_p
H.as_emitter_new.argtypes = [C.c_int]
H.as_emit_batch.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(Batch), C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
ix = lm.capi.AsHostIndex()
names = (C.c_char_p * 1)(b"chr1")
mta = (lm.capi.AsMta * 1)(lm.capi.AsMta(0, 4_600_000))
ix.names = names
ix.mta = C.cast(mta, C.c_void_p)
ix.view.mta_len = 1
for T in threads:
    t_in = t_out = 0.0
    total = 0
    for rep in range(2):
        t_in = t_out = 0.0
        total = 0
        in_ = H.as_ingest_open(fq.encode(), T, 0)
        em = H.as_emitter_new(T)
        while True:
            b = Batch()
            t0 = time.perf_counter()
            rc = H.as_ingest_next(in_, C.byref(b), None, None)
            t_in += time.perf_counter() - t0
            assert rc == 0
            if b.n == 0:
                break
            rec = np.zeros(b.n, lm.COMPACT_DTYPE)
            rec["pos"] = np.arange(b.n, dtype=np.uint64) % 4_000_000 | (np.uint64(1) << np.uint64(40))
            rec["pos"][::2] |= np.uint64(16) << np.uint64(56)
            rec["n_runs"] = 3
            rec["n_runs"][::2] |= 0x80000000
            runs = np.tile(np.array([(63 << 2), (40 << 2) | 0, (1 << 2) | 1], np.uint8), b.n)
            sb = (C.c_size_t * 2)(0, b.n)
            rp = (C.c_void_p * 1)(runs.ctypes.data)
            rl = (C.c_uint64 * 1)(runs.size)
            seg = (C.c_void_p * 64)()
            sl = (C.c_size_t * 64)()
            t0 = time.perf_counter()
            ns = H.as_emit_batch(em, C.byref(ix), C.byref(b), rec.ctypes.data, 1, sb, rp, rl, seg, sl)
            t_out += time.perf_counter() - t0
            assert ns == T
            total += b.n
            H.as_packed_batch_free(C.byref(b), None)
        H.as_ingest_close(in_)
    print(f"{T:2d} threads: ingest {total / t_in / 1e6:6.2f} M reads/s ({t_in:.3f} s)   emit {total / t_out / 1e6:6.2f} M reads/s ({t_out:.3f} s)", flush=True)
