"""as_edit_distance_device_noalloc: the edit-distance kernel with its scratch in a static device array, for callers that
cannot allocate (no counterpart in the reference).  Called through the C ABI
on device buffers and compared with as_edit_distance_batch on the same pairs -- integers, so bit for bit -- on two calls in a
row (the second on other pairs: the array keeps the first call's contents).  Collected last on purpose (pytest -x)."""
import ctypes as C

import numpy as np
import pytest

AS_ERANGE = -34


def _pairs(native, seed, n, max_len):
    rng = np.random.default_rng(seed)
    g = native.sim_genome(seed, 60_000)
    qs, ts = [], []
    for i in range(n):
        L = int(rng.integers(1, max_len + 1))
        p = int(rng.integers(0, g.size - max_len))
        t = g[p:p + L].copy()
        q = t.copy()
        for _ in range(int(rng.integers(0, max(1, L // 8)))):          # substitutions, then a few indels
            q[int(rng.integers(0, q.size))] = rng.choice(list(b"ACGT"))
        for _ in range(int(rng.integers(0, 4))):
            x = int(rng.integers(0, q.size))
            q = np.delete(q, x) if (q.size > 1 and rng.random() < 0.5) else np.insert(q, x, rng.choice(list(b"ACGT")))
        q = q[:max_len]
        if i % 97 == 0:
            q = q[:0]                                                   # empty query
        qs.append(np.ascontiguousarray(q, dtype=np.uint8))
        ts.append(t)
    return qs, ts


def _noalloc(native, torch, qs, ts, ks, stream):
    L = native.cuda_lib()
    n = len(qs)
    stride = 4 * ((max(max(len(q) for q in qs), max(len(t) for t in ts)) + 4) // 4)
    qb = np.zeros((n, stride), np.uint8)
    tb = np.zeros((n, stride), np.uint8)
    for i, (q, t) in enumerate(zip(qs, ts)):
        qb[i, :len(q)] = q
        tb[i, :len(t)] = t
    dev = torch.device("cuda", 0)
    d_q, d_t = torch.from_numpy(qb).to(dev), torch.from_numpy(tb).to(dev)
    d_ql = torch.tensor([len(q) for q in qs], dtype=torch.int32, device=dev)
    d_tl = torch.tensor([len(t) for t in ts], dtype=torch.int32, device=dev)
    d_k = None if ks is None else torch.from_numpy(np.ascontiguousarray(ks, dtype=np.int32)).to(dev)
    d_out = torch.full((n,), 12345, dtype=torch.int32, device=dev)
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    rc = L.as_edit_distance_device_noalloc(C.c_void_p(stream.cuda_stream), d_q.data_ptr(), stride, d_ql.data_ptr(), d_t.data_ptr(), stride,
                                           d_tl.data_ptr(), n, d_k.data_ptr() if d_k is not None else None, d_out.data_ptr())
    stream.synchronize()
    return rc, d_out.cpu().numpy(), free0 - torch.cuda.mem_get_info()[0]


@pytest.mark.gpu
def test_no_alloc_edit_distance_equals_batched_entry(native):
    import torch
    torch.cuda.set_device(0)
    L = native.cuda_lib()
    m = native.GpuMapper.build([native.sim_genome(5, 20_000)], 0)        # any handle serves as_edit_distance_batch
    stream = torch.cuda.Stream(device=torch.device("cuda", 0))
    try:
        for call, (seed, n, mx) in enumerate([(31, 3000, 256), (32, 5000, 150)]):
            qs, ts = _pairs(native, seed, n, mx)
            ks = None if call == 0 else np.where(np.arange(n) % 3 == 0, 2, -1).astype(np.int32)      # some bounded: -1 when exceeded
            exp = m.edit_distance_batch(qs, ts, ks)
            rc, got, grew = _noalloc(native, torch, qs, ts, ks, stream)
            assert rc == 0, L.as_last_error()
            assert np.array_equal(got, exp), np.flatnonzero(got != exp)[:10]
            if call == 1:
                assert grew <= (1 << 20), grew        # device memory in use did not grow across the (warmed-up) call
        # a pair beyond AS_NOALLOC_MAX_LEN is flagged, its neighbours are not disturbed
        qs, ts = _pairs(native, 33, 64, 200)
        exp = m.edit_distance_batch(qs, ts)
        ts[5] = np.frombuffer(b"ACGT" * 80, np.uint8)                                                # 320 bases
        rc, got, _ = _noalloc(native, torch, qs, ts, None, stream)
        assert rc == 0 and got[5] == AS_ERANGE
        keep = np.arange(64) != 5
        assert np.array_equal(got[keep], exp[keep])
        # more pairs than the static scratch holds: refused, said so, nothing launched
        d = torch.zeros(8, dtype=torch.int32, device="cuda:0")
        rc = L.as_edit_distance_device_noalloc(None, d.data_ptr(), 4, d.data_ptr(), d.data_ptr(), 4, d.data_ptr(), 32769, None, d.data_ptr())
        assert rc == AS_ERANGE and b"at most 32768" in L.as_last_error()
    finally:
        m.close()
