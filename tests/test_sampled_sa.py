"""csa_access (fmidx.c:315-331) through the C ABI (as_csa_access_batch): the LF walk to the sampled suffix array against the
oracle's restatement (pinned on the reference's object code in test_oracle_vs_ref.py) and against SA[row] itself."""
import numpy as np
import pytest

import oracle_lib

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gpu_uniform(native, idx_uniform):
    ix, _ = idx_uniform
    m = oracle_lib.gpu_mapper(native, ix, 0)
    yield m
    m.close()


@pytest.fixture(scope="module")
def gpu_repeat(native, idx_repeat):
    ix, _ = idx_repeat
    m = oracle_lib.gpu_mapper(native, ix, 0)
    yield m
    m.close()


@pytest.mark.parametrize("ratio", [1, 4, 5, 32, 64])
def test_csa_access(gpu_uniform, idx_uniform, gpu_repeat, idx_repeat, ratio):
    """csa_access (fmidx.c:315-331): the LF walk to the sampled SA returns SA[row] at every row, and the walk as the
    reference wrote it equals the oracle's literal restatement (pinned on the reference's object code in
    tests/test_oracle_vs_ref.py).  Rows past the index give 0; ratios alternate on one handle (resampling)."""
    for mapper, (ix, _) in ((gpu_uniform, idx_uniform), (gpu_repeat, idx_repeat)):
        rows = np.arange(ix.n, dtype=np.uint64) if ix.n <= 400000 else \
            np.random.default_rng(ratio).integers(0, ix.n, 300000).astype(np.uint64)
        rows = np.concatenate([rows, np.array([0, ix.n - 1, ix.n, ix.n + 7, 2**40], np.uint64)])
        want = ix.csa_access_batch(rows, ratio)
        got = mapper.csa_access_batch(rows, ratio)
        assert np.array_equal(got, want)
        inside = rows < ix.n
        assert np.array_equal(got[inside], ix.sa[rows[inside]]) and not got[~inside].any()
        assert np.array_equal(mapper.csa_access_batch(rows, ratio, literal=True),
                              ix.csa_access_batch(rows, ratio, literal=True))
        assert np.array_equal(got[inside], mapper.sa_access_batch(rows[inside]))


def test_csa_access_errors(gpu_uniform):
    with pytest.raises(Exception):
        gpu_uniform.csa_access_batch(np.zeros(4, np.uint64), 0)
    assert gpu_uniform.csa_access_batch(np.zeros(0, np.uint64), 4).size == 0


def test_csa_access_human_scale(native):
    """40-bit rows (3.1 Gbp, 6.2 G rows): the 32 + 8-bit sample and the super-block counts of the walk.  Needs a whole
    B200; skipped on anything smaller."""
    import torch
    free_b, _ = torch.cuda.mem_get_info(0)
    try:
        host_kb = int(next(l for l in open("/proc/meminfo") if l.startswith("MemAvailable")).split()[1])
    except (OSError, StopIteration, ValueError):
        host_kb = 0
    if free_b < 150 * 2**30 or host_kb < 48 * 2**20:
        pytest.skip("needs 150 GB of free HBM and 48 GB of free host memory")
    g = native.sim_genome(11, 3_100_000_000)
    m = native.GpuMapper.build([g], 0)
    try:
        N = 2 * g.size + 1
        rows = np.random.default_rng(17).integers(0, N, 2_000_000).astype(np.uint64)
        rows[:4] = (0, N - 1, (1 << 32) - 1, 1 << 32)
        sa = m.sa_access_batch(rows)
        assert sa.max() >= 1 << 32
        assert np.array_equal(m.csa_access_batch(rows, 4), sa)
        assert np.array_equal(m.csa_access_batch(rows, 32), sa)
    finally:
        m.close()


# ---------------------------------------------------------------------------------------------------
# PART 1 over the sampled suffix array (as_open_opts.sampled_sa): the handle keeps SA[r * i] only (_build_csa,
# fmidx.c:152-163), no full SA and no SA + context table, and resolves every vote position by csa_access' LF walk
# ---------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("ratio", [4, 32, 7])
def test_map_over_sampled_sa(native, idx_repeat, ratio):
    from test_gpu_parity import check_vote, compare_map
    from conftest import reads_from
    ix, gs = idx_repeat
    for opts in (native.open_opts(sampled_sa=ratio), native.open_opts(sampled_sa=ratio, lc2_hlen=13)):
        m = oracle_lib.gpu_mapper(native, ix, 0, opts)
        try:
            rows = np.random.default_rng(ratio).integers(0, ix.n, 20000).astype(np.uint64)
            assert np.array_equal(m.sa_access_batch(rows), ix.sa[rows])          # served by the walk
            for k, (lmin, lmax, e, sl, th) in enumerate([(150, 150, 0.02, 20, 300), (100, 250, 0.08, 20, 300), (60, 260, 0.05, 24, 2),
                                                        (80, 200, 0.03, 32, 1000), (80, 200, 0.03, 12, 300)]):
                reads, lens, _ = reads_from(native, 1500 + k, gs, 300, lmin, lmax, e, e, 0.7)
                for wide in (0, native.AS_F_WIDE_INDEX):
                    for ectx in (0, 1):
                        check_vote(m, ix, reads, lens, sl, th, flags=wide, vote_ectx=ectx)
            reads, lens, _ = reads_from(native, 1510, gs, 30, 1000, 5000, 0.05, 0.15, 0.34)
            check_vote(m, ix, reads, lens)
            reads, lens, stride = reads_from(native, 1520, gs, 1200, 100, 300, 0.0, 0.08, 0.6)
            compare_map(native, m, ix, reads, lens, stride)
        finally:
            m.close()


def test_built_index_over_sampled_sa(native, idx_repeat):
    """as_gpu_build with sampled_sa: same records as the handle that keeps the whole array; the export still returns the
    reference's arrays (the .sa5 rows come from the walk)"""
    from conftest import reads_from
    ix, gs = idx_repeat
    full = native.GpuMapper.build(gs, 0)
    samp = native.GpuMapper.build(gs, 0, opts=native.open_opts(sampled_sa=8))
    try:
        reads, lens, stride = reads_from(native, 1530, gs, 1500, 100, 400, 0.0, 0.1, 0.5)
        p = full.params(flags=native.AS_F_MYERS)
        a, aops = full.map_batch(reads.copy(), lens, params=p)
        b, bops = samp.map_batch(reads.copy(), lens, params=samp.params(flags=native.AS_F_MYERS))
        assert np.array_equal(a, b) and np.array_equal(aops, bops)
        ex = samp.export(want=("sa", "csa"))
        assert np.array_equal(ex["sa"], ix.sa)
        rec_a, runs_a = full.map_batch_packed(native.pack_reads(reads, lens), lens)
        rec_b, runs_b = samp.map_batch_packed(native.pack_reads(reads, lens), lens)
        assert np.array_equal(rec_a, rec_b) and np.array_equal(runs_a, runs_b)
    finally:
        full.close()
        samp.close()
