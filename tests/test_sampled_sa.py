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
