"""Shared fixtures.  `-m "not gpu"` runs here on CPU (oracle vs the compiled reference, host logic, ABI
symbols); `-m gpu` runs on a B200 and compares the CUDA path, called through the C ABI, with the oracle."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run by the driver with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def native():
    import longreadmapper_b200 as lm
    if not (os.path.exists(lm.capi.CUDA_LIB) and os.path.exists(lm.capi.HOST_LIB)):
        lm.build_native()
    return lm


@pytest.fixture(scope="session")
def orc():
    import oracle_lib
    oracle_lib.lib()
    return oracle_lib


def make_uniform(lm, seed, n):
    return lm.sim_genome(seed, n)


def make_repeat_rich(lm, seed):
    """Two sequences, ~220 kbp, dispersed diverged + exact repeats, tandem blocks, homopolymers
    (the set on which tie-break bugs show)."""
    a = lm.sim_genome(seed, 150_000)
    b = lm.sim_genome(seed + 1, 70_000)
    lm.sim_plant_repeats(seed + 2, a, n_families=6, unit_len=5000, copies=5, divergence=0.01)
    lm.sim_plant_repeats(seed + 3, a, n_families=8, unit_len=1500, copies=6, divergence=0.0)
    lm.sim_plant_repeats(seed + 4, b, n_families=4, unit_len=1500, copies=4, divergence=0.0)
    # share a family between the sequences
    b[1000:6000] = a[20_000:25_000]
    return a, b


def make_high_copy(lm, seed):
    """~400 kbp with exact repeat families of 12, 50 and 250 copies: reads inside them vote for more distinct buckets in
    one phase than the shared-memory table of PART 1 holds (8 slots per lane), so they go through the overflow pass."""
    g = lm.sim_genome(seed, 400_000)
    lm.sim_plant_repeats(seed + 1, g, n_families=3, unit_len=2000, copies=12, divergence=0.0)
    lm.sim_plant_repeats(seed + 2, g, n_families=2, unit_len=1000, copies=50, divergence=0.0)
    lm.sim_plant_repeats(seed + 3, g, n_families=1, unit_len=400, copies=250, divergence=0.0)
    lm.sim_plant_repeats(seed + 4, g, n_families=2, unit_len=6000, copies=14, divergence=0.002)
    return g


@pytest.fixture(scope="session")
def idx_high_copy(native, orc):
    g = make_high_copy(native, 71)
    ix = orc.OracleIndex.build(["rep"], [g.tobytes()])
    yield ix, [g]
    ix.close()


@pytest.fixture(scope="session")
def idx_uniform(native, orc):
    g = make_uniform(native, 11, 200_000)
    ix = orc.OracleIndex.build(["chr1"], [g.tobytes()])
    yield ix, [g]
    ix.close()


@pytest.fixture(scope="session")
def idx_repeat(native, orc):
    a, b = make_repeat_rich(native, 23)
    ix = orc.OracleIndex.build(["seqA", "seqB"], [a.tobytes(), b.tobytes()])
    yield ix, [a, b]
    ix.close()


def reads_from(native, seed, genomes, n_each, len_min, len_max, err_min, err_max, sub_frac):
    """Reads sampled from every sequence, one common stride."""
    stride = len_max + 1
    outs, lens = [], []
    for gi, g in enumerate(genomes):
        r, l, _, _ = native.sim_reads(seed + gi, g, n_each, len_min, len_max, err_min, err_max, sub_frac)
        outs.append(r)
        lens.append(l)
    return np.ascontiguousarray(np.concatenate(outs)), np.concatenate(lens), stride
