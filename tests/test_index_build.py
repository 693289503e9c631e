"""Index construction on the GPU (as_gpu_build) against the oracle's restatement of accidx
(asindex.c:78-153, fmidx.c:76-199, lchash.c:52-73): every exported array must be identical -- the suffix
array of a text is unique, so the files are byte-identical to the reference's."""
import numpy as np
import pytest

import oracle_lib

from conftest import reads_from

pytestmark = pytest.mark.gpu


def check_export(ix, built):
    ex = built.export()
    n = ix.n
    assert np.array_equal(ex["sa"], ix.sa), "suffix array"
    assert np.array_equal(ex["bwt"], ix.bwt), "BWT"
    assert np.array_equal(ex["C"], ix.Ctab), "C table"
    o = ix.O
    m = o.size if n % 32 else o.size - 4     # the reference leaves the block at row N unwritten when N % 32 == 0
    assert np.array_equal(ex["O"][:m], o[:m]), "O table"
    assert np.array_equal(ex["lc"], ix.lc), "lc table"
    from oracle_lib import _np_view
    csa = _np_view(ix.s.csa, np.uint64, ix.s.csa_len)
    assert np.array_equal(ex["csa"], csa), "CSA"
    sa5 = built.export(sa_entry_bytes=5)["sa"].reshape(-1, 5)
    lo = sa5[:, :4].copy().view("<u4").ravel().astype(np.uint64)
    assert np.array_equal(lo | (sa5[:, 4].astype(np.uint64) << 32), ix.sa)


def test_build_uniform(native, idx_uniform):
    ix, gs = idx_uniform
    built = native.GpuMapper.build(gs, 0)
    assert built.N == ix.n
    check_export(ix, built)
    built.close()


def test_build_repeat_rich_two_sequences(native, idx_repeat):
    """long exact repeats force several prefix-doubling rounds"""
    ix, gs = idx_repeat
    built = native.GpuMapper.build(gs, 0)
    check_export(ix, built)
    # and the handle maps exactly like one opened from the reference layout
    opened = oracle_lib.gpu_mapper(native, ix, 0)
    reads, lens, stride = reads_from(native, 31, gs, 500, 100, 400, 0.0, 0.1, 0.5)
    p = built.params(flags=native.AS_F_MYERS)
    a, aops = built.map_batch(reads.copy(), lens, params=p)
    b, bops = opened.map_batch(reads.copy(), lens, params=p)
    assert np.array_equal(a, b) and np.array_equal(aops, bops)
    built.close()
    opened.close()


@pytest.mark.parametrize("text", [b"A", b"ACGT", b"AAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAAA",
                                  b"ACGTACGTACGTACGTACGTACGTACGTACGTACGTACGTACGTACGTACGTACGTACGTACGTACGTACGT" * 3,
                                  b"GATTACA" * 19])
def test_build_degenerate_texts(native, orc, text):
    """tiny and maximally repetitive texts (every suffix ties with another for a long stretch; N % 32 == 0 too)"""
    ix = orc.OracleIndex.build(["t"], [text], hlen=4)
    built = native.GpuMapper.build([text], 0, hlen=4)
    ex = built.export(hlen=4)
    assert np.array_equal(ex["sa"], ix.sa) and np.array_equal(ex["bwt"], ix.bwt) and np.array_equal(ex["lc"], ix.lc)
    assert np.array_equal(ex["C"], ix.Ctab)
    m = ix.O.size if ix.n % 32 else ix.O.size - 4
    assert np.array_equal(ex["O"][:m], ix.O[:m])
    built.close()
    ix.close()


def test_known_suffix_array(native):
    """test/test-fmidx.cc:76-99: BWT `ard$rcaaaabb` <-> SA {11,10,7,0,3,5,8,1,4,6,9,2} of abracadabra$ -- same
    structure on a DNA text: the SA must sort the suffixes and the BWT be the preceding characters."""
    t = b"GATTACAGATTACACATTAG"
    built = native.GpuMapper.build([t], 0, hlen=3)
    ex = built.export(hlen=3)
    cat, _ = native.cat_image([t])
    cat = bytes(cat)
    sa = [int(x) for x in ex["sa"]]
    assert sorted(range(len(cat)), key=lambda i: cat[i:]) == sa
    assert bytes(ex["bwt"]) == bytes(cat[i - 1] if i else ord("$") for i in sa)
    built.close()


def test_build_rejects_bad_text(native):
    with pytest.raises(native.AccSeqError):
        native.GpuMapper.build([b"ACGTNNACGT" * 5], 0)     # cat_image maps N -> N: not ACGT


# ---------------------------------------------------------------------------------------------------
# partitioned builder (as_build_big.cu): the path 3.1 Gbp genomes take, forced onto small texts
# ---------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("part_max", [1, 5000, 200000])
def test_partitioned_build_matches_oracle(native, idx_repeat, part_max):
    """as_open_opts.build_part_max caps the suffixes sorted at once: 1 = one partition per 4-character bin; the planted repeats
    need many 21-character refinement rounds inside their partitions"""
    ix, gs = idx_repeat
    built = native.GpuMapper.build(gs, 0, opts=native.open_opts(build_part_max=part_max))
    assert built.N == ix.n
    check_export(ix, built)
    built.close()


@pytest.mark.parametrize("text", [b"A", b"ACGT", b"A" * 300, b"ACGTACGT" * 40, b"GATTACA" * 19])
def test_partitioned_build_degenerate_texts(native, orc, text):
    ix = orc.OracleIndex.build(["t"], [text], hlen=4)
    built = native.GpuMapper.build([text], 0, hlen=4, opts=native.open_opts(build_part_max=7))
    ex = built.export(hlen=4)
    assert np.array_equal(ex["sa"], ix.sa) and np.array_equal(ex["bwt"], ix.bwt) and np.array_equal(ex["lc"], ix.lc)
    assert np.array_equal(ex["C"], ix.Ctab)
    built.close()
    ix.close()
