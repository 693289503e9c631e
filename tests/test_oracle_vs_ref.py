"""CPU tests (-m "not gpu"): pin the oracle (oracle/oracle.c) against

  * the golden fixtures under tests/golden/, which were produced by the reference's OWN object code
    (tests/golden/make_golden.py; they travel, /root/reference does not), and
  * oracle/_ref live (libref.so = fmidx.c/lchash.c/histo.c objects, libedlib_ref.so = edlib.cpp), when present.

Known-answer vectors of the reference that exist for this path: edlib/ed_test.c:14-19,
test/test-fmidx.cc:76-99 (abracadabra$ BWT/SA), test/test-ui40read.cc:12-36 (ui40 byte layout), and the
invariants of test/test-lchash.cc:30-48 / test/test-fmidx.cc:30-41.
"""
import ctypes as C
import gzip
import hashlib
import json
import os

import numpy as np
import pytest

import oracle_lib

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = os.path.join(HERE, "golden")

needs_ref = pytest.mark.skipif(not oracle_lib.have_ref(), reason="oracle/_ref not built (no /root/reference here)")


def read_fasta_gz(path):
    names, seqs = [], []
    for line in gzip.open(path, "rb").read().split(b"\n"):
        if line.startswith(b">"):
            names.append(line[1:].decode())
            seqs.append(b"")
        elif line:
            seqs[-1] += line
    return names, seqs


def read_fastq_gz(path):
    lines = gzip.open(path, "rb").read().split(b"\n")
    return [(lines[i][1:].decode(), lines[i + 1], lines[i + 3]) for i in range(0, len(lines) - 3, 4)]


@pytest.fixture(scope="module")
def gold_index(orc):
    names, seqs = read_fasta_gz(os.path.join(GOLD, "repeat.fa.gz"))
    ix = orc.OracleIndex.build(names, seqs)
    yield ix, names, seqs
    ix.close()


# ------------------------------------------------------------------------------------------------
# index files: byte-identical to what the reference's accidx wrote
# ------------------------------------------------------------------------------------------------
def test_index_files_match_reference_sha256(gold_index, tmp_path):
    ix, names, seqs = gold_index
    fa = str(tmp_path / "repeat.fa")
    ix.write(fa)
    want = json.load(open(os.path.join(GOLD, "index.sha256.json")))
    for ext, sha in want.items():
        got = hashlib.sha256(open(fa + ext, "rb").read()).hexdigest()
        assert got == sha, f"{ext} differs from the file accidx_ref wrote"
    # and the loader round-trips them
    ld = oracle_lib.OracleIndex.load(fa)
    assert ld.n == ix.n and np.array_equal(ld.sa, ix.sa) and np.array_equal(ld.bwt, ix.bwt)
    assert np.array_equal(ld.lc, ix.lc) and np.array_equal(ld.Ctab, ix.Ctab) and ld.mta == ix.mta
    ld.close()


# ------------------------------------------------------------------------------------------------
# lookups: lc_aln / fmi_aln of the reference's object code on 4000 seeds
# ------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("qlen", [12, 20, 32, 64])
def test_lookups_match_reference(gold_index, qlen):
    ix, _, _ = gold_index
    d = np.load(os.path.join(GOLD, "lookups.npz"))
    seeds, want = d[f"seeds{qlen}"], d[f"res{qlen}"]
    k, l, r = ix.lc_aln_batch(seeds, qlen)
    assert np.array_equal(r, want[:, 2])
    hit = want[:, 2] > 0
    assert np.array_equal(k[hit], want[hit, 0]) and np.array_equal(l[hit], want[hit, 1])
    for i in range(0, len(seeds), 7):
        k2, l2, r2 = ix.fmi_aln(seeds[i].tobytes())
        assert r2 == int(want[i, 5])
        if r2:
            assert (k2, l2) == (int(want[i, 3]), int(want[i, 4]))
            # test-fmidx.cc:30-41: the text at every SA row of the interval spells the query
            for row in range(k2, min(l2, k2 + 3) + 1):
                p = int(ix.sa[row])
                assert ix.cat[p:p + qlen].tobytes() == seeds[i].tobytes()
    # test-lchash.cc:30-42: table-accelerated == plain backward search whenever there is a hit
    both = (want[:, 2] > 0)
    assert np.array_equal(want[both, 0], want[both, 3]) and np.array_equal(want[both, 1], want[both, 4])


def test_histo_traces_match_reference(orc):
    L = orc.lib()
    d = np.load(os.path.join(GOLD, "lookups.npz"))
    keys, lens, top = d["histo_keys"], d["histo_lens"], d["histo_top"]

    class E(C.Structure):
        _fields_ = [("key", C.c_uint64), ("val", C.c_uint64), ("bucket", C.c_uint64)]

    class H(C.Structure):
        _fields_ = [("e", C.c_void_p), ("cap", C.c_uint32), ("size", C.c_uint32)]

    L.orc_histo_init.argtypes = [C.POINTER(H), C.c_uint32]
    L.orc_histo_add.argtypes = [C.POINTER(H), C.c_uint64]
    L.orc_histo_top2.argtypes = [C.POINTER(H), C.POINTER(E)]
    L.orc_histo_top2.restype = C.c_uint64
    L.orc_histo_free.argtypes = [C.POINTER(H)]
    off = 0
    for t, n in enumerate(lens):
        h = H()
        L.orc_histo_init(C.byref(h), 300)
        for kx in keys[off:off + n]:
            L.orc_histo_add(C.byref(h), int(kx))
        out = (E * 2)()
        v = L.orc_histo_top2(C.byref(h), out)
        L.orc_histo_free(C.byref(h))
        got = [out[0].key, out[0].val, out[0].bucket, out[1].key, out[1].val, out[1].bucket, v]
        assert got == [int(x) for x in top[t]], f"histo trace {t}"
        off += int(n)


# ------------------------------------------------------------------------------------------------
# whole path: SAM written by the compiled reference `accaln` on the golden reads
# ------------------------------------------------------------------------------------------------
def test_map_batch_matches_reference_sam(gold_index):
    ix, names, _ = gold_index
    recs = read_fastq_gz(os.path.join(GOLD, "reads.fq.gz"))
    want = oracle_lib.parse_sam(gzip.open(os.path.join(GOLD, "expected.sam.gz"), "rt").read())
    assert len(want) == len(recs)
    lens = np.array([len(s) for _, s, _ in recs], np.uint32)
    stride = int(lens.max()) + 1
    reads = np.zeros((len(recs), stride), np.uint8)
    for i, (_, s, _) in enumerate(recs):
        reads[i, :len(s)] = np.frombuffer(s, np.uint8)
    res, ops, _ = ix.map_batch(reads, stride, lens, myers=False)
    n_checked = 0
    for i, w in enumerate(want):
        r = res[i]
        assert w["qname"] == recs[i][0]
        if r["seq_id"] < 0 or lens[i] <= 20:
            # seq_lookup failed (reference: uninitialised m[], SURVEY 8(c) exclusions) or len <= seed_len (UB)
            if r["seq_id"] < 0:
                assert r["flag"] == 4 and r["mapq"] == 0
            continue
        n_checked += 1
        assert (w["flag"], w["mapq"]) == (int(r["flag"]), int(r["mapq"])), f"read {i} flag/mapq"
        assert w["rname"] == names[r["seq_id"]] and w["pos"] == int(r["off"]) + 1, f"read {i} position"
        assert w["seq"] == reads[i, :lens[i]].tobytes().decode(), f"read {i} SEQ (strand fix)"
        # CIGAR/ED pin the GACT stand-in only (submodule absent): same code linked behind gact.h
        assert w["ed"] == int(r["ed"])
        assert w["cigar"] == oracle_lib.cigar_text(ops[i, :r["n_ops"]])
    assert n_checked > 600


# ------------------------------------------------------------------------------------------------
# known answers of the reference tree
# ------------------------------------------------------------------------------------------------
ED_S1 = b"AACTCCGACTGCTCAACAAGAGGTCACACCGTTAGGTCGACCTCAGCCCCGTACTGGCCGAAAGCGTGAGATGACACGGGGCAAGCTTGATGTTCCGAAC"
ED_S2 = b"AACTGCGACTGCTCAACAAGATGTCACACCGTTAGGTCGACCTCAGCCCCGTACTGGCCGAAAGCGTGAGATGACACGGGGCAAGCTTGATGTTCCGAAC"


def test_edlib_known_answer():
    """edlib/ed_test.c:14-19 prints 2 and 100M"""
    assert oracle_lib.edit_distance(np.frombuffer(ED_S1, np.uint8), np.frombuffer(ED_S2, np.uint8), 5000000) == 2
    score, ops, _ = oracle_lib.gact(np.frombuffer(ED_S1, np.uint8), np.frombuffer(ED_S2, np.uint8))
    assert score == 2 and oracle_lib.cigar_text(ops) == "100M"


def test_abracadabra_suffix_array(orc):
    """test/test-fmidx.cc:76-99: SA {11,10,7,0,3,5,8,1,4,6,9,2}, BWT ard$rcaaaabb of abracadabra$"""
    t = np.frombuffer(b"abracadabra$", np.uint8)
    sa = np.zeros(t.size, np.uint64)
    assert orc.lib().orc_suffix_array(t.ctypes.data, t.size, sa.ctypes.data) == 0
    assert sa.tolist() == [11, 10, 7, 0, 3, 5, 8, 1, 4, 6, 9, 2]
    assert bytes(t[(sa.astype(np.int64) - 1) % t.size]) == b"ard$rcaaaabb"


def test_ui40_layout(gold_index, tmp_path):
    """test/test-ui40read.cc:12-36: an entry is u32 low then u8 high, 5 bytes, little endian"""
    ix, _, _ = gold_index
    fa = str(tmp_path / "x.fa")
    ix.write(fa)
    raw = np.fromfile(fa + ".cat.sa5", np.uint8).reshape(-1, 5)
    assert raw.shape[0] == ix.n
    lo = raw[:, :4].copy().view("<u4").ravel().astype(np.uint64)
    assert np.array_equal(lo | (raw[:, 4].astype(np.uint64) << 32), ix.sa)
    assert int(ix.sa[0]) == ix.n - 1          # '$' sorts first


# ------------------------------------------------------------------------------------------------
# invariants on the oracle itself
# ------------------------------------------------------------------------------------------------
def test_occ_definition(gold_index):
    ix, _, _ = gold_index
    L = oracle_lib.lib()
    bwt = ix.bwt
    rng = np.random.default_rng(3)
    for loc in rng.integers(0, ix.n, 300):
        for c in b"ACGT":
            assert L.orc_occ(ix.h, bytes([c]), int(loc)) == int((bwt[:loc + 1] == c).sum())


def test_gact_single_tile_equals_nw_distance():
    rng = np.random.default_rng(5)
    acgt = np.frombuffer(b"ACGT", np.uint8)
    for _ in range(60):
        n = int(rng.integers(1, 300))
        q = acgt[rng.integers(0, 4, n)]
        t = q.copy()
        m = rng.random(n) < 0.08
        t[m] = acgt[rng.integers(0, 4, int(m.sum()))]
        if n > 10 and rng.random() < 0.5:
            cut = int(rng.integers(1, n - 1))
            t = np.concatenate([t[:cut], t[cut + 1:], acgt[rng.integers(0, 4, 1)]])
        score, ops, cells = oracle_lib.gact(q, t)
        assert score == oracle_lib.edit_distance(q, t) and cells == q.size * t.size
        assert int((ops != 2).sum()) == q.size and int((ops != 1).sum()) == t.size
        assert int((ops != 0).sum()) == score


def test_gact_tiled_consumes_everything():
    rng = np.random.default_rng(6)
    acgt = np.frombuffer(b"ACGT", np.uint8)
    for T, O in ((64, 16), (128, 48), (320, 120)):
        q = acgt[rng.integers(0, 4, 2000)]
        keep = rng.random(2000) > 0.05
        t = q[keep].copy()
        m = rng.random(t.size) < 0.05
        t[m] = acgt[rng.integers(0, 4, int(m.sum()))]
        score, ops, _ = oracle_lib.gact(q, t, T, O)
        assert score >= oracle_lib.edit_distance(q, t) and score >= 0
        assert int((ops != 2).sum()) == q.size and int((ops != 1).sum()) == t.size
        # the ops replay to the two sequences
        qi = ti = 0
        for o in ops:
            if o in (0, 3):
                assert (q[qi] == t[ti]) == (o == 0)
            qi += o != 2
            ti += o != 1


def test_parse_cigar():
    assert oracle_lib.cigar_text(np.array([0, 0, 3, 1, 1, 2, 0], np.uint8)) == "3M2I1D1M"
    assert oracle_lib.cigar_text(np.array([], np.uint8)) == ""


# ------------------------------------------------------------------------------------------------
# live against oracle/_ref (the reference's object code), when it is present
# ------------------------------------------------------------------------------------------------
@needs_ref
def test_live_lookups_against_libref(gold_index, tmp_path):
    ix, _, _ = gold_index
    fa = str(tmp_path / "r.fa")
    ix.write(fa)
    R = oracle_lib.ref()
    fmi, lch = oracle_lib.RefFmi(), oracle_lib.RefLch()
    R.fmi_read(C.byref(fmi), (fa + ".cat").encode())
    R.lc_read((fa + ".cat.lch").encode(), C.byref(lch))
    rng = np.random.default_rng(77)
    cat = ix.cat
    for qlen in (13, 20, 27):
        pos = rng.integers(0, ix.n - 40, 300)
        seeds = np.stack([cat[p:p + qlen] for p in pos]).copy()
        mut = rng.random(300) < 0.4
        seeds[mut, rng.integers(0, qlen, 300)[mut]] = np.frombuffer(b"ACGT", np.uint8)[rng.integers(0, 4, int(mut.sum()))]
        k, l, r = ix.lc_aln_batch(seeds, qlen)
        for i in range(300):
            kk, ll = C.c_uint64(1), C.c_uint64(ix.n - 1)
            rr = R.lc_aln(seeds[i].tobytes(), qlen, C.byref(kk), C.byref(ll), C.byref(fmi), C.byref(lch))
            assert rr == int(r[i])
            if rr:
                assert (kk.value, ll.value) == (int(k[i]), int(l[i]))


@needs_ref
def test_live_csa_access_against_libref(gold_index, tmp_path):
    """csa_access (fmidx.c:315-331) of the reference's object code on the .mfi it reads (csa_ratio 4, fmidx.c:194)
    against the oracle's literal restatement, at every row."""
    ix, _, _ = gold_index
    fa = str(tmp_path / "r.fa")
    ix.write(fa)
    R = oracle_lib.ref()
    fmi = oracle_lib.RefFmi()
    R.fmi_read(C.byref(fmi), (fa + ".cat").encode())
    assert fmi.csa_ratio == 4
    rows = np.arange(ix.n, dtype=np.uint64)
    lit = ix.csa_access_batch(rows, 4, literal=True)
    step = max(1, ix.n // 20000)
    for r in range(0, ix.n, step):
        assert R.csa_access(C.byref(fmi), r) == int(lit[r]), r


def test_csa_access_walk_returns_the_suffix_array(gold_index):
    """The property the reference's test states (csa_access(i) == sa[i], test/test-fmidx.cc:104-106) holds for the
    LF walk at every ratio; sampled rows are exact in the literal form too."""
    ix, _, _ = gold_index
    rows = np.arange(ix.n, dtype=np.uint64)
    for ratio in (1, 4, 7, 32):
        assert np.array_equal(ix.csa_access_batch(rows, ratio), ix.sa)
        lit = ix.csa_access_batch(rows, ratio, literal=True)
        assert np.array_equal(lit[::ratio], ix.sa[::ratio])
    assert ix.csa_access_batch(np.array([ix.n, ix.n + 5], np.uint64), 4).tolist() == [0, 0]


@needs_ref
def test_live_edit_distance_against_edlib():
    rng = np.random.default_rng(8)
    acgt = np.frombuffer(b"ACGT", np.uint8)
    for it in range(80):
        n, m = int(rng.integers(1, 400)), int(rng.integers(1, 400))
        q = acgt[rng.integers(0, 4, n)]
        t = acgt[rng.integers(0, 4, m)] if it % 2 else np.roll(q, 3)[:max(1, min(n, m))].copy()
        for k in (-1, 5, 40):
            want, _ = oracle_lib.edlib_nw(q.tobytes(), t.tobytes(), k)
            assert oracle_lib.edit_distance(q, t, k) == want


def _pairs_for_alignment(rng, count):
    acgt = np.frombuffer(b"ACGT", np.uint8)
    for it in range(count):
        n = int(rng.integers(1, 257))
        q = acgt[rng.integers(0, 4, n)]
        kind = it % 4
        if kind == 0:                                   # unrelated
            t = acgt[rng.integers(0, 4, int(rng.integers(1, 300)))]
        elif kind == 1:                                 # mutated copy: substitutions, insertions, deletions
            out = []
            for c in q:
                r = rng.random()
                if r < 0.06:
                    out.append(acgt[rng.integers(0, 4)])
                elif r < 0.10:
                    continue
                elif r < 0.14:
                    out += [c, acgt[rng.integers(0, 4)]]
                else:
                    out.append(c)
            t = np.array(out or [q[0]], np.uint8)
        elif kind == 2:                                 # low complexity: many ties
            unit = acgt[rng.integers(0, 4, int(rng.integers(1, 4)))]
            q = np.resize(unit, n)
            t = np.resize(unit, max(1, n + int(rng.integers(-6, 7))))
            t = t.copy()
            t[rng.integers(0, t.size)] = acgt[rng.integers(0, 4)]
        else:                                           # shifted copy
            t = np.roll(q, int(rng.integers(1, 9)))[: max(1, n - int(rng.integers(0, 5)))].copy()
        yield q, t


@needs_ref
def test_live_edit_alignment_against_edlib():
    """orc_edit_alignment_nw against the reference's own edlibAlign(EDLIB_MODE_NW, EDLIB_TASK_PATH) (libedlib_ref.so =
    edlib/edlib.cpp compiled as it lies): distance and every op of the path, on unrelated, mutated, low-complexity (tie-rich)
    and shifted pairs, with and without a bound k; and ed_test.c's known answer"""
    rng = np.random.default_rng(31)
    n_paths = 0
    for q, t in _pairs_for_alignment(rng, 400):
        for k in (-1, 12):
            want_d, want_ops = oracle_lib.edlib_nw_ops(q.tobytes(), t.tobytes(), k)
            got_d, got_ops = oracle_lib.edit_alignment(q, t, k)
            assert got_d == want_d
            if want_d >= 0:
                assert np.array_equal(got_ops, want_ops), (q.tobytes(), t.tobytes(), k)
                n_paths += 1
            else:
                assert got_ops.size == 0
    assert n_paths > 400
    d, ops = oracle_lib.edit_alignment(np.frombuffer(b"A" * 100, np.uint8), np.frombuffer(b"A" * 98 + b"TT", np.uint8))
    assert d == 2 and ops.size == 100                                        # edlib/ed_test.c:14-19: 2, 100M


def edlib_golden_pairs():
    z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "edlib_paths.npz"))
    qo, to, oo = np.concatenate([[0], np.cumsum(z["qlen"])]), np.concatenate([[0], np.cumsum(z["tlen"])]), np.concatenate([[0], np.cumsum(z["n_ops"])])
    for i in range(z["qlen"].size):
        yield (z["q"][qo[i]:qo[i + 1]], z["t"][to[i]:to[i + 1]], int(z["k"][i]), int(z["dist"][i]), z["ops"][oo[i]:oo[i + 1]])


def test_edit_alignment_golden():
    """the oracle against tests/golden/edlib_paths.npz: alignment paths the reference's compiled edlib produced
    (tests/golden/make_golden_edlib.py); runs without /root/reference"""
    n = 0
    for q, t, k, want_d, want_ops in edlib_golden_pairs():
        d, ops = oracle_lib.edit_alignment(q, t, k)
        assert d == want_d and np.array_equal(ops, want_ops)
        n += 1
    assert n == 120
