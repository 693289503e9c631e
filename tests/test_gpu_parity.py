"""Parity of the CUDA path (through the C ABI, libaccseq_b200.so) with the oracle: bit-exact on every
integer field -- SA intervals, SA values, voted locus, phases, strand, offset, FLAG, MAPQ, ED, CIGAR ops,
Myers distance.  The oracle itself is pinned against the compiled reference in test_oracle_vs_ref.py."""
import numpy as np
import pytest

import oracle_lib

from conftest import reads_from

pytestmark = pytest.mark.gpu

GACT_IMPL = {"warp": 4, "bitpar": 8}      # AS_F_GACT_WARP / AS_F_GACT_BITPAR: both kernels must agree with the oracle
FIELDS = ("loc", "off", "seq_id", "strand", "flag", "mapq", "ed", "valid", "myers", "phases", "from_outer")


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


def compare_map(native, mapper, ix, reads, lens, stride, **kw):
    seed_len = kw.get("seed_len", 20)
    thres = kw.get("thres", 300)
    T = kw.get("T", 320)
    O = kw.get("O", 120)
    r_cpu = reads.copy()
    r_gpu = reads.copy()
    exp, exp_ops, cells = ix.map_batch(r_cpu, stride, lens, seed_len=seed_len, thres=thres, T=T, O=O, myers=True)
    p = mapper.params(seed_len=seed_len, thres=thres, T=T, O=O, flags=native.AS_F_MYERS | GACT_IMPL.get(kw.get("gact"), 0),
                      **kw.get("knobs", {}))
    mapper.stats_reset()
    got, got_ops = mapper.map_batch(r_gpu, lens, params=p)
    for f in FIELDS:
        bad = np.nonzero(exp[f] != got[f])[0]
        assert bad.size == 0, f"{f}: {bad.size} reads differ, first {bad[0]}: exp {exp[f][bad[0]]} got {got[f][bad[0]]}"
    assert np.array_equal(exp["n_ops"], got["n_cigar_op"])
    for i in range(len(lens)):
        n = int(exp["n_ops"][i])
        assert np.array_equal(exp_ops[i, :n], got_ops[i, :n]), f"CIGAR ops differ on read {i}"
    assert np.array_equal(r_cpu, r_gpu), "in-place reverse complement differs"
    st = mapper.stats()
    assert st["total"] == len(lens)
    assert st["valid"] == int(exp["valid"].sum())
    assert st["dp_cells"] == cells
    return exp, st


# ---------------------------------------------------------------------------------------------------
# lookups (lchash.c:89-104, fmidx.c:295-313, fmidx.c:18-33)
# ---------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("qlen", [20, 12, 13, 32, 8])
def test_lc_aln_matches_oracle(gpu_uniform, idx_uniform, qlen):
    ix, (g,) = idx_uniform
    rng = np.random.default_rng(qlen)
    n = 4000
    pos = rng.integers(0, g.size - 64, n)
    seeds = np.stack([g[p:p + qlen] for p in pos])
    # half of them get one substitution so that many intervals die part-way
    mut = rng.integers(0, 2, n).astype(bool)
    where = rng.integers(0, qlen, n)
    seeds[mut, where[mut]] = np.frombuffer(b"ACGT", np.uint8)[rng.integers(0, 4, int(mut.sum()))]
    k, l, r = gpu_uniform.lc_aln_batch(seeds, qlen)
    ek, el, er = ix.lc_aln_batch(seeds, qlen)
    assert np.array_equal(r, er)
    live = er > 0
    assert np.array_equal(k[live], ek[live]) and np.array_equal(l[live], el[live])
    # dead intervals: the reference leaves (k,l) wherever the search stopped; same here
    assert np.array_equal(k, ek) and np.array_equal(l, el)


def test_lc_aln_equals_fmi_aln(gpu_uniform, idx_uniform):
    """The invariant test/test-lchash.cc:30-48 checks: table-accelerated == plain backward search, and the
    SA position spells the query."""
    ix, (g,) = idx_uniform
    rng = np.random.default_rng(5)
    pos = rng.integers(0, g.size - 64, 2000)
    q = np.stack([g[p:p + 64] for p in pos])
    k1, l1, r1 = gpu_uniform.lc_aln_batch(q, 64)
    k2, l2, r2 = gpu_uniform.fmi_aln_batch(q, 64)
    assert np.array_equal(r1, r2) and np.array_equal(k1, k2) and np.array_equal(l1, l2)
    assert (r1 >= 1).all()
    sa = gpu_uniform.sa_access_batch(k1)
    cat = ix.cat
    for i in range(0, 2000, 97):
        assert bytes(cat[int(sa[i]):int(sa[i]) + 64]) == bytes(q[i])


def test_sa_access(gpu_uniform, idx_uniform):
    ix, _ = idx_uniform
    rows = np.random.default_rng(1).integers(0, ix.n, 100000).astype(np.uint64)
    assert np.array_equal(gpu_uniform.sa_access_batch(rows), ix.sa[rows])
    assert np.array_equal(gpu_uniform.sa_access_batch(np.arange(ix.n, dtype=np.uint64)), ix.sa)


# ---------------------------------------------------------------------------------------------------
# PART 1 (alnmain.c:333-405)
# ---------------------------------------------------------------------------------------------------
def check_vote(mapper, ix, reads, lens, seed_len=20, thres=300, flags=0, **knobs):
    key, val, ph, fo = mapper.vote_batch(reads, lens, mapper.params(seed_len=seed_len, thres=thres, flags=flags, **knobs))
    for i in range(len(lens)):
        v = ix.vote(reads[i, : lens[i]], seed_len, thres)
        assert (int(key[i]), int(val[i]), int(ph[i]), int(fo[i])) == (v.best_key, v.best_val, v.phases, v.from_outer), i


def test_vote_uniform_short(native, gpu_uniform, idx_uniform):
    ix, gs = idx_uniform
    reads, lens, _ = reads_from(native, 100, gs, 3000, 150, 150, 0.01, 0.01, 0.9)
    check_vote(gpu_uniform, ix, reads, lens)


@pytest.mark.parametrize("wide", [0, 16])     # AS_F_WIDE_INDEX: the 64-bit path human-scale indexes take
def test_vote_repeat_rich(native, gpu_repeat, idx_repeat, wide):
    ix, gs = idx_repeat
    for k, (lmin, lmax, e) in enumerate([(100, 100, 0.0), (150, 150, 0.02), (250, 250, 0.05), (100, 250, 0.08)]):
        reads, lens, _ = reads_from(native, 200 + k, gs, 750, lmin, lmax, e, e, 0.7)
        check_vote(gpu_repeat, ix, reads, lens, flags=wide)


@pytest.mark.parametrize("wide", [0, 16])
def test_vote_overflow_pass_high_copy_repeats(native, idx_high_copy, wide):
    """exact repeat families of 12, 50 and 250 copies: a phase of a read inside one votes for more distinct buckets than the
    shared-memory table holds (8 per lane of the group), the read is queued and redone by the overflow pass
    (k_vote<32, BIG>) over a table in global memory.  Same votes as the oracle, and the queue was really used."""
    ix, gs = idx_high_copy
    m = oracle_lib.gpu_mapper(native, ix, 0)
    try:
        queued = 0
        for k, (n, lmin, lmax, e, sl, th) in enumerate([(1500, 150, 150, 0.01, 20, 300), (600, 150, 150, 0.02, 20, 1000),
                                                          (400, 100, 250, 0.03, 16, 300), (60, 1000, 6000, 0.08, 20, 300),
                                                          (40, 1000, 4000, 0.05, 20, 1000)]):
            reads, lens, _ = reads_from(native, 1300 + k, gs, n, lmin, lmax, e, e, 0.7)
            m.stats_reset()
            check_vote(m, ix, reads, lens, sl, th, flags=wide)
            st = m.stats()
            assert st["overflow_reads"] == 0
            queued += st["overflow_queued"]
            if k < 2:
                assert st["overflow_queued"] > 0, "no read took the overflow pass: the test does not reach it"
        assert queued > 50
        # and the whole path on them
        reads, lens, stride = reads_from(native, 1350, gs, 1200, 100, 300, 0.0, 0.05, 0.6)
        compare_map(native, m, ix, reads, lens, stride)
        assert m.stats()["overflow_queued"] > 0
    finally:
        m.close()


@pytest.mark.parametrize("h2", [0, 13])
@pytest.mark.parametrize("wide", [0, 16])
def test_vote_thread_kernel_equals_group_kernel(native, idx_repeat, idx_high_copy, h2, wide):
    """Short reads take the thread-per-read PART 1 kernel (k_vote_t), which hands what a thread cannot hold (a seed with more
    than two hits, more than 4 distinct buckets in a phase or in the outer list) to the group kernel through a
    queue; as_params.vote_thread = 0 sends everything through the group kernel.  Both equal the oracle, on the repeat-rich
    genome (with and without the deep table) and on the high-copy families (thread -> group -> overflow pass), and a larger
    batch gives the same votes either way."""
    for (ix, gs), seed in ((idx_repeat, 1500), (idx_high_copy, 1600)):
        m = oracle_lib.gpu_mapper(native, ix, 0, native.open_opts(lc2_hlen=h2) if h2 else None)
        try:
            for k, (lmin, lmax, e, sl, th) in enumerate([(150, 150, 0.02, 20, 300), (100, 150, 0.05, 20, 300), (36, 120, 0.03, 16, 50),
                                                        (150, 187, 0.01, 20, 1000), (21, 40, 0.0, 20, 300), (150, 150, 0.02, 24, 2)]):
                reads, lens, _ = reads_from(native, seed + k, gs, 250, lmin, lmax, e, e, 0.7)
                m.stats_reset()
                l0 = m.launch_count()
                check_vote(m, ix, reads, lens, sl, th, flags=wide, vote_thread=1)     # 1: also on the 64-bit path (default: small indexes)
                d_thread = m.launch_count() - l0            # pack, k_vote_t, group kernel over the queue, overflow pass
                assert m.stats()["vote_deferred"] < len(lens)
                l0 = m.launch_count()
                check_vote(m, ix, reads, lens, sl, th, flags=wide, vote_thread=0)
                assert d_thread == m.launch_count() - l0 + 1, "the thread-per-read kernel did not run"
            reads, lens, _ = reads_from(native, seed + 50, gs, 20000, 150, 150, 0.02, 0.02, 0.8)
            m.stats_reset()
            a = m.vote_batch(reads, lens, m.params(flags=wide, vote_thread=1))
            st = m.stats()
            b = m.vote_batch(reads, lens, m.params(flags=wide, vote_thread=0))
            for x, y in zip(a, b):
                assert np.array_equal(x, y)
            assert 0 < st["vote_deferred"] < len(lens)
        finally:
            m.close()


@pytest.mark.parametrize("seed_len,thres", [(12, 300), (16, 50), (24, 300), (32, 1000), (20, 2)])
def test_vote_parameters(native, gpu_repeat, idx_repeat, seed_len, thres):
    ix, gs = idx_repeat
    reads, lens, _ = reads_from(native, 300 + seed_len, gs, 300, 60, 260, 0.0, 0.06, 0.8)
    check_vote(gpu_repeat, ix, reads, lens, seed_len, thres)


@pytest.mark.parametrize("wide", [0, 16])
def test_vote_long_reads(native, gpu_repeat, idx_repeat, wide):
    ix, gs = idx_repeat
    reads, lens, _ = reads_from(native, 400, gs, 40, 1000, 6000, 0.05, 0.15, 0.34)
    check_vote(gpu_repeat, ix, reads, lens, flags=wide)


# deep k-mer table (as_lc2.cu): PART 1 must give the reference's votes whether a seed starts from the 12-mer table
# or from the h2-mer table, for every h2 (as_open_opts.lc2_hlen forces a depth the small test genomes would not get)
@pytest.mark.parametrize("h2", [13, 14, 16])
@pytest.mark.parametrize("wide", [0, 16])
def test_vote_deep_table(native, idx_repeat, h2, wide):
    ix, gs = idx_repeat
    m = oracle_lib.gpu_mapper(native, ix, 0, native.open_opts(lc2_hlen=h2))
    ectx = dict(vote_ectx=1)      # entries decide from their context bases (default only for large indexes)
    try:
        for k, (lmin, lmax, e, sl, th) in enumerate([(100, 100, 0.0, 20, 300), (150, 150, 0.02, 20, 300), (100, 250, 0.08, 20, 300),
                                                    (60, 260, 0.05, h2, 50), (60, 260, 0.05, 24, 2), (80, 200, 0.03, 32, 1000),
                                                    (80, 200, 0.03, 12, 300)]):
            reads, lens, _ = reads_from(native, 900 + k, gs, 300, lmin, lmax, e, e, 0.7)
            check_vote(m, ix, reads, lens, sl, th, flags=wide, **ectx)                          # deep table when sl >= h2
            check_vote(m, ix, reads, lens, sl, th, flags=wide | native.AS_F_NO_LC2, **ectx)     # the reference's 12-mer table
        reads, lens, _ = reads_from(native, 950, gs, 30, 1000, 5000, 0.05, 0.15, 0.34)
        check_vote(m, ix, reads, lens, flags=wide, **ectx)
    finally:
        m.close()


@pytest.mark.parametrize("wide", [0, 16])
@pytest.mark.parametrize("ectx", ["0", "1"])
@pytest.mark.parametrize("no_sax", ["0", "1"])
def test_vote_sa_context_table(native, idx_repeat, no_sax, wide, ectx):
    """sax (SA value + the 12 bases in front of it, as_lc2.cu) against the plain SA + text gathers: same votes, also for
    hits within 12 bases of the text start (context shorter than the entry), seeds whose tail is longer than the entry
    holds (seed_len - h2 > 12: the text gather stays) and exact repeats (two-row intervals).  vote_ectx=1 forces the
    path that decides one- and two-row deep-table entries from the context bases they carry (on by itself only for
    indexes larger than L2)."""
    ix, gs = idx_repeat
    m = oracle_lib.gpu_mapper(native, ix, 0, native.open_opts(lc2_hlen=13, no_sax=int(no_sax)))
    ectx = dict(vote_ectx=int(ectx))
    try:
        a = gs[0]
        L = 120
        starts = list(range(0, 40)) + [len(a) - L - k for k in range(0, 20)]
        reads = np.zeros((len(starts), L + 1), dtype=np.uint8)
        for i, p0 in enumerate(starts):
            reads[i, :L] = a[p0:p0 + L]
        lens = np.full(len(starts), L, dtype=np.uint32)
        for sl, th in [(20, 300), (14, 300), (24, 300), (28, 50), (32, 1000)]:
            check_vote(m, ix, reads, lens, sl, th, flags=wide, **ectx)
        for k, (lmin, lmax, e, sl, th) in enumerate([(100, 250, 0.05, 20, 300), (80, 200, 0.03, 32, 1000), (60, 260, 0.05, 27, 2)]):
            r2, l2, _ = reads_from(native, 970 + k, gs, 300, lmin, lmax, e, e, 0.7)
            check_vote(m, ix, r2, l2, sl, th, flags=wide, **ectx)
    finally:
        m.close()


def test_map_deep_table_built_index(native, idx_repeat):
    """the same through as_gpu_build and the whole path"""
    ix, gs = idx_repeat
    m = native.GpuMapper.build(gs, 0, opts=native.open_opts(lc2_hlen=15))
    try:
        reads, lens, stride = reads_from(native, 960, gs, 1200, 100, 250, 0.0, 0.08, 0.6)
        compare_map(native, m, ix, reads, lens, stride, knobs=dict(vote_ectx=1))
    finally:
        m.close()


# ---------------------------------------------------------------------------------------------------
# whole path (alnmain.c:330-479)
# ---------------------------------------------------------------------------------------------------
def test_map_uniform_short(native, gpu_uniform, idx_uniform):
    ix, gs = idx_uniform
    reads, lens, stride = reads_from(native, 500, gs, 4000, 150, 150, 0.01, 0.01, 0.9)
    exp, st = compare_map(native, gpu_uniform, ix, reads, lens, stride)
    assert exp["valid"].mean() > 0.95


@pytest.mark.parametrize("gact", ["warp", "bitpar"])
def test_map_repeat_rich_mixed(native, gpu_repeat, idx_repeat, gact):
    ix, gs = idx_repeat
    reads, lens, stride = reads_from(native, 600, gs, 1500, 100, 250, 0.0, 0.08, 0.6)
    compare_map(native, gpu_repeat, ix, reads, lens, stride, gact=gact)


@pytest.mark.parametrize("gact", ["warp", "bitpar"])
def test_map_long_reads_tiled(native, gpu_repeat, idx_repeat, gact):
    """reads longer than the GACT tile: several tiles per read, indels."""
    ix, gs = idx_repeat
    reads, lens, stride = reads_from(native, 700, gs, 60, 400, 5000, 0.05, 0.15, 0.34)
    compare_map(native, gpu_repeat, ix, reads, lens, stride, gact=gact)


@pytest.mark.parametrize("gact", ["warp", "bitpar"])
@pytest.mark.parametrize("T,O", [(64, 16), (128, 120), (320, 0), (512, 128), (32, 31)])
def test_map_tile_parameters(native, gpu_repeat, idx_repeat, T, O, gact):
    ix, gs = idx_repeat
    reads, lens, stride = reads_from(native, 800 + T, gs, 100, 50, 900, 0.0, 0.12, 0.4)
    compare_map(native, gpu_repeat, ix, reads, lens, stride, T=T, O=O, gact=gact)


@pytest.mark.parametrize("gact", ["warp", "bitpar"])
def test_map_edge_cases(native, gpu_repeat, idx_repeat, gact):
    """ragged batch: reads at the sequence boundaries (seq_lookup straddles -> unmapped), reads shorter than
    one phase (num_seeds == 0), reads of pure garbage (no votes -> key 0), a read of len == seed_len + 1."""
    ix, (a, b) = idx_repeat
    rng = np.random.default_rng(9)
    seqs = [a[:150], a[-150:], b[:150], b[-150:],                    # at the ends
            np.concatenate([a[-75:], b[:75]]),                       # straddles A.fwd|A.rev? (no: A.fwd end + junk)
            a[1000:1021], a[1000:1022], a[1000:1041], a[1000:1042],   # 21, 22, 41, 42 bases
            np.frombuffer(b"ACGT", np.uint8)[rng.integers(0, 4, 150)],  # random: no hits
            a[5000:5150][::-1].copy(),                               # reversed, not complemented
            np.full(150, ord("A"), np.uint8),                        # homopolymer
            np.tile(np.frombuffer(b"ACGTTGCA", np.uint8), 18)]       # tandem unit
    comp = np.zeros(256, np.uint8)
    comp[list(b"ACGT")] = list(b"TGCA")
    seqs.append(comp[a[-150:]][::-1].copy())                         # start of A's reverse copy
    seqs.append(comp[a[:150]][::-1].copy())                          # end of A's reverse copy
    stride = 151
    reads = np.zeros((len(seqs), stride), np.uint8)
    lens = np.zeros(len(seqs), np.uint32)
    for i, s in enumerate(seqs):
        reads[i, : len(s)] = s
        lens[i] = len(s)
    compare_map(native, gpu_repeat, ix, reads, lens, stride, gact=gact)


@pytest.mark.parametrize("gact", ["warp", "bitpar"])
def test_map_compact_output(native, gpu_repeat, idx_repeat, gact):
    """AS_F_RLE_CIGAR | AS_F_KEEP_READS: same results, the CIGAR as run bytes that print the same text, host reads
    left as sequenced (the emitter reverse-complements strand-1 records, alnmain.c:433-438)"""
    ix, gs = idx_repeat
    reads, lens, stride = reads_from(native, 640, gs, 900, 30, 700, 0.0, 0.12, 0.5)
    full, fops = gpu_repeat.map_batch(reads.copy(), lens, params=gpu_repeat.params(flags=GACT_IMPL[gact]))
    host = reads.copy()
    comp, runs = gpu_repeat.map_batch(host, lens, params=gpu_repeat.params(
        flags=GACT_IMPL[gact] | native.AS_F_RLE_CIGAR | native.AS_F_KEEP_READS))
    assert np.array_equal(host, reads)                                   # nothing copied back over the reads
    for f in full.dtype.names:
        if f != "n_cigar_op":
            assert np.array_equal(full[f], comp[f]), f
    assert (full["strand"] == 1).sum() > 100
    for i in range(len(lens)):
        n, m = int(full["n_cigar_op"][i]), int(comp["n_cigar_op"][i])
        assert (n > 0) == (m > 0)
        if n > 0:
            assert m <= n and native.cigar_text_rle(runs[i, :m]) == native.cigar_text(fops[i, :n]), i
            assert int((runs[i, :m] >> 2).sum()) == n


def check_packed_equals_full(native, m, ix, reads, lens, **knobs):
    """as_map_batch_packed (2-bit reads in, compact records + dense run bytes out) against as_map_batch on the same reads:
    every as_result field a record carries, and the run bytes equal to the AS_F_RLE_CIGAR rows read by read"""
    flags = native.AS_F_RLE_CIGAR | native.AS_F_KEEP_READS
    full, rows = m.map_batch(reads.copy(), lens, params=m.params(flags=flags))
    rec, dense = m.map_batch_packed(native.pack_reads(reads, lens), lens, params=m.params(**knobs))
    back = native.results_from_compact(rec, lens, [(o, s) for (_, o, s) in ix.mta])
    mapped = full["seq_id"] >= 0
    for f in ("off", "seq_id", "flag", "mapq", "ed", "valid"):
        assert np.array_equal(back[f], full[f]), f
    assert np.array_equal(back["strand"], full["strand"])
    assert np.array_equal(back["loc"][mapped], full["loc"][mapped])
    want_n = np.where(mapped, np.maximum(full["n_cigar_op"], 0), 0)
    n_runs = rec["n_runs"] & 0x07FFFFFF
    assert np.array_equal(n_runs, want_n.astype(np.uint32))
    assert np.array_equal((rec["n_runs"] >> 31).astype(np.int32), full["strand"])
    offs = np.concatenate([[0], np.cumsum(n_runs.astype(np.int64))])
    assert offs[-1] == dense.size
    for i in range(len(lens)):
        assert np.array_equal(dense[offs[i]:offs[i + 1]], rows[i, : want_n[i]]), i
    return rec, dense


@pytest.mark.parametrize("chunk", [0, 777])
def test_map_packed_compact_short(native, gpu_repeat, idx_repeat, chunk):
    ix, gs = idx_repeat
    reads, lens, _ = reads_from(native, 641, gs, 2500, 30, 300, 0.0, 0.1, 0.5)
    rec, dense = check_packed_equals_full(native, gpu_repeat, ix, reads, lens, chunk_reads=chunk)
    assert (rec["pos"] >> 56 == 16).sum() > 500
    h2d, d2h = gpu_repeat.last_transfer()
    assert h2d == len(lens) * (4 * ((int(lens.max()) + 15) // 16) + 4) and d2h < len(lens) * 16 + dense.size + 1024


@pytest.mark.parametrize("knobs", [dict(), dict(gact_runs=0), dict(chunk_reads=3000), dict(T=64, O=24), dict(band=-1), dict(band=16)])
def test_map_packed_runs_straight_from_gact(native, gpu_repeat, idx_repeat, knobs):
    """Reads of one tile in the short-read kernels (<= 160 bases) leave their run bytes straight from GACT's traceback events
    (GactArgs.emit_runs: no op bytes, no cleared rows, no run-length pass); the reads the one-word band cannot prove exact go
    through the retry queue (a few: warp-per-read kernel; many: bit-parallel), and with T = 64 most reads take several tiles
    and keep the op-byte path.  Run bytes and records equal as_map_batch's AS_F_RLE_CIGAR rows in every case, also with the
    direct path switched off (gact_runs = 0) -- including indel runs next to substitutions and runs longer than 63 columns."""
    ix, gs = idx_repeat
    tk = {k: knobs[k] for k in ("T", "O", "band") if k in knobs}
    kn = {k: v for k, v in knobs.items() if k not in tk}
    gpu_repeat.stats_reset()
    for seed, (lmin, lmax, emin, emax, sub) in enumerate([(30, 150, 0.0, 0.03, 0.5), (100, 160, 0.05, 0.15, 0.34), (150, 150, 0.0, 0.0, 1.0)]):
        reads, lens, _ = reads_from(native, 660 + seed, gs, 2200, lmin, lmax, emin, emax, sub)
        flags = native.AS_F_RLE_CIGAR | native.AS_F_KEEP_READS
        full, rows = gpu_repeat.map_batch(reads.copy(), lens, params=gpu_repeat.params(flags=flags, **tk))
        rec, dense = gpu_repeat.map_batch_packed(native.pack_reads(reads, lens), lens, params=gpu_repeat.params(**tk, **kn))
        mapped = full["seq_id"] >= 0
        want_n = np.where(mapped, np.maximum(full["n_cigar_op"], 0), 0)
        n_runs = (rec["n_runs"] & 0x07FFFFFF).astype(np.int64)
        assert np.array_equal(n_runs, want_n)
        assert np.array_equal(rec["ed"], full["ed"])
        offs = np.concatenate([[0], np.cumsum(n_runs)])
        assert offs[-1] == dense.size
        for i in range(len(lens)):
            assert np.array_equal(dense[offs[i]:offs[i + 1]], rows[i, : want_n[i]]), (seed, i)
    if tk == dict(band=16):     # the one-word band cannot prove all of the 5-15 % reads exact
        assert gpu_repeat.stats()["gact_queued"] > 0, "no read took GACT's retry queue: the test does not reach it"


def test_map_packed_compact_long_and_edge(native, gpu_repeat, idx_repeat):
    """reads of several tiles (packed reverse complement through the per-warp scratch row), reads with a byte outside ACGT,
    reads that fail seq_lookup, a one-read batch, a too-small runs buffer"""
    ix, (a, b) = idx_repeat
    reads, lens, _ = reads_from(native, 701, [a, b], 70, 400, 5000, 0.05, 0.15, 0.34)
    reads[3, 100] = ord("N")
    reads[5, :60] = np.frombuffer(b"acgt" * 15, np.uint8)
    check_packed_equals_full(native, gpu_repeat, ix, reads, lens)
    check_packed_equals_full(native, gpu_repeat, ix, reads, lens, chunk_reads=9)
    edge = np.zeros((4, 151), np.uint8)
    edge[0, :150] = np.concatenate([a[-75:], b[:75]])
    edge[1, :150] = np.frombuffer(b"ACGT", np.uint8)[np.random.default_rng(1).integers(0, 4, 150)]
    edge[2, :21] = a[1000:1021]
    edge[3, :150] = a[500:650]
    el = np.array([150, 150, 21, 150], np.uint32)
    check_packed_equals_full(native, gpu_repeat, ix, edge, el)
    check_packed_equals_full(native, gpu_repeat, ix, edge[3:], el[3:])
    with pytest.raises(native.AccSeqError):
        gpu_repeat.map_batch_packed(native.pack_reads(reads, lens), lens, runs_cap=10)
    pk = native.pack_reads(edge, el)
    bad = el.copy()
    bad[1] = 16 * pk.shape[1] + 1                       # a length its row cannot hold: found on the device, the call fails
    with pytest.raises(native.AccSeqError, match="longer than their row"):
        gpu_repeat.map_batch_packed(pk, bad)
    check_packed_equals_full(native, gpu_repeat, ix, edge, el)      # and the handle is fine afterwards


def test_map_batch_reports_how_far_it_got(native, gpu_uniform, idx_uniform):
    """as_gpu_last_reads_done: n after a good call (whatever the chunking); after an argument error found before any copy back
    (a length its row cannot hold) 0, with the caller's reads and results untouched"""
    ix, (g,) = idx_uniform
    reads, lens, _ = reads_from(native, 915, [g], 900, 100, 150, 0.0, 0.03, 0.5)
    for kn in (dict(), dict(chunk_reads=128)):
        gpu_uniform.map_batch(reads.copy(), lens, params=gpu_uniform.params(**kn))
        assert gpu_uniform.last_reads_done() == (len(lens), len(lens))
        gpu_uniform.map_batch(reads.copy(), lens, params=gpu_uniform.params(**kn), want_ops=False)
        assert gpu_uniform.last_reads_done() == (len(lens), len(lens))
    bad = lens.copy()
    bad[700] = reads.shape[1] + 5
    mine = reads.copy()
    with pytest.raises(native.AccSeqError):
        gpu_uniform.map_batch(mine, bad)
    assert gpu_uniform.last_reads_done() == (0, 0)
    assert np.array_equal(mine, reads)
    compare_map(native, gpu_uniform, ix, reads, lens, reads.shape[1])            # and the handle is fine afterwards


def test_map_empty_and_single(native, gpu_uniform, idx_uniform):
    ix, (g,) = idx_uniform
    res, ops = gpu_uniform.map_batch(np.zeros((0, 151), np.uint8), np.zeros(0, np.uint32))
    assert res.size == 0
    reads = np.zeros((1, 151), np.uint8)
    reads[0, :150] = g[777:927]
    exp, _ = compare_map(native, gpu_uniform, ix, reads, np.array([150], np.uint32), 151)
    assert exp["off"][0] == 777 and exp["flag"][0] == 0 and exp["ed"][0] == 0


def test_map_batch_split_invariance(native, gpu_uniform, idx_uniform):
    """results do not depend on batching (rid/batch never enter the arithmetic)."""
    ix, gs = idx_uniform
    reads, lens, stride = reads_from(native, 900, gs, 1000, 150, 150, 0.02, 0.02, 0.8)
    p = gpu_uniform.params()
    whole, wops = gpu_uniform.map_batch(reads.copy(), lens, params=p)
    parts = [gpu_uniform.map_batch(reads[s:e].copy(), lens[s:e], params=p) for s, e in ((0, 1), (1, 500), (500, 1000))]
    assert np.array_equal(whole, np.concatenate([x[0] for x in parts]))
    assert np.array_equal(wops, np.concatenate([x[1] for x in parts]))


def test_map_chunked_pipeline(native, gpu_uniform, idx_uniform):
    """as_map_batch pipelines chunks over three streams and trims the CIGAR-column copy per chunk: a ragged last
    chunk, chunk-local scratch and the trimmed 2D copy must leave results identical to the oracle's."""
    ix, gs = idx_uniform
    reads, lens, stride = reads_from(native, 950, gs, 4099, 100, 180, 0.0, 0.05, 0.6)
    compare_map(native, gpu_uniform, ix, reads, lens, stride, knobs=dict(chunk_reads=1000))
    h2d, d2h = gpu_uniform.last_transfer()
    assert h2d == 4099 * (stride + 4) and d2h < 4099 * (56 + stride + 2 * 180)
    tm = gpu_uniform.last_timing()
    assert tm["total"] > 0 and tm["part1"] > 0 and tm["gact"] > 0
    compare_map(native, gpu_uniform, ix, reads[:100], lens[:100], stride, knobs=dict(chunk_reads=5))   # 20 chunks of 5 reads


def test_truth_recovered(native, gpu_uniform, idx_uniform):
    """size-independent property: substitution-only reads land on the simulated position and strand."""
    ix, (g,) = idx_uniform
    reads, lens, pos, strand = native.sim_reads(77, g, 5000, 150, 150, 0.02, 0.02, 1.0)
    res, _ = gpu_uniform.map_batch(reads, lens, params=gpu_uniform.params(flags=native.AS_F_MYERS))
    ok = res["valid"] == 1
    assert ok.mean() > 0.99
    assert (res["off"][ok] == pos[ok]).mean() > 0.99
    assert np.array_equal(res["strand"][ok] == 1, strand[ok] == 1)
    # single-tile reads: GACT ED equals the exact NW distance
    both = ok & (res["myers"] >= 0)
    assert both.sum() > 4000 and np.array_equal(res["ed"][both], res["myers"][both])


# ---------------------------------------------------------------------------------------------------
# aligners on explicit pairs (mutils.c:94-105; edlib.cpp:694-895)
# ---------------------------------------------------------------------------------------------------
def mutate(rng, s, err, sub_frac=0.34):
    out = []
    alpha = b"ACGT"
    for c in s:
        x = rng.random()
        if x < err:
            y = rng.random()
            if y < sub_frac:
                out.append(alpha[rng.integers(0, 4)])
            elif y < sub_frac + (1 - sub_frac) / 2:
                out.append(alpha[rng.integers(0, 4)])
                out.append(c)
            else:
                continue
        else:
            out.append(c)
    return bytes(out)


def random_pairs(seed, n, lmin, lmax, emax):
    rng = np.random.default_rng(seed)
    qs, ts = [], []
    for _ in range(n):
        L = int(rng.integers(lmin, lmax + 1))
        t = bytes(np.frombuffer(b"ACGT", np.uint8)[rng.integers(0, 4, L)])
        q = mutate(rng, t, rng.random() * emax)
        qs.append(q)
        ts.append(t)
    return qs, ts


@pytest.mark.parametrize("gact", ["warp", "bitpar"])
@pytest.mark.parametrize("T,O,lmax", [(320, 120, 300), (320, 120, 2500), (64, 20, 700), (32, 8, 300), (512, 200, 3000)])
def test_gact_pairs(orc, gpu_uniform, T, O, lmax, gact):
    qs, ts = random_pairs(T + lmax, 120, 1, lmax, 0.2)
    qs += [b"A", b"ACGT", b"A" * 50, b"ACGT" * 30]
    ts += [b"C", b"ACGT", b"A" * 80, b"TGCA" * 30]
    qs += [b"", b"ACG", b""]
    ts += [b"ACGTA", b"", b""]
    score, ops = gpu_uniform.cigar_align_batch(qs, ts, gpu_uniform.params(T=T, O=O, flags=GACT_IMPL[gact]))
    for i, (q, t) in enumerate(zip(qs, ts)):
        es, eops, _ = orc.gact(np.frombuffer(q, np.uint8), np.frombuffer(t, np.uint8), T, O)
        assert es == score[i], (i, len(q), len(t))
        assert np.array_equal(eops, ops[i]), i
        if es >= 0:   # every path consumes exactly the two sequences
            assert int(np.isin(ops[i], (0, 1, 3)).sum()) == len(q) and int(np.isin(ops[i], (0, 2, 3)).sum()) == len(t)


def tie_rich_pairs(seed, n, lmax):
    """low-complexity pairs (homopolymer runs, short tandem repeats, shifted copies): many equal-cost paths, so the
    diag > up > left priority and the banded pass's exactness argument are both exercised"""
    rng = np.random.default_rng(seed)
    qs, ts = [], []
    for _ in range(n):
        L = int(rng.integers(8, lmax + 1))
        unit = bytes(np.frombuffer(b"ACGT", np.uint8)[rng.integers(0, 4, int(rng.integers(1, 5)))])
        t = (unit * (L // len(unit) + 1))[:L]
        kind = rng.integers(0, 3)
        if kind == 0:
            q = mutate(rng, t, 0.05)
        elif kind == 1:                       # the window starts a few bases off, as a voted locus does (alnmain.c:427)
            sh = int(rng.integers(1, 14))
            q = mutate(rng, t[sh:] + t[:sh][::-1], 0.03)
        else:
            q = t[: L // 2] + mutate(rng, t[L // 2:], 0.3)
        qs.append(q)
        ts.append(t)
    return qs, ts


@pytest.mark.parametrize("band", [0, 3, 16, 48, 128, -1])
@pytest.mark.parametrize("T,O,lmax", [(320, 120, 300), (320, 120, 2500), (64, 20, 700), (32, 8, 300), (512, 200, 3000)])
def test_gact_banded_pass_is_exact(orc, gpu_uniform, T, O, lmax, band):
    """bit-parallel kernel: banded first pass + full-matrix second pass give the oracle's ops for every half-width"""
    qs, ts = random_pairs(7 * T + lmax, 90, 1, lmax, 0.25)
    q2, t2 = tie_rich_pairs(T + lmax, 60, min(lmax, 900))
    qs, ts = qs + q2 + [b"", b"ACG", b"A" * 40], ts + t2 + [b"ACGTA", b"", b"A" * 90]
    gpu_uniform.stats_reset()
    score, ops = gpu_uniform.cigar_align_batch(qs, ts, gpu_uniform.params(T=T, O=O, flags=GACT_IMPL["bitpar"], band=band))
    st = gpu_uniform.stats()
    for i, (q, t) in enumerate(zip(qs, ts)):
        es, eops, _ = orc.gact(np.frombuffer(q, np.uint8), np.frombuffer(t, np.uint8), T, O)
        assert es == score[i], (i, len(q), len(t))
        assert np.array_equal(eops, ops[i]), i
    if band == 3:
        assert st["gact_redone"] > 0          # the narrow band cannot prove most of these: the second pass ran
    if band == -1:
        assert st["gact_redone"] == 0


@pytest.mark.parametrize("gact", ["warp", "bitpar"])
def test_gact_single_tile_is_exact_nw(orc, gpu_uniform, gact):
    qs, ts = random_pairs(3, 200, 20, 320, 0.25)
    score, _ = gpu_uniform.cigar_align_batch(qs, ts, gpu_uniform.params(flags=GACT_IMPL[gact]))
    dist = gpu_uniform.edit_distance_batch(qs, ts)
    assert np.array_equal(score, dist)


def test_edlib_known_answer(gpu_uniform):
    """edlib/ed_test.c:14-19: two 100-mers differing at 2 positions -> 2 / 100M."""
    rng = np.random.default_rng(0)
    t = np.frombuffer(b"ACGT", np.uint8)[rng.integers(0, 4, 100)].copy()
    q = t.copy()
    q[10] = ord("A") if q[10] != ord("A") else ord("C")
    q[60] = ord("G") if q[60] != ord("G") else ord("T")
    d = gpu_uniform.edit_distance_batch([q.tobytes()], [t.tobytes()], [5000000])
    assert d[0] == 2
    s, ops = gpu_uniform.cigar_align_batch([q.tobytes()], [t.tobytes()])
    assert s[0] == 2 and len(ops[0]) == 100 and set(ops[0].tolist()) <= {0, 3}


@pytest.mark.parametrize("lmax,emax", [(256, 0.3), (2000, 0.15), (12000, 0.12)])
def test_myers_pairs(orc, gpu_uniform, lmax, emax):
    n = 150 if lmax <= 2000 else 24
    qs, ts = random_pairs(lmax, n, 1, lmax, emax)
    rng = np.random.default_rng(lmax)
    # unbounded
    d = gpu_uniform.edit_distance_batch(qs, ts)
    exp = np.array([orc.edit_distance(np.frombuffer(q, np.uint8), np.frombuffer(t, np.uint8), -1) for q, t in zip(qs, ts)])
    assert np.array_equal(d, exp)
    # bounded: k around the true distance exercises the -1 verdict and the band
    ks = np.maximum(0, exp + rng.integers(-5, 6, len(qs))).astype(np.int32)
    d = gpu_uniform.edit_distance_batch(qs, ts, ks)
    assert np.array_equal(d, np.where(exp > ks, -1, exp))


def test_edit_alignment_path(native, orc, gpu_uniform):
    """as_edit_alignment_batch = edlib's alignment path (EDLIB_MODE_NW, EDLIB_TASK_PATH; edlib.cpp:909-1108): distance and
    every op equal (1) the oracle's restatement on seeded pairs -- unrelated, mutated, tie-rich low-complexity, shifted, up to
    256-base queries against targets of up to 2000 bases, with and without k -- and (2) the golden paths the reference's own
    compiled edlib produced (tests/golden/edlib_paths.npz); the ops re-read as an alignment consume both sequences and cost
    the distance; out-of-range pairs are refused"""
    from test_oracle_vs_ref import _pairs_for_alignment, edlib_golden_pairs
    rng = np.random.default_rng(77)
    pairs = list(_pairs_for_alignment(rng, 600))
    acgt = np.frombuffer(b"ACGT", np.uint8)
    for L in (1, 2, 63, 64, 65, 128, 129, 255, 256):                     # block boundaries of the query
        q = acgt[rng.integers(0, 4, L)]
        pairs.append((q, np.concatenate([q[: L // 2], acgt[rng.integers(0, 4, 3)], q[L // 2:]])))
        pairs.append((q, acgt[rng.integers(0, 4, 2000)]))
    qs, ts = [p[0].tobytes() for p in pairs], [p[1].tobytes() for p in pairs]
    for ks in (None, np.full(len(pairs), 9, np.int32)):
        dist, ops = gpu_uniform.edit_alignment_batch(qs, ts, ks)
        for i, (q, t) in enumerate(pairs):
            d, want = orc.edit_alignment(q, t, -1 if ks is None else int(ks[i]))
            assert dist[i] == d and np.array_equal(ops[i], want), (i, q.tobytes(), t.tobytes())
            if d >= 0:
                o = ops[i]
                assert (o != 2).sum() == q.size and (o != 1).sum() == t.size and (o != 0).sum() == d
    gold = list(edlib_golden_pairs())
    dist, ops = gpu_uniform.edit_alignment_batch([g[0].tobytes() for g in gold], [g[1].tobytes() for g in gold],
                                                 np.array([g[2] for g in gold], np.int32))
    for i, g in enumerate(gold):
        assert dist[i] == g[3] and np.array_equal(ops[i], g[4]), i
    with pytest.raises(native.AccSeqError, match="beyond the traceback"):
        gpu_uniform.edit_alignment_batch([b"A" * 257], [b"A" * 10])
    with pytest.raises(native.AccSeqError, match="beyond the traceback"):
        gpu_uniform.edit_alignment_batch([b"A" * 256], [b"A" * 12000])
    dist, ops = gpu_uniform.edit_alignment_batch(qs[:3], ts[:3])            # and the handle is fine afterwards
    assert all(np.array_equal(ops[i], orc.edit_alignment(pairs[i][0], pairs[i][1], -1)[1]) for i in range(3))


# ---------------------------------------------------------------------------------------------------
# BASELINE config 2 at its full size (4.6 Mbp genome, 1 M reads of 150 bp at 2 %): size-independent properties, plus the
# oracle on a random sample of the same batch
# ---------------------------------------------------------------------------------------------------
@pytest.mark.gpu
def test_c2_full_size_properties(native, orc):
    g = native.sim_genome(11, 4_600_000)
    n = 1_000_000
    reads, lens, pos, strand = native.sim_reads(5, g, n, 150, 150, 0.02, 0.02, 0.8)     # bench.py's c2 mix
    stride = reads.shape[1]
    m = native.GpuMapper.build([g], 0)
    try:
        p = m.params(flags=native.AS_F_MYERS)
        r_all = reads.copy()
        res, ops = m.map_batch(r_all, lens, params=p)
        ok = res["valid"] == 1
        # (1) the simulated position and strand come back
        assert ok.mean() > 0.99
        near = np.abs(res["off"][ok].astype(np.int64) - pos[ok].astype(np.int64)) <= 50   # indels move the voted start
        assert near.mean() > 0.999
        assert ((res["strand"][ok] == 1) == (strand[ok] == 1)).mean() > 0.999
        # (2) CIGAR consistency of every valid read: query-consuming ops (=, I, X) add up to the read, ED = non-'=' ops,
        #     and one tile is exact NW (reads are shorter than T): ED equals the Myers distance
        nops = res["n_cigar_op"].astype(np.int64)
        live = np.arange(ops.shape[1])[None, :] < np.maximum(nops, 0)[:, None]
        qcons = (((ops == 0) | (ops == 1) | (ops == 3)) & live).sum(axis=1)
        nonmatch = ((ops != 0) & live).sum(axis=1)
        assert np.array_equal(qcons[ok], lens[ok].astype(np.int64))
        assert np.array_equal(nonmatch[ok], res["ed"][ok].astype(np.int64))
        both = ok & (res["myers"] >= 0)
        assert both.sum() > 0.8 * n and np.array_equal(res["ed"][both], res["myers"][both])
        # (3) strand-1 reads were reverse-complemented in place, the others left alone (alnmain.c:433-438)
        s1 = ok & (res["strand"] == 1)
        assert np.array_equal(r_all[~s1], reads[~s1])
        comp = np.zeros(256, np.uint8)
        comp[list(b"ACGT")] = list(b"TGCA")
        k = np.nonzero(s1)[0][:20000]
        assert np.array_equal(r_all[k, :150], comp[reads[k, :150]][:, ::-1])
        # (4) batch-split invariance: two unequal halves give the rows of the whole batch
        cut = 400_003
        for lo, hi in ((0, cut), (cut, n)):
            r_part = reads[lo:hi].copy()
            res_p, ops_p = m.map_batch(r_part, lens[lo:hi], params=p)
            assert np.array_equal(res_p, res[lo:hi])
            assert np.array_equal(ops_p * live[lo:hi], ops[lo:hi] * live[lo:hi])
            assert np.array_equal(r_part, r_all[lo:hi])
        # (5) compact output: same records, run lengths add up to the columns
        r_c = reads.copy()
        res_c, runs = m.map_batch(r_c, lens, params=m.params(flags=native.AS_F_MYERS | native.AS_F_RLE_CIGAR | native.AS_F_KEEP_READS))
        assert np.array_equal(r_c, reads)
        for f in res.dtype.names:
            if f != "n_cigar_op":
                assert np.array_equal(res_c[f], res[f]), f
        nr = res_c["n_cigar_op"].astype(np.int64)
        lr = np.arange(runs.shape[1])[None, :] < np.maximum(nr, 0)[:, None]
        assert np.array_equal(((runs >> 2).astype(np.int64) * lr).sum(axis=1)[ok], nops[ok])
        # (6) the oracle on a random sample of the batch: bit-identical records and CIGAR columns
        ix = orc.OracleIndex.build(["chr1"], [g.tobytes()])
        try:
            pick = np.sort(np.random.default_rng(3).choice(n, 3000, replace=False))
            r_cpu = reads[pick].copy()
            exp, exp_ops, _ = ix.map_batch(r_cpu, stride, lens[pick], seed_len=20, thres=300, T=320, O=120, myers=True)
            for f in FIELDS:
                assert np.array_equal(exp[f], res[f][pick]), f
            assert np.array_equal(exp["n_ops"], res["n_cigar_op"][pick])
            for i, r in enumerate(pick):
                k = int(exp["n_ops"][i])
                assert np.array_equal(exp_ops[i, :k], ops[r, :k])
            assert np.array_equal(r_cpu, r_all[pick])
        finally:
            ix.close()
    finally:
        m.close()
