"""The oracle -- and, where it runs in test time, the compiled reference itself -- against the CUDA path at the FULL sizes of
BASELINE.json's configs, so that the code only large indexes reach (deep table at its default depth 15/16, SA + context
table larger than L2, entry contexts on by default, 40-bit SA rows, super blocks, 64-bit backward search, partitioned suffix
sort) is compared with the reference's algorithm and not just with itself.

  c1  1 Mbp, 10 k x 150 bp @ 1 %      the reference's programs end to end: accidx_ref -> accaln_ref  vs  accidx_b200 -> accaln_b200,
                                      index files byte for byte, SAM line for line (alnmain.c:277-551); the oracle on the same reads
  c3  100 Mbp, 10 kbp @ 10 %          index built on the GPU, exported in the reference's layouts (as_gpu_export), handed to the
                                      oracle (OracleIndex.from_arrays): every record field and every CIGAR column of a read sample;
                                      and accaln_ref itself on the files accidx_b200 wrote
  c5  1-50 kbp @ 5-15 %               the same on mixed-length long reads (100 Mbp genome always, 3.1 Gbp when the box has the memory)
  c4  3.1 Gbp, 150 bp @ 2 %           the same at human scale (needs a whole B200 and ~80 GB of host memory; skipped otherwise)

The size-independent property tests of c3 / c4 (truth recovery, CIGAR consistency, alternative PART 1 / GACT paths) live here too
and share the indexes."""
import os
import subprocess

import numpy as np
import pytest

import oracle_lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ACCALN = os.path.join(ROOT, "longreadmapper_b200", "accaln_b200")
ACCIDX = os.path.join(ROOT, "longreadmapper_b200", "accidx_b200")
FIELDS = ("loc", "off", "seq_id", "strand", "flag", "mapq", "ed", "valid", "myers", "phases", "from_outer")
INDEX_EXTS = (".mta", ".cat", ".cat.sa5", ".cat.mfi", ".cat.lch")

needs_ref = pytest.mark.skipif(not oracle_lib.have_ref(), reason="oracle/_ref not built (no /root/reference here)")


def write_fasta(path, names, seqs):
    with open(path, "wb") as f:
        for n, s in zip(names, seqs):
            f.write(b">" + n.encode() + b"\n")
            f.write(s.tobytes() if isinstance(s, np.ndarray) else bytes(s))
            f.write(b"\n")


def write_fastq(path, reads, lens, prefix="r"):
    with open(path, "wb") as f:
        for i in range(len(lens)):
            L = int(lens[i])
            f.write(b"@%s%d\n" % (prefix.encode(), i) + reads[i, :L].tobytes() + b"\n+\n" + b"I" * L + b"\n")


def mem_available_gb():
    try:
        return int(next(l for l in open("/proc/meminfo") if l.startswith("MemAvailable")).split()[1]) / 2**20
    except (OSError, StopIteration, ValueError):
        return 0.0


def sam_records(text):
    return [l for l in text.splitlines() if l and l[0] != "@"]


def assert_sam_equal(got_text, want_text, seed_len=20):
    """accaln_b200's SAM against accaln_ref's, line for line.  Exclusions: the time-based @RG id, and
    records where seq_lookup failed or len <= seed_len -- the reference prints uninitialised memory there."""
    gh = [l for l in got_text.splitlines() if l.startswith("@") and not l.startswith("@RG")]
    wh = [l for l in want_text.splitlines() if l.startswith("@") and not l.startswith("@RG")]
    assert gh == wh
    got, want = sam_records(got_text), sam_records(want_text)
    assert len(got) == len(want)
    same = undefined = 0
    for g, w in zip(got, want):
        gf = g.split("\t")
        assert gf[0] == w.split("\t", 1)[0]
        if gf[2] == "*" or len(gf[9]) <= seed_len:
            undefined += 1
            assert gf[2] != "*" or (gf[1], gf[4]) == ("4", "0")
            continue
        assert g == w
        same += 1
    return same, undefined


def run(cmd, **env):
    p = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, env=dict(os.environ, ACCSEQ_RG_ID="1", **env),
                       timeout=1800)
    assert p.returncode == 0, (cmd, p.stderr.decode()[-2000:])
    return p.stdout.decode()


def assert_oracle_equal(native, m, ox, reads, lens, params=None, myers=True):
    """every as_result field, every CIGAR column and the in-place reverse complement, GPU vs oracle, on the same reads"""
    stride = reads.shape[1]
    r_cpu, r_gpu = reads.copy(), reads.copy()
    exp, exp_ops, _ = ox.map_batch(r_cpu, stride, lens, seed_len=20, thres=300, T=320, O=120, myers=myers)
    p = params or m.params(flags=native.AS_F_MYERS if myers else 0)
    got, got_ops = m.map_batch(r_gpu, lens, params=p)
    for f in FIELDS:
        if f == "myers" and not myers:
            continue
        bad = np.nonzero(exp[f] != got[f])[0]
        assert bad.size == 0, f"{f}: {bad.size} reads differ, first {bad[0]}: oracle {exp[f][bad[0]]} gpu {got[f][bad[0]]}"
    assert np.array_equal(exp["n_ops"], got["n_cigar_op"])
    for i in range(len(lens)):
        k = int(exp["n_ops"][i])
        assert np.array_equal(exp_ops[i, :k], got_ops[i, :k]), f"CIGAR columns differ on read {i}"
    assert np.array_equal(r_cpu, r_gpu), "in-place reverse complement differs"
    return exp


def cigar_consistent(res, ops, lens):
    ok = res["valid"] == 1
    nops = res["n_cigar_op"].astype(np.int64)
    live = np.arange(ops.shape[1])[None, :] < np.maximum(nops, 0)[:, None]
    qcons = (((ops == 0) | (ops == 1) | (ops == 3)) & live).sum(axis=1)
    nonmatch = ((ops != 0) & live).sum(axis=1)
    assert np.array_equal(qcons[ok], lens[ok].astype(np.int64))
    assert np.array_equal(nonmatch[ok], res["ed"][ok].astype(np.int64))
    return ok, live


# ---------------------------------------------------------------------------------------------------
# c1: BASELINE configs[0], "run on the reference's CPU build"
# ---------------------------------------------------------------------------------------------------
@pytest.fixture(scope="module")
def c1(native, tmp_path_factory):
    d = tmp_path_factory.mktemp("c1")
    g = native.sim_genome(101, 1_000_000)
    reads, lens, pos, strand = native.sim_reads(102, g, 10_000, 150, 150, 0.01, 0.01, 0.9)
    fa, fq = str(d / "c1.fa"), str(d / "c1.fq")
    write_fasta(fa, ["chr1"], [g])
    write_fastq(fq, reads, lens)
    out = dict(g=g, reads=reads, lens=lens, pos=pos, strand=strand, fa=fa, fq=fq, dir=d)
    if oracle_lib.have_ref():
        oracle_lib.run_ref_accidx(fa)                       # the reference's accidx, pSAscan included
        out["ref_sam"] = oracle_lib.run_ref_accaln(fa, fq, batch=1_000_000)
    return out


@needs_ref
def test_c1_oracle_index_files_equal_reference(c1, orc, tmp_path):
    """at c1's size too, the oracle writes the bytes the reference's accidx (pSAscan suffix sort) writes"""
    ix = orc.OracleIndex.build(["chr1"], [c1["g"].tobytes()])
    fa2 = str(tmp_path / "o.fa")
    ix.write(fa2)
    ix.close()
    for ext in INDEX_EXTS:
        assert open(fa2 + ext, "rb").read() == open(c1["fa"] + ext, "rb").read(), ext


@needs_ref
def test_c1_oracle_records_equal_reference_sam(c1, orc):
    """all 10 000 reads of c1: the oracle's records against the SAM the compiled reference printed"""
    ix = orc.OracleIndex.load(c1["fa"])                     # the files accidx_ref wrote
    try:
        r = c1["reads"].copy()
        res, ops, _ = ix.map_batch(r, r.shape[1], c1["lens"], myers=False)
        recs = oracle_lib.parse_sam(c1["ref_sam"])
        assert len(recs) == 10_000
        n_cmp = 0
        for i, rec in enumerate(recs):
            if res["seq_id"][i] < 0:
                continue                                    # seq_lookup failed: reference output undefined
            n_cmp += 1
            assert (rec["flag"], rec["rname"], rec["pos"], rec["mapq"], rec["ed"]) == \
                (int(res["flag"][i]), "chr1", int(res["off"][i]) + 1, int(res["mapq"][i]), int(res["ed"][i])), i
            assert rec["cigar"] == oracle_lib.cigar_text(ops[i, : res["n_ops"][i]]), i
            assert rec["seq"].encode() == r[i, :150].tobytes(), i
        assert n_cmp > 9_900
        ok = res["valid"] == 1
        near = np.abs(res["off"][ok].astype(np.int64) - c1["pos"][ok].astype(np.int64)) <= 50   # indels shift the voted start
        assert ok.mean() > 0.99 and near.mean() > 0.999
    finally:
        ix.close()


@pytest.mark.gpu
@needs_ref
def test_c1_driver_programs_equal_reference(c1, native, tmp_path):
    """accidx_b200 writes accidx_ref's five files byte for byte; accaln_b200 prints accaln_ref's SAM line for line"""
    fa2 = str(tmp_path / "b.fa")
    write_fasta(fa2, ["chr1"], [c1["g"]])
    run([ACCIDX, fa2])
    for ext in INDEX_EXTS:
        assert open(fa2 + ext, "rb").read() == open(c1["fa"] + ext, "rb").read(), ext
    for batch in ("1000000", "3333"):
        sam = run([ACCALN, fa2, c1["fq"], batch, "20", "300"])
        same, undefined = assert_sam_equal(sam, c1["ref_sam"])
        assert same > 9_900 and undefined < 100
    sam = run([ACCALN, fa2, c1["fq"], "1000000", "20", "300"], ACCALN_FULL_COPY="1")     # the reference's buffer layout
    assert assert_sam_equal(sam, c1["ref_sam"])[0] > 9_900


# ---------------------------------------------------------------------------------------------------
# c3 / c5 on the 100 Mbp genome
# ---------------------------------------------------------------------------------------------------
@pytest.fixture(scope="module")
def c3_index(native):
    g = native.sim_genome(11, 100_000_000)
    cat, mta = native.cat_image([g])
    m = native.GpuMapper.build_cat(cat, mta, 0)
    ox = oracle_lib.OracleIndex.from_gpu_export(m, cat, mta, names=["chr1"])
    yield dict(g=g, cat=cat, mta=mta, m=m, ox=ox)
    ox.close()
    m.close()


@pytest.mark.gpu
def test_c3_oracle_sample(native, c3_index):
    """200 reads of 10 kbp at 10 % error (BASELINE configs[2]) through the oracle on the GPU-built 100 Mbp index: the default
    depth-15 deep table, entry contexts (on by default above L2 size), several GACT tiles per read"""
    c = c3_index
    reads, lens, pos, strand = native.sim_reads(9, c["g"], 200, 10_000, 10_000, 0.10, 0.10, 1 / 3)
    exp = assert_oracle_equal(native, c["m"], c["ox"], reads, lens)
    ok = exp["valid"] == 1
    assert ok.mean() > 0.98
    # and the other ways through PART 1: all equal to the oracle, not only to one another
    for kw in (dict(flags=native.AS_F_NO_LC2), dict(flags=native.AS_F_WIDE_INDEX), dict(vote_ectx=0), dict(gact_order=0)):
        flags = kw.pop("flags", 0)
        assert_oracle_equal(native, c["m"], c["ox"], reads[:60], lens[:60],
                            params=c["m"].params(flags=native.AS_F_MYERS | flags, **kw))


@pytest.mark.gpu
def test_c3_oracle_short_reads(native, c3_index):
    """2 000 reads of 150 bp at 2 % on the same index (the c4 read shape on a large-index code path)"""
    c = c3_index
    reads, lens, _, _ = native.sim_reads(19, c["g"], 2000, 150, 150, 0.02, 0.02, 0.8)
    assert_oracle_equal(native, c["m"], c["ox"], reads, lens)


@pytest.mark.gpu
def test_c5_oracle_sample_100mbp(native, c3_index):
    """BASELINE configs[4]'s read mix -- lengths uniform in 1-50 kbp, 5-15 % error -- on the 100 Mbp genome: 100 reads through
    the oracle.  Reports how many land within 50 bp of the simulated start: indel-rich long reads shift the voted start
    (best.key is the minimum key of the winning 16-bp bucket, alnmain.c:360-368), in the reference's algorithm as here."""
    c = c3_index
    reads, lens, pos, strand = native.sim_reads(29, c["g"], 100, 1_000, 50_000, 0.05, 0.15, 1 / 3)
    assert int(lens.max()) > 40_000 and int(lens.min()) < 5_000
    exp = assert_oracle_equal(native, c["m"], c["ox"], reads, lens, myers=False)
    ok = exp["valid"] == 1
    near = np.abs(exp["off"][ok].astype(np.int64) - pos[ok].astype(np.int64)) <= 50
    print(f"c5 sample: valid {ok.mean():.3f}, within 50 bp of truth (oracle == gpu) {near.mean():.3f}")
    assert ok.mean() > 0.95


@pytest.mark.gpu
@needs_ref
def test_c3_reference_program(native, c3_index, tmp_path):
    """accaln_ref ITSELF on c3: accidx_b200 writes the 100 Mbp index files (2.4 GB), the compiled reference loads them and maps
    300 reads of 10 kbp; accaln_b200 on the same files prints the same SAM"""
    c = c3_index
    fa, fq = str(tmp_path / "c3.fa"), str(tmp_path / "c3.fq")
    write_fasta(fa, ["chr1"], [c["g"]])
    run([ACCIDX, fa])
    assert open(fa + ".cat", "rb").read() == c["cat"].tobytes()
    assert os.path.getsize(fa + ".cat.sa5") == 5 * c["cat"].size                       # 1 000 000 005 B
    reads, lens, _, _ = native.sim_reads(39, c["g"], 300, 10_000, 10_000, 0.10, 0.10, 1 / 3)
    write_fastq(fq, reads, lens)
    want = oracle_lib.run_ref_accaln(fa, fq, batch=1000)
    got = run([ACCALN, fa, fq, "1000", "20", "300"])
    same, undefined = assert_sam_equal(got, want)
    assert same >= 295
    for ext in INDEX_EXTS:
        os.remove(fa + ext)


@pytest.mark.gpu
def test_c3_scale_properties(native, c3_index):
    """size-independent properties on 3 000 reads of c3 (truth recovery, CIGAR consistency, batch-split invariance) and
    equality of every alternative PART 1 / GACT path with the default one on ALL of them"""
    c = c3_index
    m, g = c["m"], c["g"]
    n = 3000
    reads, lens, pos, strand = native.sim_reads(9, g, n, 10_000, 10_000, 0.10, 0.10, 1 / 3)
    p = m.params()
    r0 = reads.copy()
    res, ops = m.map_batch(r0, lens, params=p)
    ok, live = cigar_consistent(res, ops, lens)
    assert ok.mean() > 0.99
    near = np.abs(res["off"][ok].astype(np.int64) - pos[ok].astype(np.int64)) <= 200
    assert near.mean() > 0.99
    assert ((res["strand"][ok] == 1) == (strand[ok] == 1)).mean() > 0.99
    for kw in (dict(vote_ectx=0), dict(flags=native.AS_F_NO_LC2), dict(gact_order=0)):
        r1 = reads.copy()
        res1, ops1 = m.map_batch(r1, lens, params=m.params(**kw))
        assert np.array_equal(res1, res), kw
        assert np.array_equal(ops1 * live, ops * live), kw
        assert np.array_equal(r1, r0)
    cut = 1234
    res_a, ops_a = m.map_batch(reads[:cut].copy(), lens[:cut], params=p)
    res_b, ops_b = m.map_batch(reads[cut:].copy(), lens[cut:], params=p)
    assert np.array_equal(np.concatenate([res_a, res_b]), res)
    assert np.array_equal(np.concatenate([ops_a, ops_b]) * live, ops * live)


# ---------------------------------------------------------------------------------------------------
# c4 / c5 at human scale
# ---------------------------------------------------------------------------------------------------
@pytest.fixture(scope="module")
def c4_index(native):
    import torch
    free_b, _ = torch.cuda.mem_get_info(0)
    avail = mem_available_gb()
    if free_b < 150 * 2**30 or avail < 48:
        pytest.skip(f"human scale needs 150 GB of free HBM and 48 GB of free host memory (have {free_b / 2**30:.0f} / {avail:.0f})")
    g = native.sim_genome(11, 3_100_000_000)
    cat, mta = native.cat_image([g])
    m = native.GpuMapper.build_cat(cat, mta, 0)
    yield dict(g=g, cat=cat, mta=mta, m=m)
    m.close()


@pytest.mark.gpu
def test_c4_human_scale_properties(native, c4_index):
    """3.1 Gbp: 6.2 G rows, partitioned suffix sort, 40-bit SA, super blocks, 64-bit backward search, deep table of 4^16
    entries; 200 000 reads of 150 bp at 2 %: truth recovery, CIGAR consistency, ED = Myers, and the 12-mer-table walk and the
    no-entry-context path equal to the default path"""
    m, g = c4_index["m"], c4_index["g"]
    n = 200_000
    reads, lens, pos, strand = native.sim_reads(13, g, n, 150, 150, 0.02, 0.02, 0.8)
    p = m.params(flags=native.AS_F_MYERS)
    r0 = reads.copy()
    res, ops = m.map_batch(r0, lens, params=p)
    ok, live = cigar_consistent(res, ops, lens)
    assert ok.mean() > 0.99
    near = np.abs(res["off"][ok].astype(np.int64) - pos[ok].astype(np.int64)) <= 50
    assert near.mean() > 0.995
    assert ((res["strand"][ok] == 1) == (strand[ok] == 1)).mean() > 0.995
    both = ok & (res["myers"] >= 0)
    assert both.sum() > 0.8 * n and np.array_equal(res["ed"][both], res["myers"][both])
    for kw in (dict(vote_ectx=0), dict(flags=native.AS_F_NO_LC2)):
        flags = kw.pop("flags", 0)
        r1 = reads.copy()
        res1, ops1 = m.map_batch(r1, lens, params=m.params(flags=native.AS_F_MYERS | flags, **kw))
        assert np.array_equal(res1, res), kw
        assert np.array_equal(ops1 * live, ops * live), kw
        assert np.array_equal(r1, r0)


@pytest.mark.gpu
def test_c4_c5_oracle_sample_human_scale(native, c4_index):
    """the oracle at human scale: the GPU-built 6.2 G-row index exported in the reference's layouts (C, O, BWT, lc, the raw
    .sa5 rows: ~50 GB of host memory) and handed to orc_map_batch -- 2 000 reads of c4 (150 bp at 2 %) and 40 reads of c5
    (1-50 kbp at 5-15 %), every record field and CIGAR column"""
    avail = mem_available_gb()
    if avail < 75:
        pytest.skip(f"the exported human-scale index needs ~60 GB of host memory beside the genome (MemAvailable {avail:.0f} GB)")
    c = c4_index
    ox = oracle_lib.OracleIndex.from_gpu_export(c["m"], c["cat"], c["mta"], names=["chr1"], sa_entry_bytes=5)
    try:
        reads, lens, _, _ = native.sim_reads(14, c["g"], 2000, 150, 150, 0.02, 0.02, 0.8)
        assert_oracle_equal(native, c["m"], ox, reads, lens)
        assert_oracle_equal(native, c["m"], ox, reads[:500], lens[:500], params=c["m"].params(flags=native.AS_F_MYERS | native.AS_F_NO_LC2))
        reads, lens, pos, _ = native.sim_reads(15, c["g"], 40, 1_000, 50_000, 0.05, 0.15, 1 / 3)
        exp = assert_oracle_equal(native, c["m"], ox, reads, lens, myers=False)
        ok = exp["valid"] == 1
        near = np.abs(exp["off"][ok].astype(np.int64) - pos[ok].astype(np.int64)) <= 50
        print(f"c5 human-scale sample: valid {ok.mean():.3f}, within 50 bp of truth (oracle == gpu) {near.mean():.3f}")
    finally:
        ox.close()


@pytest.mark.gpu
def test_c4_sampled_sa_human_scale(native, c4_index):
    """SURVEY 8(f)-4 at the size it is meant for: a second human-scale handle that keeps SA[8 i] only (3.9 GB instead of the
    31 GB suffix array + 50 GB SA + context table) maps 200 000 reads of c4 to the same records and CIGARs as the full one"""
    import torch
    m, g = c4_index["m"], c4_index["g"]
    free0, _ = torch.cuda.mem_get_info(0)
    if free0 < 75 * 2**30:
        pytest.skip("not enough free HBM for a second human-scale index build")
    reads, lens, _, _ = native.sim_reads(16, g, 200_000, 150, 150, 0.02, 0.02, 0.8)
    res, ops = m.map_batch(reads.copy(), lens, params=m.params())
    samp = native.GpuMapper.build_cat(c4_index["cat"], c4_index["mta"], 0, opts=native.open_opts(sampled_sa=8))
    try:
        free1, _ = torch.cuda.mem_get_info(0)
        used_gb = (free0 - free1) / 2**30
        print(f"human-scale index with sampled_sa=8: {used_gb:.1f} GB of HBM")
        assert used_gb < 48
        res2, ops2 = samp.map_batch(reads.copy(), lens, params=samp.params())
        assert np.array_equal(res, res2)
        live = np.arange(ops.shape[1])[None, :] < np.maximum(res["n_cigar_op"], 0)[:, None]
        assert np.array_equal(ops * live, ops2 * live)
    finally:
        samp.close()
