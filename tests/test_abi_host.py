"""CPU tests (-m "not gpu") of the drop-in boundary and the host helpers: the C-ABI libraries load and export
every symbol include/*.h declares, compute entries fail loudly without a GPU (no CPU fallback), and the plain-C
host side (index files, FASTQ ingest, SAM emit) agrees with the files/records the compiled reference wrote."""
import ctypes as C
import gzip
import os
import re

import numpy as np
import pytest

import oracle_lib
from test_oracle_vs_ref import GOLD, read_fasta_gz, read_fastq_gz

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared(header):
    text = open(os.path.join(ROOT, "include", header)).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(as_[a-z0-9_]+)\s*\(", text)))


def test_cuda_library_exports_every_declared_symbol(native):
    L = native.cuda_lib()
    names = declared("accseq_b200.h")
    assert len(names) >= 24
    for n in names:
        assert hasattr(L, n), f"libaccseq_b200.so does not export {n}"


def test_host_library_exports_every_declared_symbol(native):
    L = native.host_lib()
    names = declared("accseq_host.h")
    assert len(names) >= 15
    for n in names:
        assert hasattr(L, n), f"libaccseq_host.so does not export {n}"


def test_gact_header_names_are_defined_by_the_oracle_glue():
    """the reference's own sources compile against include/gact/gact.h (oracle/Makefile); the names it needs"""
    text = open(os.path.join(ROOT, "include", "gact", "gact.h")).read()
    for n in ("mmstring", "cigar", "ms_borrow", "simple_gact", "parse_cigar", "n_cigar_op", "score"):
        assert n in text


def test_no_cpu_fallback(native):
    """without a device every compute entry fails with AS_ENODEV and says why"""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    L = native.cuda_lib()
    assert native.device_count() == 0
    g = native.sim_genome(1, 2000)
    cat, mta = native.cat_image([g])
    arr = (native.capi.AsMta * 1)(native.capi.AsMta(mta[0][0], mta[0][1]))
    h = C.c_void_p()
    rc = L.as_gpu_build(C.byref(h), 0, cat.ctypes.data, cat.size, C.cast(arr, C.c_void_p), 1, 12)
    assert rc == -19 and not h.value
    assert b"device" in L.as_last_error().lower() or b"cuda" in L.as_last_error().lower()
    with pytest.raises(native.AccSeqError):
        native.GpuMapper.build([g], 0)


def test_pack_reads_layout(native):
    """as_pack_reads: base j of a read at bits [2(j&15), 2(j&15)+1] of word j>>4, A0 C1 G2 T3 for either case, other bytes the
    code ((c>>1)&3)^((c>>2)&1) the device gives them, bits past the length zero"""
    g = native.sim_genome(1, 5000)
    reads, lens, _, _ = native.sim_reads(2, g, 60, 1, 333, 0.0, 0.0, 1.0)
    reads[3, 0] = ord("N")
    reads[4, :10] = np.frombuffer(b"acgtnACGTN", np.uint8)
    pk = native.pack_reads(reads, lens, wstride=25)
    for r in range(60):
        L = int(lens[r])
        c = reads[r, :L].astype(np.uint32)
        code = ((c >> 1) & 3) ^ ((c >> 2) & 1)
        got = (pk[r, np.arange(L) >> 4] >> (2 * (np.arange(L) & 15)).astype(np.uint32)) & 3
        assert np.array_equal(got, code), r
        if L & 15:
            assert pk[r, L >> 4] >> (2 * (L & 15)) == 0
        assert not pk[r, (L + 15) // 16:].any()
    with pytest.raises(native.AccSeqError):
        native.pack_reads(reads, lens, wstride=2)


@pytest.mark.parametrize("threads,block", [(1, 5000), (3, 40000), (8, 7000), (5, 0)])
def test_parallel_ingest_and_emit_equal_the_serial_reader_and_printer(native, tmp_path, threads, block):
    """tests/compat/host_pipeline_check.c: ingest.c against fastq.c (kseq's grammar) record by record and word by word,
    emit.c against sam.c byte by byte, on the golden FASTQ (plain and gzip) and on a FASTQ of mixed-length reads with
    lower-case bases, N, CR LF line ends, comments after the name and no final newline"""
    import subprocess
    exe = os.path.join(ROOT, "tests", "_build", "host_pipeline_check")
    if not os.path.exists(exe):
        subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "tests"), "all"], check=True)
    plain = str(tmp_path / "g.fq")
    open(plain, "wb").write(gzip.open(os.path.join(GOLD, "reads.fq.gz"), "rb").read())
    rng = np.random.default_rng(threads)
    mixed = str(tmp_path / "m.fq")
    with open(mixed, "wb") as f:
        for i in range(900):
            L = int(rng.integers(1, 4000)) if i % 7 else int(rng.integers(1, 40))
            s = np.frombuffer(b"ACGTacgtN", np.uint8)[rng.integers(0, 9 if i % 5 == 0 else 4, L)].tobytes()
            q = bytes(rng.integers(33, 74, L).astype(np.uint8))       # quality strings may start with '@' or '+'
            eol = b"\r\n" if i % 11 == 0 else b"\n"
            f.write(b"@r%d some comment\tx" % i + eol + s + eol + b"+" + (b"r%d" % i if i % 3 == 0 else b"") + eol + q)
            if i != 899:
                f.write(eol)
    for path in (plain, os.path.join(GOLD, "reads.fq.gz"), mixed):
        p = subprocess.run([exe, path, str(threads), str(block)], stdout=subprocess.PIPE, stderr=subprocess.PIPE)
        assert p.returncode == 0, p.stderr.decode()
        assert b"host_pipeline_check ok" in p.stdout


def test_pack_reads_every_byte_and_length(native):
    """as_pack_reads (pack.c: 16 bases per step with SSSE3 where the CPU has it, else 8 per multiply) against the per-byte rule of
    include/accseq_b200.h -- code ((c >> 1) & 3) ^ ((c >> 2) & 1) for EVERY byte value, base j at bits 2(j & 15) of word j >> 4,
    zero past the length -- on lengths 0..600, row strides wider than needed, and rows that start at any alignment"""
    rng = np.random.default_rng(12)
    n, stride = 700, 641
    reads = rng.integers(0, 256, (n, stride)).astype(np.uint8)
    acgt = np.frombuffer(b"ACGTacgtN", np.uint8)
    reads[::2] = acgt[rng.integers(0, 9, (reads[::2].shape))]
    lens = rng.integers(0, 601, n).astype(np.uint32)
    lens[:40] = np.arange(40)
    for wstride in (38, 41):
        got = native.pack_reads(reads, lens, wstride)
        code = ((reads >> 1) & 3) ^ ((reads >> 2) & 1)
        want = np.zeros((n, wstride), np.uint32)
        for j in range(600):
            live = lens > j
            want[live, j >> 4] |= code[live, j].astype(np.uint32) << np.uint32(2 * (j & 15))
        assert np.array_equal(got, want)


def test_params_default(native):
    p = native.Params()
    native.cuda_lib().as_params_default(C.byref(p))
    assert (p.seed_len, p.thres, p.gact_T, p.gact_O, p.flags) == (20, 300, 320, 120, 0)   # alnmain.c:577-580


# ------------------------------------------------------------------------------------------------
# host helpers
# ------------------------------------------------------------------------------------------------
def test_cat_image_matches_oracle(native, orc):
    names, seqs = read_fasta_gz(os.path.join(GOLD, "repeat.fa.gz"))
    ix = orc.OracleIndex.build(names, seqs)
    cat, mta = native.cat_image(seqs)
    assert np.array_equal(cat, ix.cat)
    assert mta == [(o, s) for (_, o, s) in ix.mta]
    # and the C builder (asindex.c:97-110)
    H = native.host_lib()
    n = len(seqs)
    out_mta = (native.capi.AsMta * n)()
    con_len = C.c_uint64(0)
    p = H.as_cat_build(n, (C.c_char_p * n)(*seqs), (C.c_uint64 * n)(*[len(s) for s in seqs]), C.cast(out_mta, C.c_void_p),
                       C.byref(con_len))
    assert p and con_len.value == ix.n
    assert C.string_at(p, con_len.value) == ix.cat.tobytes()
    assert [(m.offset, m.seq_len) for m in out_mta] == mta
    C.CDLL(None).free(C.c_void_p(p))
    ix.close()


def test_index_loader_reads_reference_files(native, orc, tmp_path):
    """as_index_load reads the five files in the reference's formats (fmidx.c:246-275, lchash.c:114-127,
    alnmain.c:125-140,237-252); the files here are sha256-identical to accidx_ref's (test_oracle_vs_ref)."""
    names, seqs = read_fasta_gz(os.path.join(GOLD, "repeat.fa.gz"))
    ix = orc.OracleIndex.build(names, seqs)
    fa = str(tmp_path / "g.fa")
    ix.write(fa)
    H = native.host_lib()
    p = C.POINTER(native.capi.AsHostIndex)()
    assert H.as_index_load(fa.encode(), C.byref(p)) == 0
    v = p.contents.view
    assert v.length == ix.n and v.o_ratio == 32 and v.hlen == 12 and v.mta_len == 2 and v.sa_entry_bytes == 5
    view = oracle_lib._np_view
    assert np.array_equal(view(v.bwt, np.uint8, ix.n), ix.bwt)
    assert np.array_equal(view(v.C, np.uint64, 256), ix.Ctab)
    assert np.array_equal(view(v.O, np.uint64, v.o_len), ix.O)
    assert np.array_equal(view(v.lc, np.uint64, v.lc_len), ix.lc)
    assert np.array_equal(view(v.content, np.uint8, v.con_len), ix.cat)
    sa5 = view(v.sa, np.uint8, 5 * ix.n).reshape(-1, 5)
    assert np.array_equal(sa5[:, :4].copy().view("<u4").ravel().astype(np.uint64) | (sa5[:, 4].astype(np.uint64) << 32), ix.sa)
    assert [p.contents.names[i].decode() for i in range(2)] == names
    # write them back with the C writer: identical bytes
    fb = str(tmp_path / "h.fa")
    assert H.as_index_write(fb.encode(), p) == 0
    for ext in (".mta", ".cat", ".cat.sa5", ".cat.mfi", ".cat.lch"):
        assert open(fa + ext, "rb").read() == open(fb + ext, "rb").read(), ext
    H.as_index_free(p)
    assert H.as_index_load(str(tmp_path / "missing.fa").encode(), C.byref(p)) != 0
    ix.close()


class ReadBatch(C.Structure):
    _fields_ = [("n", C.c_size_t), ("stride", C.c_size_t), ("max_len", C.c_uint32), ("seqs", C.c_void_p),
                ("lens", C.POINTER(C.c_uint32)), ("names", C.POINTER(C.c_char_p)), ("quals", C.POINTER(C.c_char_p))]


def fastq_api(H):
    H.as_fastq_open.restype = C.c_void_p
    H.as_fastq_open.argtypes = [C.c_char_p]
    H.as_fastq_next_batch.argtypes = [C.c_void_p, C.c_size_t, C.POINTER(ReadBatch), C.c_void_p]
    H.as_read_batch_free.argtypes = [C.POINTER(ReadBatch), C.c_void_p]
    H.as_read_batch_free.restype = None
    H.as_fastq_close.argtypes = [C.c_void_p]
    H.as_fastq_close.restype = None
    return H


@pytest.mark.parametrize("gz", [True, False])
def test_fastq_batches(native, tmp_path, gz):
    """reads_load + refactor_reads_seq (accaln.c:45-58, alnmain.c:87-103): batches of <= max_reads, one zero
    padded buffer of stride max_len+1"""
    H = fastq_api(native.host_lib())
    recs = read_fastq_gz(os.path.join(GOLD, "reads.fq.gz"))
    path = os.path.join(GOLD, "reads.fq.gz")
    if not gz:
        path = str(tmp_path / "r.fq")
        open(path, "wb").write(gzip.open(os.path.join(GOLD, "reads.fq.gz"), "rb").read())
    fq = H.as_fastq_open(path.encode())
    assert fq
    seen = 0
    while True:
        b = ReadBatch()
        assert H.as_fastq_next_batch(fq, 100, C.byref(b), None) == 0
        if b.n == 0:
            break
        assert b.n <= 100 and b.stride == b.max_len + 1
        buf = oracle_lib._np_view(b.seqs, np.uint8, b.n * b.stride).reshape(b.n, b.stride)
        for i in range(b.n):
            name, s, q = recs[seen + i]
            assert b.lens[i] == len(s) and buf[i, :len(s)].tobytes() == s and not buf[i, len(s):].any()
            assert b.names[i].decode() == name and b.quals[i] == q
        assert b.max_len == max(len(recs[seen + i][1]) for i in range(b.n))
        seen += b.n
        H.as_read_batch_free(C.byref(b), None)
    assert seen == len(recs)
    H.as_fastq_close(fq)
    assert not H.as_fastq_open(str(tmp_path / "nope.fq").encode())


def test_cigar_text(native):
    assert native.cigar_text(np.array([0, 3, 0, 1, 2, 2, 0], np.uint8)) == "3M1I2D1M"      # edlib.cpp:262-311 STANDARD
    assert native.cigar_text(np.array([], np.uint8)) == ""
    buf = C.create_string_buffer(4)
    ops = np.array([0, 1] * 10, np.uint8)
    assert native.host_lib().as_cigar_text(ops.ctypes.data, ops.size, buf, 4) == -1


def test_sim_reads_are_seeded_and_truthful(native):
    g = native.sim_genome(3, 50_000)
    assert set(np.unique(g)) <= set(b"ACGT") and np.array_equal(g, native.sim_genome(3, 50_000))
    r1, l1, p1, s1 = native.sim_reads(9, g, 200, 100, 200, 0.0, 0.0, 1.0)
    r2, l2, p2, s2 = native.sim_reads(9, g, 200, 100, 200, 0.0, 0.0, 1.0)
    assert np.array_equal(r1, r2) and np.array_equal(p1, p2)
    comp = np.zeros(256, np.uint8)
    comp[list(b"ACGT")] = list(b"TGCA")
    for i in range(200):
        ref = g[int(p1[i]):int(p1[i]) + int(l1[i])]
        rd = r1[i, :l1[i]]
        assert np.array_equal(rd, ref if s1[i] == 0 else comp[ref][::-1])


def test_cigar_text_rle_host():
    """run bytes of AS_F_RLE_CIGAR print like the column ops they stand for (alnmain.c:497-498 parse_cigar)"""
    import longreadmapper_b200 as lm
    rng = np.random.default_rng(5)
    for _ in range(50):
        cls = rng.integers(0, 3, rng.integers(1, 40))
        lens = rng.integers(1, 200, cls.size)
        ops = np.concatenate([np.full(l, (0, 1, 2)[c], np.uint8) for c, l in zip(cls, lens)])
        ops[(ops == 0) & (rng.random(ops.size) < 0.1)] = 3                 # X prints as M
        runs = []
        for c, l in zip(cls, lens):
            while l > 0:
                t = min(63, int(l))
                runs.append((t << 2) | int(c))
                l -= t
        assert lm.cigar_text_rle(np.array(runs, np.uint8)) == lm.cigar_text(ops)
    with pytest.raises(lm.AccSeqError):
        lm.cigar_text_rle(np.array([3], np.uint8))
