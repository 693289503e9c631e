"""libaccseq_compat.so -- the reference's SCALAR C surface (lc_aln lchash.h:25-27, lc_access :22, fmi_aln
fmidx.h:37-38, sa_access :30, csa_access :31, cigar_align mutils.h:56-57, histo_* histo.h:32-37, and the gact.h names
ms_borrow / simple_gact / parse_cigar) as thin wrappers over the batched C ABI -- and the INTEGRATION.md patch applied to the
reference's own alnmain.c, linked against the PRODUCT libraries only.

CPU: the library exports every name, the patch applies to the pristine file and the patched program links without oracle/.
GPU: compat_check (tests/compat/compat_check.c: the invariants of test/test-lchash.cc:19-48, test/test-fmidx.cc:17-41,104-106,
edlib/ed_test.c) compiled against this repo's header AND against the reference's own headers; the scalar wrappers against the
golden lookups the reference's object code produced; the patched reference program's SAM against tests/golden/expected.sam.gz."""
import ctypes as C
import gzip
import os
import re
import subprocess

import numpy as np
import pytest

import oracle_lib
from test_oracle_vs_ref import GOLD, read_fasta_gz

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "longreadmapper_b200", "libaccseq_compat.so")
BUILD = os.path.join(ROOT, "tests", "_build")
HAVE_REF_TREE = os.path.isdir("/root/reference")


def make_artifacts():
    subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "tests"), "all"], check=True)


@pytest.fixture(scope="module")
def artifacts(native):
    if not os.path.exists(os.path.join(BUILD, "compat_check")):
        make_artifacts()
    return BUILD


def test_compat_library_exports_the_reference_names(native):
    L = C.CDLL(LIB)
    text = open(os.path.join(ROOT, "include", "accseq_compat.h")).read() + open(os.path.join(ROOT, "include", "gact", "gact.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    names = set(re.findall(r"\b([a-z_0-9]+)\s*\([^;{]*\)\s*;", text))
    want = {"lc_access", "lc_build", "lc_length", "lc_aln", "lc_destroy", "lc_write", "lc_read", "sa_access", "csa_access",
            "sa_done_access", "fmi_build", "fmi_aln", "fmi_write", "fmi_read", "fmi_destroy", "sa_build", "histo_init",
            "histo_add", "histo_destroy", "histo_get", "histo_find_max", "histo_find_2_max", "cigar_align", "ms_borrow",
            "simple_gact", "parse_cigar"}
    assert want <= names, want - names
    for n in sorted(names):
        assert hasattr(L, n), f"libaccseq_compat.so does not export {n}"
    assert C.c_void_p.in_dll(L, "sa_buf") is not None                # the global of fmidx.c:16


@pytest.mark.skipif(not HAVE_REF_TREE, reason="/root/reference is not on this machine")
def test_patch_applies_and_links_against_product_libraries_only(native, tmp_path):
    """integration/alnmain_b200.ed on the pristine alnmain.c; the result, with the reference's own accaln.c / mutils.c /
    mlog/logger.c, links against libaccseq_compat + libaccseq_b200 + libaccseq_host -- nothing from oracle/."""
    make_artifacts()
    exe = os.path.join(BUILD, "accaln_patched")
    assert os.path.exists(exe) and os.path.exists(os.path.join(BUILD, "compat_check_refhdr"))
    needed = subprocess.run(["ldd", exe], stdout=subprocess.PIPE, text=True).stdout
    assert "libaccseq_compat.so" in needed and "libaccseq_b200.so" in needed and "oracle" not in needed
    undefined = subprocess.run(["nm", "-u", exe], stdout=subprocess.PIPE, text=True).stdout
    for n in ("as_map_batch", "as_gpu_open", "fmi_read", "lc_read", "parse_cigar", "simple_gact"):
        assert re.search(rf"\b{n}\b", undefined), n
    for n in ("orc_", "histo_add", "lc_aln"):                          # the loop that called them is gone
        assert n not in undefined
    src = str(tmp_path / "patched.c")
    subprocess.run(["python", os.path.join(ROOT, "integration", "apply_ed.py"), "/root/reference/alnmain.c", src], check=True)
    text = open(src).read()
    assert "as_map_batch(as_b200" in text and "histo_find_2_max" not in text.split("static inline int single_end")[1].split("static inline int pair_end")[0]
    import torch
    if not torch.cuda.is_available():                                  # no device: refuses, loudly
        p = subprocess.run([exe, "/nonexistent.fa", "/nonexistent.fq", "10", "20", "300"], stdout=subprocess.PIPE, stderr=subprocess.PIPE)
        assert p.returncode != 0


@pytest.fixture(scope="module")
def golden_files(native, orc, tmp_path_factory):
    d = tmp_path_factory.mktemp("compat")
    names, seqs = read_fasta_gz(os.path.join(GOLD, "repeat.fa.gz"))
    fa = str(d / "repeat.fa")
    with open(fa, "wb") as f:
        for n, s in zip(names, seqs):
            f.write(b">" + n.encode() + b"\n" + s + b"\n")
    ix = orc.OracleIndex.build(names, seqs)
    ix.write(fa)
    ix.close()
    fq = str(d / "reads.fq")
    open(fq, "wb").write(gzip.open(os.path.join(GOLD, "reads.fq.gz"), "rb").read())
    return fa, fq


@pytest.mark.gpu
@pytest.mark.parametrize("exe", ["compat_check", "compat_check_refhdr"])
def test_reference_unit_test_invariants(native, artifacts, tmp_path, exe):
    path = os.path.join(artifacts, exe)
    if not os.path.exists(path):
        pytest.skip(f"{exe} is built only where /root/reference exists")
    g = native.sim_genome(3, 60_000)
    cat, _ = native.cat_image([g])
    text = str(tmp_path / "input2-cap.txt")
    open(text, "wb").write(cat.tobytes())
    p = subprocess.run([path, text], stdout=subprocess.PIPE, stderr=subprocess.PIPE, timeout=600)
    assert p.returncode == 0, p.stderr.decode()[-2000:]
    assert b"compat_check ok" in p.stdout


class Fmi(C.Structure):
    _fields_ = [("length", C.c_uint64), ("o_len", C.c_uint64), ("csa_len", C.c_uint64), ("c", C.c_void_p), ("o", C.c_void_p),
                ("csa", C.c_void_p), ("o_ratio", C.c_int), ("csa_ratio", C.c_int), ("bwt", C.c_void_p)]


class Lch(C.Structure):
    _fields_ = [("lc", C.c_void_p), ("len", C.c_uint64), ("hlen", C.c_int)]


class Cigar(C.Structure):
    _fields_ = [("cigar", C.c_void_p), ("n_cigar_op", C.c_int), ("score", C.c_int)]


@pytest.mark.gpu
def test_scalar_wrappers_match_reference_object_code(native, orc, golden_files):
    """fmi_read / lc_read the files, then lc_aln / fmi_aln seed by seed: the values libref.so (the reference's own fmidx.c /
    lchash.c) produced for the golden seeds; csa_access against the oracle's literal form (pinned on libref.so);
    cigar_align against orc_gact"""
    fa, _ = golden_files
    L = C.CDLL(LIB)
    u64p = C.POINTER(C.c_uint64)
    L.fmi_read.argtypes = [C.POINTER(Fmi), C.c_char_p]
    L.lc_read.argtypes = [C.c_char_p, C.POINTER(Lch)]
    L.fmi_aln.restype = C.c_uint64
    L.fmi_aln.argtypes = [C.POINTER(Fmi), C.c_char_p, C.c_int, u64p, u64p]
    L.lc_aln.restype = C.c_uint64
    L.lc_aln.argtypes = [C.c_char_p, C.c_int, u64p, u64p, C.POINTER(Fmi), C.POINTER(Lch)]
    L.csa_access.restype = C.c_uint64
    L.csa_access.argtypes = [C.POINTER(Fmi), C.c_uint64]
    L.cigar_align.restype = Cigar
    L.cigar_align.argtypes = [C.c_char_p, C.c_int, C.c_char_p, C.c_int, C.POINTER(C.c_int), C.c_void_p]
    L.fmi_destroy.argtypes = [C.POINTER(Fmi)]
    fmi, lch = Fmi(), Lch()
    L.fmi_read(C.byref(fmi), (fa + ".cat").encode())
    L.lc_read((fa + ".cat.lch").encode(), C.byref(lch))
    d = np.load(os.path.join(GOLD, "lookups.npz"))
    for qlen in (20, 12, 32, 64):
        seeds, want = d[f"seeds{qlen}"], d[f"res{qlen}"]
        for i in range(0, 1000, 5):
            q = seeds[i].tobytes()
            k, l = C.c_uint64(1), C.c_uint64(int(fmi.length) - 1)
            r = L.lc_aln(q, qlen, C.byref(k), C.byref(l), C.byref(fmi), C.byref(lch))
            assert (k.value, l.value, r) == tuple(int(x) for x in want[i, :3]), (qlen, i)
            k, l = C.c_uint64(1), C.c_uint64(int(fmi.length) - 1)
            r = L.fmi_aln(C.byref(fmi), q, qlen, C.byref(k), C.byref(l))
            assert (k.value, l.value, r) == tuple(int(x) for x in want[i, 3:]), (qlen, i)
    names, seqs = read_fasta_gz(os.path.join(GOLD, "repeat.fa.gz"))
    ix = orc.OracleIndex.build(names, seqs)
    try:
        rows = np.random.default_rng(4).integers(0, ix.n, 300).astype(np.uint64)
        want = ix.csa_access_batch(rows, ratio=4, literal=True)
        got = np.array([L.csa_access(C.byref(fmi), int(r)) for r in rows], dtype=np.uint64)
        assert np.array_equal(got, want)
    finally:
        ix.close()
    rng = np.random.default_rng(5)
    for _ in range(20):
        n = int(rng.integers(30, 700))
        t = np.frombuffer(b"ACGT", np.uint8)[rng.integers(0, 4, n)]
        q = t.copy()
        q[rng.integers(0, n, 6)] = ord("A")
        q = np.delete(q, rng.integers(0, n, 2))
        ops = np.zeros(q.size + t.size + 2, np.uint8)
        lim = C.c_int(0)
        c = L.cigar_align(q.tobytes(), q.size, t.tobytes(), t.size, C.byref(lim), ops.ctypes.data)
        es, eops, _ = oracle_lib.gact(q, t)
        assert (c.score, lim.value, c.n_cigar_op) == (es, es, len(eops)) and np.array_equal(ops[: c.n_cigar_op], eops)
    L.fmi_destroy(C.byref(fmi))
    L.as_compat_release()


@pytest.mark.gpu
def test_patched_reference_program_prints_the_golden_sam(native, artifacts, golden_files):
    """INTEGRATION.md section 2 for real: the reference's alnmain.c with the ed script applied, its own accaln.c (reads_load,
    refactor_reads_seq), mutils.c and SAM loop, the product libraries underneath -- against the SAM the unpatched reference
    printed (tests/golden/expected.sam.gz)"""
    exe = os.path.join(artifacts, "accaln_patched")
    if not os.path.exists(exe):
        pytest.skip("accaln_patched is built only where /root/reference exists")
    fa, fq = golden_files
    for batch in ("100000", "97"):
        p = subprocess.run([exe, fa, fq, batch, "20", "300"], stdout=subprocess.PIPE, stderr=subprocess.PIPE, timeout=600)
        assert p.returncode == 0, p.stderr.decode()[-2000:]
        got_all = p.stdout.decode().splitlines()
        want_all = gzip.open(os.path.join(GOLD, "expected.sam.gz"), "rt").read().splitlines()
        assert [l for l in got_all if l[0] == "@" and not l.startswith("@RG")] == [l for l in want_all if l[0] == "@"]
        got = [l for l in got_all if l[0] != "@"]
        want = [l for l in want_all if l[0] != "@"]
        assert len(got) == len(want) == 653
        same = 0
        for g, w in zip(got, want):
            gf = g.split("\t")
            if gf[2] == "*" or len(gf[9]) <= 20:      # seq_lookup failed / len <= seed_len: the reference prints uninitialised memory
                assert gf[2] != "*" or (gf[1], gf[4]) == ("4", "0")
                continue
            assert g == w
            same += 1
        assert same > 600
        assert b"Sensitivity: " in p.stderr
