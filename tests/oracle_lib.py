"""ctypes view of oracle/liboracle.so (the CPU restatement) and oracle/_ref (the reference's own object
code).  TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's CPU legs."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
REF_DIR = os.path.join(ORACLE_DIR, "_ref")

u64p = C.POINTER(C.c_uint64)


class OrcVote(C.Structure):
    _fields_ = [("best_key", C.c_uint64), ("best_val", C.c_uint64), ("phases", C.c_int32),
                ("from_outer", C.c_int32), ("n_seeds", C.c_uint64), ("n_hits", C.c_uint64),
                ("n_steps", C.c_uint64)]


RESULT_DTYPE = np.dtype([("loc", "<u8"), ("off", "<u8"), ("seq_id", "<i4"), ("strand", "<i4"),
                         ("flag", "<i4"), ("mapq", "<i4"), ("ed", "<i4"), ("valid", "<i4"),
                         ("n_ops", "<i4"), ("myers", "<i4"), ("phases", "<i4"), ("from_outer", "<i4")])
assert RESULT_DTYPE.itemsize == 56


def build_oracle():
    subprocess.run(["make", "-s", "-C", ORACLE_DIR, "oracle"], check=True)
    if os.path.isdir("/root/reference"):
        subprocess.run(["make", "-s", "-C", ORACLE_DIR, "ref"], check=True)


_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    path = os.path.join(ORACLE_DIR, "liboracle.so")
    if not os.path.exists(path):
        build_oracle()
    L = C.CDLL(path)
    L.orc_index_build.restype = C.c_void_p
    L.orc_index_build.argtypes = [C.c_int, C.POINTER(C.c_char_p), C.POINTER(C.c_char_p), u64p, C.c_int, C.c_int]
    L.orc_index_write.argtypes = [C.c_void_p, C.c_char_p]
    L.orc_index_load.restype = C.c_void_p
    L.orc_index_load.argtypes = [C.c_char_p]
    L.orc_index_free.argtypes = [C.c_void_p]
    L.orc_suffix_array.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p]
    L.orc_occ.restype = C.c_uint64
    L.orc_occ.argtypes = [C.c_void_p, C.c_char, C.c_uint64]
    L.orc_fmi_aln.restype = C.c_uint64
    L.orc_fmi_aln.argtypes = [C.c_void_p, C.c_char_p, C.c_int, u64p, u64p]
    L.orc_lc_aln.restype = C.c_uint64
    L.orc_lc_aln.argtypes = [C.c_void_p, C.c_char_p, C.c_int, u64p, u64p]
    L.orc_lc_aln_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p]
    L.orc_sa_access.restype = C.c_uint64
    L.orc_sa_access.argtypes = [C.c_void_p, C.c_uint64]
    L.orc_csa_access_batch.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p]
    L.orc_csa_access_batch.restype = None
    L.orc_vote_read.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.POINTER(OrcVote)]
    L.orc_seq_lookup.argtypes = [C.c_void_p, C.c_uint64, C.c_uint32, C.POINTER(C.c_int), C.POINTER(C.c_int), u64p, u64p]
    L.orc_gact.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p,
                           C.POINTER(C.c_int), u64p]
    L.orc_edit_distance_nw.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int]
    L.orc_edit_alignment_nw.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    L.orc_parse_cigar.argtypes = [C.c_void_p, C.c_int, C.c_char_p, C.c_size_t]
    L.orc_map_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_uint32,
                                C.c_uint32, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_size_t, u64p]
    _lib = L
    return L


class OrcIndexStruct(C.Structure):
    _fields_ = [("n", C.c_uint64), ("cat", C.c_void_p), ("C", C.c_uint64 * 256), ("o_ratio", C.c_int),
                ("o_len", C.c_uint64), ("O", C.c_void_p), ("bwt", C.c_void_p), ("csa_ratio", C.c_int),
                ("csa_len", C.c_uint64), ("csa", C.c_void_p), ("hlen", C.c_int), ("lc_len", C.c_uint64),
                ("lc", C.c_void_p), ("sa", C.c_void_p), ("n_mta", C.c_int), ("mta", C.c_void_p),
                ("sa5", C.c_void_p)]


class OrcMta(C.Structure):
    _fields_ = [("name", C.c_char_p), ("name_len", C.c_uint64), ("offset", C.c_uint64), ("seq_len", C.c_uint64)]


def _np_view(addr, dtype, count):
    buf = (C.c_char * (int(count) * np.dtype(dtype).itemsize)).from_address(addr)
    return np.frombuffer(buf, dtype=dtype)


class OracleIndex:
    """An in-memory index in the reference's layout (orc_index)."""

    def __init__(self, handle):
        self.h = handle
        self.s = OrcIndexStruct.from_address(handle)
        self.n = int(self.s.n)
        self._borrowed = None

    @classmethod
    def from_arrays(cls, *, cat, C_table, O, o_ratio, bwt, lc, hlen, sa, mta, names=None):
        """An orc_index over numpy arrays the caller already holds in the reference's layouts -- e.g. what
        as_gpu_export returned for an index built on the GPU at a BASELINE config's full size, where the oracle's own
        builder would take too long.  `sa` is uint64[N] or the raw .sa5 bytes uint8[5N]; mta = [(offset, seq_len)].
        Nothing is copied; the arrays stay referenced by the object."""
        lib()
        cat, C_table, O, bwt, lc, sa = (np.ascontiguousarray(x) for x in (cat, C_table, O, bwt, lc, sa))
        assert cat.dtype == np.uint8 and bwt.dtype == np.uint8 and C_table.dtype == np.uint64 and C_table.size == 256
        assert O.dtype == np.uint64 and lc.dtype == np.uint64 and lc.size == 2 * 4 ** hlen
        n = int(cat.size)
        s = OrcIndexStruct()
        s.n, s.cat, s.o_ratio, s.o_len, s.O, s.bwt = n, cat.ctypes.data, int(o_ratio), int(O.size), O.ctypes.data, bwt.ctypes.data
        C.memmove(s.C, C_table.ctypes.data, 256 * 8)
        s.csa_ratio, s.csa_len, s.csa = 0, 0, None
        s.hlen, s.lc_len, s.lc = int(hlen), int(lc.size), lc.ctypes.data
        if sa.dtype == np.uint64:
            assert sa.size == n
            s.sa, s.sa5 = sa.ctypes.data, None
        else:
            assert sa.dtype == np.uint8 and sa.size == 5 * n
            s.sa, s.sa5 = None, sa.ctypes.data
        names = list(names) if names else [f"seq{i}" for i in range(len(mta))]
        name_bufs = [C.create_string_buffer(x.encode()) for x in names]
        marr = (OrcMta * max(1, len(mta)))()
        for i, (o, l) in enumerate(mta):
            marr[i].name = C.cast(name_bufs[i], C.c_char_p)
            marr[i].name_len, marr[i].offset, marr[i].seq_len = len(names[i]), int(o), int(l)
        s.n_mta, s.mta = len(mta), C.cast(marr, C.c_void_p)
        self = cls.__new__(cls)
        self.s, self.h, self.n = s, C.addressof(s), n
        self._borrowed = (cat, C_table, O, bwt, lc, sa, marr, name_bufs)
        return self

    @classmethod
    def from_gpu_export(cls, mapper, cat, mta, names=None, sa_entry_bytes=8):
        """The index a GpuMapper holds, exported in the reference's layouts (as_gpu_export) and wrapped for the oracle."""
        ex = mapper.export(sa_entry_bytes=sa_entry_bytes, want=("C", "O", "bwt", "lc", "sa"))
        return cls.from_arrays(cat=cat, C_table=ex["C"], O=ex["O"], o_ratio=32, bwt=ex["bwt"], lc=ex["lc"], hlen=12,
                               sa=ex["sa"], mta=mta, names=names)

    @classmethod
    def build(cls, names, seqs, o_ratio=32, hlen=12):
        L = lib()
        n = len(seqs)
        names_b = [x.encode() if isinstance(x, str) else x for x in names]
        seqs_b = [x.encode() if isinstance(x, str) else bytes(x) for x in seqs]
        h = L.orc_index_build(n, (C.c_char_p * n)(*names_b), (C.c_char_p * n)(*seqs_b),
                              (C.c_uint64 * n)(*[len(s) for s in seqs_b]), o_ratio, hlen)
        if not h:
            raise RuntimeError("orc_index_build failed")
        return cls(h)

    @classmethod
    def load(cls, fasta_path):
        h = lib().orc_index_load(fasta_path.encode())
        if not h:
            raise RuntimeError("orc_index_load failed for " + fasta_path)
        return cls(h)

    def write(self, fasta_path):
        if lib().orc_index_write(self.h, fasta_path.encode()) != 0:
            raise RuntimeError("orc_index_write failed")

    # numpy views (no copies)
    @property
    def cat(self): return _np_view(self.s.cat, np.uint8, self.n)
    @property
    def bwt(self): return _np_view(self.s.bwt, np.uint8, self.n)
    @property
    def Ctab(self): return np.frombuffer(self.s.C, dtype=np.uint64)
    @property
    def O(self): return _np_view(self.s.O, np.uint64, self.s.o_len)
    @property
    def lc(self): return _np_view(self.s.lc, np.uint64, self.s.lc_len)
    @property
    def sa(self): return _np_view(self.s.sa, np.uint64, self.n)
    @property
    def mta(self):
        arr = (OrcMta * self.s.n_mta).from_address(self.s.mta)
        return [(m.name.decode(), int(m.offset), int(m.seq_len)) for m in arr]

    def lc_aln_batch(self, seeds, qlen):
        seeds = np.ascontiguousarray(seeds, dtype=np.uint8)
        n = seeds.size // qlen
        k = np.zeros(n, np.uint64); l = np.zeros(n, np.uint64); r = np.zeros(n, np.uint64)
        lib().orc_lc_aln_batch(self.h, seeds.ctypes.data, qlen, n, k.ctypes.data, l.ctypes.data, r.ctypes.data)
        return k, l, r

    def csa_access_batch(self, rows, ratio=4, literal=False):
        rows = np.ascontiguousarray(rows, dtype=np.uint64)
        pos = np.zeros(rows.size, np.uint64)
        lib().orc_csa_access_batch(self.h, ratio, rows.ctypes.data, rows.size, int(literal), pos.ctypes.data)
        return pos

    def fmi_aln(self, q):
        k = C.c_uint64(1); l = C.c_uint64(self.n - 1)
        r = lib().orc_fmi_aln(self.h, q, len(q), C.byref(k), C.byref(l))
        return int(k.value), int(l.value), int(r)

    def vote(self, seq, seed_len=20, thres=300):
        v = OrcVote()
        seq = np.ascontiguousarray(seq, dtype=np.uint8)
        lib().orc_vote_read(self.h, seq.ctypes.data, seq.size, seed_len, thres, C.byref(v))
        return v

    def map_batch(self, reads, stride, lens, seed_len=20, thres=300, T=320, O=120, myers=True, ops_stride=None):
        """reads: uint8 array n*stride (modified in place for strand-1 reads). Returns (results, ops, cells)."""
        n = len(lens)
        lens = np.ascontiguousarray(lens, dtype=np.uint32)
        if ops_stride is None:
            ops_stride = 2 * int(lens.max()) if n else 2
        res = np.zeros(n, dtype=RESULT_DTYPE)
        ops = np.zeros(n * ops_stride, dtype=np.uint8)
        cells = C.c_uint64(0)
        lib().orc_map_batch(self.h, reads.ctypes.data, stride, lens.ctypes.data, n, seed_len, thres, T, O,
                            1 if myers else 0, res.ctypes.data, ops.ctypes.data, ops_stride, C.byref(cells))
        return res, ops.reshape(n, ops_stride) if n else ops, int(cells.value)

    def close(self):
        if self.h and self._borrowed is None:
            lib().orc_index_free(self.h)
        self.h = None
        self._borrowed = None


def gpu_mapper(lm, ox, device=0, opts=None):
    """A GpuMapper over an OracleIndex (the reference's in-memory layout) -- test convenience, so that the package itself
    never names the oracle."""
    return lm.GpuMapper(device, length=ox.n, C_table=ox.Ctab, O=ox.O, o_ratio=ox.s.o_ratio, bwt=ox.bwt, lc=ox.lc,
                        hlen=ox.s.hlen, sa=ox.sa, sa_entry_bytes=8, content=ox.cat, mta=[(o, s) for (_, o, s) in ox.mta],
                        opts=opts)


def gact(q, t, T=320, O=120):
    q = np.ascontiguousarray(q, dtype=np.uint8); t = np.ascontiguousarray(t, dtype=np.uint8)
    ops = np.zeros(q.size + t.size + 2, dtype=np.uint8)
    n = C.c_int(0); cells = C.c_uint64(0)
    score = lib().orc_gact(q.ctypes.data, q.size, t.ctypes.data, t.size, T, O, ops.ctypes.data, C.byref(n), C.byref(cells))
    return score, ops[: n.value].copy(), int(cells.value)


def edit_distance(q, t, k=-1):
    q = np.ascontiguousarray(q, dtype=np.uint8); t = np.ascontiguousarray(t, dtype=np.uint8)
    return lib().orc_edit_distance_nw(q.ctypes.data, q.size, t.ctypes.data, t.size, k)


def cigar_text(ops):
    ops = np.ascontiguousarray(ops, dtype=np.uint8)
    buf = C.create_string_buffer(4 * ops.size + 16)
    n = lib().orc_parse_cigar(ops.ctypes.data, ops.size, buf, len(buf))
    if n < 0:
        raise ValueError("bad ops")
    return buf.value.decode()


# ---------------------------------------------------------------------------------------------------
# oracle/_ref: the reference's own objects (present here; prebuilt copy travels to the GPU box)
# ---------------------------------------------------------------------------------------------------
def have_ref():
    return all(os.path.exists(os.path.join(REF_DIR, f)) for f in ("libref.so", "accaln_ref", "accidx_ref", "libedlib_ref.so"))


class RefMstring(C.Structure):
    _fields_ = [("l", C.c_uint64), ("s", C.c_char_p), ("own", C.c_int)]


class RefFmi(C.Structure):      # dna_fmi, fmidx.h:16-21
    _fields_ = [("length", C.c_uint64), ("o_len", C.c_uint64), ("csa_len", C.c_uint64), ("c", C.c_void_p),
                ("o", C.c_void_p), ("csa", C.c_void_p), ("o_ratio", C.c_int), ("csa_ratio", C.c_int),
                ("bwt", C.c_void_p)]


class RefLch(C.Structure):      # lc_hash, lchash.h:12-16
    _fields_ = [("lc", C.c_void_p), ("len", C.c_uint64), ("hlen", C.c_int)]


class RefEntry(C.Structure):    # histo.h:21-23
    _fields_ = [("key", C.c_uint64), ("val", C.c_uint64), ("bucket", C.c_uint64)]


_ref = None


def ref():
    global _ref
    if _ref is None:
        R = C.CDLL(os.path.join(REF_DIR, "libref.so"))
        R.fmi_read.argtypes = [C.POINTER(RefFmi), C.c_char_p]
        R.lc_read.argtypes = [C.c_char_p, C.POINTER(RefLch)]
        R.fmi_aln.restype = C.c_uint64
        R.fmi_aln.argtypes = [C.POINTER(RefFmi), C.c_char_p, C.c_int, u64p, u64p]
        R.lc_aln.restype = C.c_uint64
        R.lc_aln.argtypes = [C.c_char_p, C.c_int, u64p, u64p, C.POINTER(RefFmi), C.POINTER(RefLch)]
        R.csa_access.restype = C.c_uint64
        R.csa_access.argtypes = [C.POINTER(RefFmi), C.c_uint64]
        R.histo_init.restype = C.c_void_p
        R.histo_init.argtypes = [C.c_uint32]
        R.histo_add.argtypes = [C.c_void_p, C.c_uint64]
        R.histo_find_2_max.restype = C.c_uint64
        R.histo_find_2_max.argtypes = [C.c_void_p, C.POINTER(RefEntry)]
        R.histo_destroy.argtypes = [C.c_void_p]
        _ref = R
    return _ref


_edlib = None


class EdlibConfig(C.Structure):
    _fields_ = [("k", C.c_int), ("mode", C.c_int), ("task", C.c_int), ("eq", C.c_void_p), ("neq", C.c_int)]


class EdlibResult(C.Structure):
    _fields_ = [("status", C.c_int), ("editDistance", C.c_int), ("endLocations", C.c_void_p),
                ("startLocations", C.c_void_p), ("numLocations", C.c_int), ("alignment", C.POINTER(C.c_ubyte)),
                ("alignmentLength", C.c_int), ("alphabetLength", C.c_int)]


def edlib_ref():
    global _edlib
    if _edlib is None:
        E = C.CDLL(os.path.join(REF_DIR, "libedlib_ref.so"))
        E.edlibAlign.restype = EdlibResult
        E.edlibAlign.argtypes = [C.c_char_p, C.c_int, C.c_char_p, C.c_int, EdlibConfig]
        E.edlibFreeAlignResult.argtypes = [EdlibResult]
        E.edlibAlignmentToCigar.restype = C.c_void_p
        E.edlibAlignmentToCigar.argtypes = [C.POINTER(C.c_ubyte), C.c_int, C.c_int]
        _edlib = E
    return _edlib


def edlib_nw(q, t, k=-1, path=False):
    """Reference edlib, NW mode. Returns (distance, cigar-or-None)."""
    E = edlib_ref()
    cfg = EdlibConfig(k, 0, 2 if path else 0, None, 0)
    r = E.edlibAlign(bytes(q), len(q), bytes(t), len(t), cfg)
    cig = None
    if path and r.editDistance >= 0 and r.alignment:
        p = E.edlibAlignmentToCigar(r.alignment, r.alignmentLength, 0)
        cig = C.string_at(p).decode()
    d = r.editDistance
    E.edlibFreeAlignResult(r)
    return d, cig


def edlib_nw_ops(q, t, k=-1):
    """Reference edlib, NW mode, EDLIB_TASK_PATH: (distance, alignment ops as uint8 array or None)."""
    E = edlib_ref()
    r = E.edlibAlign(bytes(q), len(q), bytes(t), len(t), EdlibConfig(k, 0, 2, None, 0))
    ops = None
    if r.editDistance >= 0 and r.alignment:
        ops = np.ctypeslib.as_array(r.alignment, shape=(r.alignmentLength,)).copy()
    d = r.editDistance
    E.edlibFreeAlignResult(r)
    return d, ops


def edit_alignment(q, t, k=-1):
    """oracle: edlib's NW alignment path (orc_edit_alignment_nw): (distance, ops forward order)"""
    q = np.ascontiguousarray(q, dtype=np.uint8)
    t = np.ascontiguousarray(t, dtype=np.uint8)
    ops = np.zeros(q.size + t.size + 1, np.uint8)
    n = C.c_int(0)
    d = lib().orc_edit_alignment_nw(q.ctypes.data, q.size, t.ctypes.data, t.size, k, ops.ctypes.data, C.byref(n))
    return d, ops[: n.value].copy()


def run_ref_accidx(fasta_path):
    subprocess.run([os.path.join(REF_DIR, "accidx_ref"), fasta_path], check=True, stdout=subprocess.DEVNULL)


def run_ref_accaln(fasta_path, fastq_path, batch=100000, seed_len=20, thres=300):
    """Returns SAM text produced by the compiled reference."""
    p = subprocess.run([os.path.join(REF_DIR, "accaln_ref"), fasta_path, fastq_path, str(batch), str(seed_len), str(thres)],
                       check=True, stdout=subprocess.PIPE, stderr=subprocess.PIPE)
    return p.stdout.decode()


def parse_sam(text):
    recs = []
    for line in text.splitlines():
        if not line or line[0] == "@":
            continue
        f = line.split("\t")
        recs.append(dict(qname=f[0], flag=int(f[1]), rname=f[2], pos=int(f[3]), mapq=int(f[4]), cigar=f[5],
                         seq=f[9], qual=f[10], ed=int(f[11].split(":")[2])))
    return recs
