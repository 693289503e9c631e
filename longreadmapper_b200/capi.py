"""ctypes bindings of libaccseq_b200.so / libaccseq_host.so (see include/*.h for the contracts)."""
import ctypes as C
import os
import subprocess

import numpy as np

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
CSRC_DIR = os.path.join(PKG_DIR, "csrc")
CUDA_LIB = os.path.join(PKG_DIR, "libaccseq_b200.so")
HOST_LIB = os.path.join(PKG_DIR, "libaccseq_host.so")

AS_F_MYERS = 1
AS_F_NO_ALIGN = 2
AS_F_GACT_WARP = 4
AS_F_GACT_BITPAR = 8
AS_F_WIDE_INDEX = 16
AS_F_NO_LC2 = 32
AS_F_RLE_CIGAR = 64
AS_F_KEEP_READS = 128

# as_result (include/accseq_b200.h)
RESULT_DTYPE = np.dtype([("loc", "<u8"), ("off", "<u8"), ("seq_id", "<i4"), ("strand", "<i4"), ("flag", "<i4"),
                         ("mapq", "<i4"), ("ed", "<i4"), ("valid", "<i4"), ("n_cigar_op", "<i4"), ("myers", "<i4"),
                         ("phases", "<i4"), ("from_outer", "<i4")])
assert RESULT_DTYPE.itemsize == 56


# as_result_compact (include/accseq_b200.h)
COMPACT_DTYPE = np.dtype([("pos", "<u8"), ("ed", "<i4"), ("n_runs", "<u4")])
assert COMPACT_DTYPE.itemsize == 16


class AccSeqError(RuntimeError):
    pass


class AsMta(C.Structure):
    _fields_ = [("offset", C.c_uint64), ("seq_len", C.c_uint64)]


class AsIndexView(C.Structure):
    _fields_ = [("length", C.c_uint64), ("C", C.c_void_p), ("O", C.c_void_p), ("o_len", C.c_uint64),
                ("o_ratio", C.c_int), ("bwt", C.c_void_p), ("lc", C.c_void_p), ("lc_len", C.c_uint64),
                ("hlen", C.c_int), ("sa", C.c_void_p), ("sa_entry_bytes", C.c_int), ("content", C.c_void_p),
                ("con_len", C.c_uint64), ("mta", C.c_void_p), ("mta_len", C.c_int)]


class AsIndexOut(C.Structure):
    _fields_ = [("C", C.c_void_p), ("O", C.c_void_p), ("o_ratio", C.c_int), ("bwt", C.c_void_p), ("lc", C.c_void_p),
                ("sa", C.c_void_p), ("sa_entry_bytes", C.c_int), ("csa", C.c_void_p), ("csa_ratio", C.c_int)]


class Params(C.Structure):
    _fields_ = [("seed_len", C.c_uint32), ("thres", C.c_uint32), ("gact_T", C.c_int32), ("gact_O", C.c_int32),
                ("flags", C.c_uint32), ("gact_band", C.c_int32), ("vote_ectx", C.c_int32), ("gact_order", C.c_int32),
                ("gact_blocks_per_sm", C.c_int32), ("chunk_reads", C.c_uint32), ("chunk_tail", C.c_int32),
                ("vote_thread", C.c_int32), ("gact_runs", C.c_int32), ("vote_blocks_per_sm", C.c_int32)]


class OpenOpts(C.Structure):
    _fields_ = [("lc2_hlen", C.c_int32), ("no_sax", C.c_int32), ("build_part_max", C.c_uint64), ("sampled_sa", C.c_int32),
                ("l2_fetch_gran", C.c_int32), ("reserved", C.c_int32 * 4)]


def open_opts(lc2_hlen=0, no_sax=0, build_part_max=0, sampled_sa=0, l2_fetch_gran=0):
    o = OpenOpts()
    o.lc2_hlen, o.no_sax, o.build_part_max, o.sampled_sa = int(lc2_hlen), int(no_sax), int(build_part_max), int(sampled_sa)
    o.l2_fetch_gran = int(l2_fetch_gran)
    return o


class AsStats(C.Structure):
    _fields_ = [(k, C.c_uint64) for k in ("total", "valid", "n_flag4", "n_flag16", "sum_ed", "dp_cells", "seeds",
                                           "sa_hits", "bs_steps", "overflow_reads", "dp_cells_eval", "gact_redone", "fm_steps_exec", "text_verified",
                                           "overflow_queued", "vote_deferred", "gact_queued")]


class AsHostIndex(C.Structure):
    _fields_ = [("view", AsIndexView), ("C", C.c_void_p), ("O", C.c_void_p), ("lc", C.c_void_p), ("csa", C.c_void_p),
                ("bwt", C.c_void_p), ("content", C.c_void_p), ("sa5", C.c_void_p), ("mta", C.c_void_p),
                ("names", C.POINTER(C.c_char_p)), ("csa_len", C.c_uint64), ("csa_ratio", C.c_int)]


def build_native(verbose=False):
    """Compile both libraries in-tree (nvcc cross-compiles sm_100a without a GPU)."""
    out = None if verbose else subprocess.DEVNULL
    subprocess.run(["make", "-s", "-j8", "-C", CSRC_DIR, "all"], check=True, stdout=out)


_cuda = None
_host = None


def cuda_lib():
    """libaccseq_b200.so; raises if it is not built -- there is no fallback path."""
    global _cuda
    if _cuda is not None:
        return _cuda
    if not os.path.exists(CUDA_LIB):
        raise AccSeqError(f"{CUDA_LIB} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                          "(make -C longreadmapper_b200/csrc). There is no CPU fallback.")
    L = C.CDLL(CUDA_LIB)
    vp, u64, sz, i32 = C.c_void_p, C.c_uint64, C.c_size_t, C.c_int
    L.as_last_error.restype = C.c_char_p
    L.as_device_count.restype = i32
    L.as_gpu_open.argtypes = [C.POINTER(vp), i32, C.POINTER(AsIndexView)]
    L.as_gpu_build.argtypes = [C.POINTER(vp), i32, vp, u64, vp, i32, i32]
    L.as_gpu_open_opts.argtypes = [C.POINTER(vp), i32, C.POINTER(AsIndexView), C.POINTER(OpenOpts)]
    L.as_gpu_build_opts.argtypes = [C.POINTER(vp), i32, vp, u64, vp, i32, i32, C.POINTER(OpenOpts)]
    L.as_gpu_export.argtypes = [vp, C.POINTER(AsIndexOut)]
    L.as_gpu_index_length.argtypes = [vp]
    L.as_gpu_index_length.restype = u64
    L.as_gpu_close.argtypes = [vp]
    L.as_gpu_close.restype = None
    L.as_gpu_set_stream.argtypes = [vp, vp]
    L.as_gpu_sync.argtypes = [vp]
    L.as_gpu_launch_count.argtypes = [vp]
    L.as_gpu_launch_count.restype = u64
    L.as_params_default.argtypes = [C.POINTER(Params)]
    L.as_params_default.restype = None
    L.as_map_batch.argtypes = [vp, C.POINTER(Params), vp, sz, vp, sz, vp, vp, sz]
    L.as_map_batch_packed.argtypes = [vp, C.POINTER(Params), vp, sz, vp, sz, vp, vp, sz, C.POINTER(u64)]
    L.as_map_batch_device.argtypes = [vp, C.POINTER(Params), vp, sz, vp, sz, C.c_uint32, vp, vp, sz]
    L.as_gpu_stats.argtypes = [vp, C.POINTER(AsStats)]
    L.as_gpu_stats_reset.argtypes = [vp]
    L.as_gpu_last_timing.argtypes = [vp, C.POINTER(C.c_float * 8)]
    L.as_gpu_last_chunk_timeline.argtypes = [vp, i32, C.POINTER(C.c_float * 8)]
    L.as_gpu_last_host_timing.argtypes = [vp, C.POINTER(C.c_double * 3)]
    L.as_gpu_last_transfer.argtypes = [vp, C.POINTER(u64), C.POINTER(u64)]
    L.as_gpu_last_reads_done.argtypes = [vp, C.POINTER(C.c_size_t)]
    L.as_gpu_last_reads_done.restype = C.c_size_t
    L.as_lc_aln_batch.argtypes = [vp, vp, i32, sz, vp, vp, vp]
    L.as_fmi_aln_batch.argtypes = [vp, vp, i32, sz, vp, vp, vp]
    L.as_sa_access_batch.argtypes = [vp, vp, sz, vp]
    L.as_csa_access_batch.argtypes = [vp, i32, i32, vp, sz, vp]
    L.as_vote_batch.argtypes = [vp, C.POINTER(Params), vp, sz, vp, sz, vp, vp, vp, vp]
    L.as_cigar_align_batch.argtypes = [vp, C.POINTER(Params), vp, sz, vp, vp, sz, vp, sz, vp, sz, vp, vp]
    L.as_edit_alignment_batch.argtypes = [vp, vp, sz, vp, vp, sz, vp, sz, vp, vp, vp, sz, vp]
    L.as_edit_distance_batch.argtypes = [vp, vp, sz, vp, vp, sz, vp, sz, vp, vp]
    L.as_edit_distance_device_noalloc.argtypes = [vp, vp, sz, vp, vp, sz, vp, sz, vp, vp]
    L.as_edit_distance_device_noalloc.restype = C.c_int
    L.as_host_alloc.argtypes = [sz]
    L.as_host_alloc.restype = vp
    L.as_host_free.argtypes = [vp]
    L.as_host_free.restype = None
    _cuda = L
    return L


def host_lib():
    global _host
    if _host is not None:
        return _host
    if not os.path.exists(HOST_LIB):
        raise AccSeqError(f"{HOST_LIB} is missing: run make -C longreadmapper_b200/csrc host")
    L = C.CDLL(HOST_LIB)
    vp, u64, sz = C.c_void_p, C.c_uint64, C.c_size_t
    L.as_sim_genome.argtypes = [u64, u64, vp]
    L.as_sim_genome.restype = None
    L.as_sim_plant_repeats.argtypes = [u64, vp, u64, C.c_int, u64, C.c_int, C.c_double]
    L.as_sim_plant_repeats.restype = None
    L.as_sim_reads.argtypes = [u64, vp, u64, sz, C.c_uint32, C.c_uint32, C.c_double, C.c_double, C.c_double, vp, sz,
                               vp, vp, vp]
    L.as_sim_reads.restype = sz
    L.as_index_load.argtypes = [C.c_char_p, C.POINTER(C.POINTER(AsHostIndex))]
    L.as_index_free.argtypes = [C.POINTER(AsHostIndex)]
    L.as_index_free.restype = None
    L.as_cat_build.argtypes = [C.c_int, C.POINTER(C.c_char_p), C.POINTER(u64), vp, C.POINTER(u64)]
    L.as_cat_build.restype = vp
    L.as_index_alloc.argtypes = [u64, C.c_int, C.c_int, C.c_int, C.c_int]
    L.as_index_alloc.restype = C.POINTER(AsHostIndex)
    L.as_index_out_of.argtypes = [C.POINTER(AsHostIndex), vp]
    L.as_index_out_of.restype = None
    L.as_index_write.argtypes = [C.c_char_p, C.POINTER(AsHostIndex)]
    L.as_pack_reads.argtypes = [vp, sz, vp, sz, vp, sz]
    L.as_result_from_compact.argtypes = [vp, C.c_uint32, vp, C.c_int, vp]
    L.as_result_from_compact.restype = None
    L.as_cigar_text.argtypes = [vp, C.c_int, C.c_char_p, sz]
    L.as_cigar_text_rle.argtypes = [vp, C.c_int, C.c_char_p, sz]
    _host = L
    return L


def device_count():
    return int(cuda_lib().as_device_count())


# ---------------------------------------------------------------------------------------------------
# synthetic inputs (distribution of the reference's scripts/dna_txt_gen.py:12, seeded)
# ---------------------------------------------------------------------------------------------------
def sim_genome(seed, n):
    g = np.empty(n, dtype=np.uint8)
    host_lib().as_sim_genome(seed, n, g.ctypes.data)
    return g


def sim_plant_repeats(seed, genome, n_families=8, unit_len=1500, copies=6, divergence=0.01):
    host_lib().as_sim_plant_repeats(seed, genome.ctypes.data, genome.size, n_families, unit_len, copies, divergence)
    return genome


def sim_reads(seed, genome, n_reads, len_min, len_max, err_min, err_max, sub_frac, out=None):
    """Returns (reads[n, stride] uint8 zero padded, lens uint32, truth_pos uint64, truth_strand uint8)."""
    stride = len_max + 1
    reads = out if out is not None else np.zeros((n_reads, stride), dtype=np.uint8)
    lens = np.zeros(n_reads, dtype=np.uint32)
    pos = np.zeros(n_reads, dtype=np.uint64)
    strand = np.zeros(n_reads, dtype=np.uint8)
    got = host_lib().as_sim_reads(seed, genome.ctypes.data, genome.size, n_reads, len_min, len_max, err_min, err_max,
                                  sub_frac, reads.ctypes.data, stride, lens.ctypes.data, pos.ctypes.data,
                                  strand.ctypes.data)
    if got != n_reads:
        raise AccSeqError("as_sim_reads: genome too short for these reads")
    return reads, lens, pos, strand


def pack_reads(reads, lens, wstride=None):
    """2-bit packing of a refactor_reads_seq buffer (uint8 [n, stride]) for as_map_batch_packed: uint32 [n, wstride]."""
    reads = np.ascontiguousarray(reads)
    lens = np.ascontiguousarray(lens, dtype=np.uint32)
    n, stride = reads.shape
    if wstride is None:
        wstride = (int(lens.max()) + 15) // 16 if n else 1
    out = np.zeros((n, wstride), dtype=np.uint32)
    rc = host_lib().as_pack_reads(reads.ctypes.data, stride, lens.ctypes.data, n, out.ctypes.data, wstride)
    if rc != 0:
        raise AccSeqError(f"as_pack_reads failed ({rc})")
    return out


def results_from_compact(rec, lens, mta):
    """as_result records (RESULT_DTYPE) recomputed from compact ones."""
    rec = np.ascontiguousarray(rec)
    out = np.zeros(rec.size, dtype=RESULT_DTYPE)
    marr = (AsMta * max(1, len(mta)))(*[AsMta(int(o), int(s)) for (o, s) in mta])
    H = host_lib()
    for i in range(rec.size):
        H.as_result_from_compact(rec[i:].ctypes.data, int(lens[i]), C.cast(marr, C.c_void_p), len(mta), out[i:].ctypes.data)
    return out


def cat_image(seqs):
    """The .cat text of a list of sequences (uint8 arrays / bytes): forward + reverse complement each, final '$'
    (asindex.c:97-110).  Returns (cat uint8[N], [(offset, seq_len)])."""
    comp = np.zeros(256, np.uint8)
    comp[:] = ord("N")
    for a, b in zip(b"ACGT", b"TGCA"):
        comp[a] = b
    parts, mta, off = [], [], 0
    for s in seqs:
        s = np.frombuffer(bytes(s), np.uint8) if not isinstance(s, np.ndarray) else s
        if s.size and int(s.max()) > 0x60:            # lower case -> upper case (asindex.c:63-67)
            s = np.where(s > 0x60, s - 0x20, s).astype(np.uint8)
        parts += [s, comp[s][::-1]]
        mta.append((off, int(s.size)))
        off += 2 * int(s.size)
    parts.append(np.frombuffer(b"$", np.uint8))
    return np.ascontiguousarray(np.concatenate(parts)), mta


def cigar_text_rle(runs):
    """CIGAR text of AS_F_RLE_CIGAR run bytes ((len << 2) | class)."""
    runs = np.ascontiguousarray(runs, dtype=np.uint8)
    buf = C.create_string_buffer(24 * runs.size + 16)
    n = host_lib().as_cigar_text_rle(runs.ctypes.data, runs.size, buf, len(buf))
    if n < 0:
        raise AccSeqError("bad run bytes")
    return buf.value.decode()


def cigar_text(ops):
    ops = np.ascontiguousarray(ops, dtype=np.uint8)
    buf = C.create_string_buffer(4 * ops.size + 16)
    n = host_lib().as_cigar_text(ops.ctypes.data, ops.size, buf, len(buf))
    if n < 0:
        raise AccSeqError("bad column ops")
    return buf.value.decode()


# ---------------------------------------------------------------------------------------------------
def _ptr(a):
    return a.ctypes.data if a is not None else None


class GpuMapper:
    """One as_gpu handle: the index resident on one B200, mirrors the calls alnmain.c makes per batch."""

    def __init__(self, device, *, length, C_table, O, o_ratio, bwt, lc, hlen, sa, sa_entry_bytes, content, mta, opts=None):
        L = cuda_lib()
        self._keep = [np.ascontiguousarray(x) for x in (C_table, O, bwt, lc, sa, content)]
        Ct, Ot, bw, lcv, sav, con = self._keep
        mta_arr = (AsMta * max(1, len(mta)))(*[AsMta(int(o), int(s)) for (o, s) in mta])
        v = AsIndexView(int(length), Ct.ctypes.data, Ot.ctypes.data, Ot.size, int(o_ratio), bw.ctypes.data,
                        lcv.ctypes.data, lcv.size, int(hlen), sav.ctypes.data, int(sa_entry_bytes), con.ctypes.data,
                        con.size, C.cast(mta_arr, C.c_void_p), len(mta))
        h = C.c_void_p()
        rc = L.as_gpu_open_opts(C.byref(h), device, C.byref(v), C.byref(opts) if opts is not None else None)
        if rc != 0:
            raise AccSeqError(f"as_gpu_open failed ({rc}): {L.as_last_error().decode()}")
        self.h = h
        self.L = L
        self.device = device
        self.N = int(length)
        self._keep = None          # the device copy is complete; host arrays are no longer referenced

    @classmethod
    def build(cls, seqs, device=0, hlen=12, names=None, opts=None):
        """Index construction on the GPU from raw sequences (the accidx path: asindex.c:129-153)."""
        cat, mta = cat_image(seqs)
        return cls.build_cat(cat, mta, device, hlen, names, opts)

    @classmethod
    def build_cat(cls, cat, mta, device=0, hlen=12, names=None, opts=None):
        """The same from a ready .cat image (uint8[N], forward + reverse complement per sequence, final '$') and its
        [(offset, seq_len)] table."""
        L = cuda_lib()
        cat = np.ascontiguousarray(cat, dtype=np.uint8)
        mta_arr = (AsMta * max(1, len(mta)))(*[AsMta(o, s) for (o, s) in mta])
        h = C.c_void_p()
        rc = L.as_gpu_build_opts(C.byref(h), device, cat.ctypes.data, cat.size, C.cast(mta_arr, C.c_void_p), len(mta), hlen,
                                 C.byref(opts) if opts is not None else None)
        if rc != 0:
            raise AccSeqError(f"as_gpu_build failed ({rc}): {L.as_last_error().decode()}")
        self = cls.__new__(cls)
        self.h, self.L, self.device, self.N, self._keep = h, L, device, int(cat.size), None
        self.names = list(names) if names else [f"seq{i}" for i in range(len(mta))]
        self.mta, self.hlen = mta, hlen
        return self

    def export(self, o_ratio=32, csa_ratio=4, hlen=None, sa_entry_bytes=8, want=("C", "O", "bwt", "lc", "sa", "csa")):
        """The index in the reference's in-memory layouts: dict of numpy arrays C, O, bwt, lc, sa, csa (those in `want`)."""
        hlen = hlen if hlen is not None else getattr(self, "hlen", 12)
        N = self.N
        shapes = dict(C=(256, np.uint64), O=(4 * (N // o_ratio + 1), np.uint64), bwt=(N, np.uint8), lc=(2 * 4 ** hlen, np.uint64),
                      sa=(N, np.uint64) if sa_entry_bytes == 8 else (N * 5, np.uint8), csa=(N // csa_ratio + 1, np.uint64))
        out = {k: np.zeros(shapes[k][0], shapes[k][1]) for k in want}
        ptr = lambda k: out[k].ctypes.data if k in out else None
        o = AsIndexOut(ptr("C"), ptr("O"), o_ratio, ptr("bwt"), ptr("lc"), ptr("sa"), sa_entry_bytes, ptr("csa"), csa_ratio)
        self._check(self.L.as_gpu_export(self.h, C.byref(o)), "as_gpu_export")
        return out

    @classmethod
    def from_index_files(cls, fasta_path, device=0, opts=None):
        """Loads <fa>.mta/.cat/.cat.mfi/.cat.lch/.cat.sa5 with the C loader (libaccseq_host.so) and uploads."""
        H = host_lib()
        L = cuda_lib()
        p = C.POINTER(AsHostIndex)()
        rc = H.as_index_load(fasta_path.encode(), C.byref(p))
        if rc != 0:
            raise AccSeqError(f"as_index_load({fasta_path}) failed: {rc}")
        try:
            h = C.c_void_p()
            rc = L.as_gpu_open_opts(C.byref(h), device, C.byref(p.contents.view), C.byref(opts) if opts is not None else None)
            if rc != 0:
                raise AccSeqError(f"as_gpu_open failed ({rc}): {L.as_last_error().decode()}")
            names = [p.contents.names[i].decode() for i in range(p.contents.view.mta_len)]
            n = int(p.contents.view.length)
        finally:
            H.as_index_free(p)
        self = cls.__new__(cls)
        self.h, self.L, self.device, self.N, self._keep, self.names = h, L, device, n, None, names
        return self

    def _check(self, rc, what):
        if rc != 0:
            raise AccSeqError(f"{what} failed ({rc}): {self.L.as_last_error().decode()}")

    def close(self):
        if getattr(self, "h", None):
            self.L.as_gpu_close(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def params(self, seed_len=20, thres=300, T=None, O=None, flags=0, band=None, **knobs):
        """as_params; knobs = vote_ectx, gact_order, gact_blocks_per_sm, chunk_reads, chunk_tail (scheduling only)."""
        p = Params()
        self.L.as_params_default(C.byref(p))
        p.seed_len, p.thres, p.flags = seed_len, thres, flags
        for k, v in knobs.items():
            if k not in ("vote_ectx", "gact_order", "gact_blocks_per_sm", "chunk_reads", "chunk_tail", "vote_thread", "gact_runs", "vote_blocks_per_sm"):
                raise TypeError(f"unknown as_params field {k}")
            setattr(p, k, int(v))
        if band is not None:
            p.gact_band = band
        if T is not None:
            p.gact_T = T
        if O is not None:
            p.gact_O = O
        return p

    # -- whole path -------------------------------------------------------------------------------
    def map_batch(self, reads, lens, params=None, want_ops=True, store_stride=None):
        """reads: uint8 [n, stride] (strand-1 reads are reverse-complemented in place). Returns (results, ops)."""
        p = params or self.params()
        reads = np.ascontiguousarray(reads)
        assert reads.dtype == np.uint8 and reads.ndim == 2
        n, stride = reads.shape
        lens = np.ascontiguousarray(lens, dtype=np.uint32)
        res = np.zeros(n, dtype=RESULT_DTYPE)
        ops = None
        if want_ops:
            if store_stride is None:
                store_stride = 2 * int(lens.max()) if n else 2
            ops = np.zeros((n, store_stride), dtype=np.uint8)
        self._check(self.L.as_map_batch(self.h, C.byref(p), reads.ctypes.data, stride, lens.ctypes.data, n,
                                        res.ctypes.data, _ptr(ops), store_stride or 0), "as_map_batch")
        return res, ops

    def map_batch_packed(self, packed, lens, params=None, runs_cap=None):
        """packed: uint32 [n, wstride] (pack_reads). Returns (compact records, dense run bytes)."""
        p = params or self.params()
        packed = np.ascontiguousarray(packed, dtype=np.uint32)
        n, wstride = packed.shape
        lens = np.ascontiguousarray(lens, dtype=np.uint32)
        rec = np.zeros(n, dtype=COMPACT_DTYPE)
        if runs_cap is None:
            runs_cap = int(2 * lens.astype(np.int64).sum()) + 16
        runs = np.zeros(runs_cap, dtype=np.uint8)
        got = C.c_uint64(0)
        rc = self.L.as_map_batch_packed(self.h, C.byref(p), packed.ctypes.data, wstride, lens.ctypes.data, n, rec.ctypes.data,
                                        runs.ctypes.data, runs_cap, C.byref(got))
        if rc != 0:
            raise AccSeqError(f"as_map_batch_packed failed ({rc}): {self.L.as_last_error().decode()}")
        return rec, runs[: int(got.value)]

    def map_batch_device(self, d_reads_ptr, stride, d_lens_ptr, n, max_len, d_res_ptr, d_store_ptr, store_stride,
                         params=None):
        p = params or self.params()
        self._check(self.L.as_map_batch_device(self.h, C.byref(p), d_reads_ptr, stride, d_lens_ptr, n, max_len,
                                               d_res_ptr, d_store_ptr, store_stride), "as_map_batch_device")

    def set_stream(self, cuda_stream_ptr):
        self._check(self.L.as_gpu_set_stream(self.h, cuda_stream_ptr), "as_gpu_set_stream")

    def sync(self):
        self._check(self.L.as_gpu_sync(self.h), "as_gpu_sync")

    def launch_count(self):
        return int(self.L.as_gpu_launch_count(self.h))

    def stats(self):
        s = AsStats()
        self._check(self.L.as_gpu_stats(self.h, C.byref(s)), "as_gpu_stats")
        return {k: int(getattr(s, k)) for k, _ in AsStats._fields_}

    def stats_reset(self):
        self._check(self.L.as_gpu_stats_reset(self.h), "as_gpu_stats_reset")

    def last_timing(self):
        ms = (C.c_float * 8)()
        self._check(self.L.as_gpu_last_timing(self.h, C.byref(ms)), "as_gpu_last_timing")
        k = ("part1", "part2", "gact", "myers", "unused", "h2d", "d2h", "total")
        return dict(zip(k, [float(x) for x in ms]))

    def last_host_timing(self):
        us = (C.c_double * 3)()
        self._check(self.L.as_gpu_last_host_timing(self.h, C.byref(us)), "as_gpu_last_host_timing")
        return dict(zip(("enqueue_us", "wait_us", "events_us"), [float(x) for x in us]))

    def last_transfer(self):
        a, b = C.c_uint64(0), C.c_uint64(0)
        self._check(self.L.as_gpu_last_transfer(self.h, C.byref(a), C.byref(b)), "as_gpu_last_transfer")
        return int(a.value), int(b.value)

    def last_reads_done(self):
        """(done, touched) of the last map_batch call: reads [0, done) were delivered, reads [touched, n) are untouched."""
        t = C.c_size_t(0)
        d = self.L.as_gpu_last_reads_done(self.h, C.byref(t))
        return int(d), int(t.value)

    # -- single stages ----------------------------------------------------------------------------
    def _aln(self, fn, seeds, qlen):
        seeds = np.ascontiguousarray(seeds, dtype=np.uint8)
        n = seeds.size // qlen if qlen else 0
        k = np.zeros(n, np.uint64); l = np.zeros(n, np.uint64); r = np.zeros(n, np.uint64)
        self._check(fn(self.h, seeds.ctypes.data, qlen, n, k.ctypes.data, l.ctypes.data, r.ctypes.data), "aln_batch")
        return k, l, r

    def lc_aln_batch(self, seeds, qlen):
        return self._aln(self.L.as_lc_aln_batch, seeds, qlen)

    def fmi_aln_batch(self, seeds, qlen):
        return self._aln(self.L.as_fmi_aln_batch, seeds, qlen)

    def sa_access_batch(self, rows):
        rows = np.ascontiguousarray(rows, dtype=np.uint64)
        pos = np.zeros(rows.size, np.uint64)
        self._check(self.L.as_sa_access_batch(self.h, rows.ctypes.data, rows.size, pos.ctypes.data), "as_sa_access_batch")
        return pos

    def csa_access_batch(self, rows, ratio=4, literal=False):
        """csa_access (fmidx.c:315-331) on a batch of rows."""
        rows = np.ascontiguousarray(rows, dtype=np.uint64)
        pos = np.zeros(rows.size, np.uint64)
        self._check(self.L.as_csa_access_batch(self.h, ratio, int(literal), rows.ctypes.data, rows.size, pos.ctypes.data),
                    "as_csa_access_batch")
        return pos

    def vote_batch(self, reads, lens, params=None):
        p = params or self.params()
        reads = np.ascontiguousarray(reads)
        n, stride = reads.shape
        lens = np.ascontiguousarray(lens, dtype=np.uint32)
        key = np.zeros(n, np.uint64); val = np.zeros(n, np.uint64)
        ph = np.zeros(n, np.int32); fo = np.zeros(n, np.int32)
        self._check(self.L.as_vote_batch(self.h, C.byref(p), reads.ctypes.data, stride, lens.ctypes.data, n,
                                         key.ctypes.data, val.ctypes.data, ph.ctypes.data, fo.ctypes.data), "as_vote_batch")
        return key, val, ph, fo

    @staticmethod
    def _pairs(seqs):
        lens = np.array([len(s) for s in seqs], dtype=np.uint32)
        stride = int(lens.max()) + 1 if len(seqs) else 1
        buf = np.zeros((len(seqs), stride), dtype=np.uint8)
        for i, s in enumerate(seqs):
            buf[i, : len(s)] = np.frombuffer(bytes(s), dtype=np.uint8) if not isinstance(s, np.ndarray) else s
        return buf, stride, lens

    def cigar_align_batch(self, queries, targets, params=None):
        """queries/targets: lists of bytes / uint8 arrays. Returns (scores, [ops arrays])."""
        p = params or self.params()
        n = len(queries)
        qb, qs, ql = self._pairs(queries)
        tb, tstr, tl = self._pairs(targets)
        ostride = int(ql.max() + tl.max()) + 2 if n else 2
        ops = np.zeros((n, ostride), dtype=np.uint8)
        n_ops = np.zeros(n, np.int32); score = np.zeros(n, np.int32)
        self._check(self.L.as_cigar_align_batch(self.h, C.byref(p), qb.ctypes.data, qs, ql.ctypes.data, tb.ctypes.data, tstr,
                                                tl.ctypes.data, n, ops.ctypes.data, ostride, n_ops.ctypes.data,
                                                score.ctypes.data), "as_cigar_align_batch")
        return score, [ops[i, : n_ops[i]].copy() for i in range(n)]

    def edit_distance_batch(self, queries, targets, ks=None):
        n = len(queries)
        qb, qs, ql = self._pairs(queries)
        tb, tstr, tl = self._pairs(targets)
        dist = np.zeros(n, np.int32)
        kk = None if ks is None else np.ascontiguousarray(ks, dtype=np.int32)
        self._check(self.L.as_edit_distance_batch(self.h, qb.ctypes.data, qs, ql.ctypes.data, tb.ctypes.data, tstr,
                                                  tl.ctypes.data, n, _ptr(kk), dist.ctypes.data), "as_edit_distance_batch")
        return dist

    def edit_alignment_batch(self, queries, targets, ks=None):
        """edlib's NW alignment path (EDLIB_TASK_PATH): (dist int32[n], list of uint8 op arrays in forward order)"""
        n = len(queries)
        qb, qs, ql = self._pairs(queries)
        tb, tstr, tl = self._pairs(targets)
        dist, n_ops = np.zeros(n, np.int32), np.zeros(n, np.int32)
        stride = int((ql.astype(np.int64) + tl.astype(np.int64)).max()) if n else 1
        stride = max(16, (stride + 15) & ~15)
        ops = np.zeros((n, stride), np.uint8)
        kk = None if ks is None else np.ascontiguousarray(ks, dtype=np.int32)
        self._check(self.L.as_edit_alignment_batch(self.h, qb.ctypes.data, qs, ql.ctypes.data, tb.ctypes.data, tstr, tl.ctypes.data,
                                                   n, _ptr(kk), dist.ctypes.data, ops.ctypes.data, stride, n_ops.ctypes.data),
                    "as_edit_alignment_batch")
        return dist, [ops[i, : n_ops[i]].copy() for i in range(n)]
