/*
 * as_index.cu -- upload of the reference's in-memory index (dna_fmi / lc_hash / sa_buf / content / mta,
 * as alnmain.c:179-256 init() leaves them) and its repack into the device layouts of as_common.cuh.
 * The repack runs on the GPU, chunk by chunk, so a human-scale index never needs a second host copy.
 */
#include "as_common.cuh"

#include <stdarg.h>
#include <stdlib.h>
#include <new>

static thread_local char g_err[512] = "";

void as_set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof g_err, fmt, ap);
    va_end(ap);
}

extern "C" const char *as_last_error(void) { return g_err; }

extern "C" int as_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}

int as_scratch(as_gpu *h, ScratchBuf *b, size_t bytes) {
    if (bytes <= b->cap && b->p) return AS_OK;
    if (b->p) {
        AS_CUDA(cudaStreamSynchronize(h->stream));
        AS_CUDA(cudaFree(b->p));
        b->p = nullptr;
        b->cap = 0;
    }
    size_t want = bytes + bytes / 8 + 256;
    AS_CUDA(cudaMalloc(&b->p, want));
    b->cap = want;
    return AS_OK;
}

int as_pinned(as_gpu *h, size_t bytes) {
    if (bytes <= h->h_pin_cap && h->h_pin) return AS_OK;
    if (h->h_pin) {
        AS_CUDA(cudaStreamSynchronize(h->stream));
        AS_CUDA(cudaFreeHost(h->h_pin));
        h->h_pin = nullptr;
    }
    size_t want = bytes + bytes / 8 + 4096;
    AS_CUDA(cudaMallocHost(&h->h_pin, want));
    h->h_pin_cap = want;
    return AS_OK;
}

/* ---------------------------------------------------------------------------------------------- */
/* repack kernels                                                                                  */
/* ---------------------------------------------------------------------------------------------- */

/* rows [row0, row0 + nrows) of the BWT, one thread per 64-row block.  `bwt` points at row0 (64-byte
 * aligned chunk start); `Oc` points at the O entries of row0 (4 u64 per o_ratio rows). */
__global__ void k_build_fm_blocks(const uint8_t *__restrict__ bwt, uint64_t row0, uint64_t nrows,
                                  const uint64_t *__restrict__ Oc, int o_ratio, const uint64_t *__restrict__ super,
                                  FmBlock *__restrict__ out, unsigned long long *primary) {
    uint64_t nblk = (nrows + 63) >> 6;
    for (uint64_t b = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; b < nblk; b += (uint64_t)gridDim.x * blockDim.x) {
        uint64_t lo = 0, hi = 0;
        uint64_t r0 = b << 6;
        uint32_t lim = (uint32_t)((nrows - r0 < 64) ? nrows - r0 : 64);
        for (uint32_t r = 0; r < lim; r++) {
            uint32_t ch = bwt[r0 + r];
            uint32_t code;
            if (ch == '$') {
                code = 0;
                *primary = row0 + r0 + r;
            } else {
                code = as_base_code(ch);
            }
            lo |= (uint64_t)(code & 1u) << r;
            hi |= (uint64_t)(code >> 1) << r;
        }
        uint64_t row = row0 + r0;
        const uint64_t *o = Oc + 4 * (r0 / (uint64_t)o_ratio);
        const uint64_t *sb = super + 4 * (row >> AS_SB_SHIFT);
        /* o_ratio > 64 (the reference's tests build with 128): the checkpoint lies up to o_ratio - 64 rows before the
         * block; count the rows in between (chunks start at multiples of o_ratio, so they are in `bwt`) */
        uint32_t extra[4] = {0, 0, 0, 0};
        for (uint64_t r = r0 & ~((uint64_t)o_ratio - 1); r < r0; r++) {
            const uint32_t ch = bwt[r];
            if (ch != '$') extra[as_base_code(ch)]++;
        }
        FmBlock blk;
        blk.cnt[0] = (uint32_t)(o[0] - sb[0]) + extra[0];
        blk.cnt[1] = (uint32_t)(o[1] - sb[1]) + extra[1];
        blk.cnt[2] = (uint32_t)(o[2] - sb[2]) + extra[2];
        blk.cnt[3] = (uint32_t)(o[3] - sb[3]) + extra[3];
        blk.lo = lo;
        blk.hi = hi;
        out[row >> 6] = blk;
    }
}

/* (k,l) pairs from MSB-first code order (lchash.c:36-49) to the LSB-first order the packed reads give */
__global__ void k_permute_lc(const ulonglong2 *__restrict__ in, uint64_t code0, uint64_t ncodes, int hlen,
                             ulonglong2 *__restrict__ out) {
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < ncodes; i += (uint64_t)gridDim.x * blockDim.x) {
        uint64_t code = code0 + i, rev = 0, c = code;
        for (int t = 0; t < hlen; t++) {
            rev = (rev << 2) | (c & 3u);
            c >>= 2;
        }
        out[rev] = in[i];
    }
}

__global__ void k_split_sa(const uint8_t *__restrict__ in, int entry_bytes, uint64_t n, uint32_t *__restrict__ lo,
                           uint8_t *__restrict__ hi) {
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
        const uint8_t *p = in + i * (uint64_t)entry_bytes;
        uint32_t v = (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24);
        lo[i] = v;
        if (hi) hi[i] = entry_bytes > 4 ? p[4] : 0;   /* bytes 5..7 of an 8-byte ui40_t are padding (sa_use.h:17-20) */
    }
}

/* text bytes [base0, base0+n) -> 2-bit words; base0 is a multiple of 16 */
__global__ void k_pack_text(const uint8_t *__restrict__ in, uint64_t n, uint32_t *__restrict__ out) {
    uint64_t nw = (n + 15) >> 4;
    for (uint64_t w = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; w < nw; w += (uint64_t)gridDim.x * blockDim.x) {
        uint32_t v = 0;
        uint64_t b0 = w << 4;
        uint32_t lim = (uint32_t)((n - b0 < 16) ? n - b0 : 16);
        for (uint32_t t = 0; t < lim; t++) {
            uint32_t ch = in[b0 + t];
            uint32_t code = (ch == '$') ? 0u : as_base_code(ch);
            v |= code << (2 * t);
        }
        out[w] = v;
    }
}

/* ---------------------------------------------------------------------------------------------- */

static int upload_chunk(void *d_dst, const void *h_src, size_t bytes, cudaStream_t s) {
    AS_CUDA(cudaMemcpyAsync(d_dst, h_src, bytes, cudaMemcpyHostToDevice, s));
    return AS_OK;
}

static void free_index(as_gpu *h) {
    void **ptrs[] = {&h->d_lc2, &h->d_sax, &h->d_fm, &h->d_super, &h->d_lc, &h->d_sa_lo, &h->d_sa_hi, &h->d_text2, &h->d_mta,
                     &h->d_csa_lo, &h->d_csa_hi};
    h->csa_ratio = 0;
    for (void **p : ptrs) {
        if (*p) cudaFree(*p);
        *p = nullptr;
    }
}

static void free_scratch(ScratchBuf *b) {
    if (b->p) cudaFree(b->p);
    b->p = nullptr;
    b->cap = 0;
}

extern "C" void as_gpu_close(as_gpu *h) {
    if (!h) return;
    cudaSetDevice(h->device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    if (h->twin) {
        as_gpu_close(h->twin);
        h->twin = nullptr;
    }
    if (!h->is_twin) free_index(h);
    ScratchBuf *bufs[] = {&h->reads, &h->lens, &h->results, &h->store, &h->qpack, &h->tpack, &h->toff,
                          &h->vote, &h->counters, &h->ovf_list, &h->ovf_table, &h->gact_planes, &h->gact_tb, &h->gact_order, &h->gact_retry, &h->path_planes, &h->vote_dlist,
                          &h->locus_tmp, &h->pack_in, &h->cres, &h->runs, &h->run_tot};
    for (ScratchBuf *b : bufs) free_scratch(b);
    for (ScratchBuf &b : h->misc) free_scratch(&b);
    if (h->h_pin) cudaFreeHost(h->h_pin);
    if (h->d_stats) cudaFree(h->d_stats);
    for (cudaEvent_t &e : h->ev)
        if (e) cudaEventDestroy(e);
    for (auto &row : h->cev)
        for (cudaEvent_t &e : row)
            if (e) cudaEventDestroy(e);
    if (h->ev_d2h0) cudaEventDestroy(h->ev_d2h0);
    if (h->d_chunk_max) cudaFree(h->d_chunk_max);
    if (h->h_chunk_max) cudaFreeHost(h->h_chunk_max);
    if (h->h_run_tot) cudaFreeHost(h->h_run_tot);
    if (h->h_band) cudaFreeHost(h->h_band);
    if (h->s_h2d) cudaStreamDestroy(h->s_h2d);
    if (h->s_d2h) cudaStreamDestroy(h->s_d2h);
    if (h->s_d2h2) cudaStreamDestroy(h->s_d2h2);
    if (h->own_stream) cudaStreamDestroy(h->own_stream);
    delete h;
}

static int build_index(as_gpu *h, const as_index_view *v) {
    cudaStream_t s = h->stream;
    const uint64_t N = v->length;
    DevIndex &ix = h->ix;
    ix.N = N;
    ix.hlen = v->hlen;
    ix.C[0] = v->C['A'];
    ix.C[1] = v->C['C'];
    ix.C[2] = v->C['G'];
    ix.C[3] = v->C['T'];

    const size_t CH = (size_t)1 << 26;   /* rows per upload chunk */
    void *d_stage = nullptr, *d_stage2 = nullptr;
    unsigned long long *d_primary = nullptr;
    int rc = AS_OK;
#define STEP(call)                 \
    do {                           \
        rc = (call);               \
        if (rc != AS_OK) goto out; \
    } while (0)
#define CU(call)                                                                                        \
    do {                                                                                                \
        cudaError_t e_ = (call);                                                                        \
        if (e_ != cudaSuccess) {                                                                        \
            as_set_error("%s:%d: %s failed: %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_));     \
            rc = (e_ == cudaErrorMemoryAllocation) ? AS_ENOMEM : AS_ECUDA;                              \
            goto out;                                                                                   \
        }                                                                                               \
    } while (0)

    {
        /* ---- FM blocks ---- */
        uint64_t nblk = (N + 63) >> 6;
        uint64_t nsuper = ((N - 1) >> AS_SB_SHIFT) + 1;
        uint64_t *h_super = (uint64_t *)malloc(nsuper * 4 * sizeof(uint64_t));
        if (!h_super) return AS_ENOMEM;
        for (uint64_t sb = 0; sb < nsuper; sb++) {
            uint64_t row = sb << AS_SB_SHIFT;
            memcpy(h_super + 4 * sb, v->O + 4 * (row / (uint64_t)v->o_ratio), 4 * sizeof(uint64_t));
        }
        cudaError_t e1 = cudaMalloc(&h->d_super, nsuper * 4 * sizeof(uint64_t));
        if (e1 == cudaSuccess) e1 = cudaMemcpy(h->d_super, h_super, nsuper * 4 * sizeof(uint64_t), cudaMemcpyHostToDevice);
        free(h_super);
        CU(e1);
        CU(cudaMalloc(&h->d_fm, nblk * sizeof(FmBlock)));
        CU(cudaMalloc(&d_primary, sizeof(unsigned long long)));
        CU(cudaMemset(d_primary, 0xff, sizeof(unsigned long long)));
        size_t o_per_chunk = 4 * (CH / (size_t)v->o_ratio);
        CU(cudaMalloc(&d_stage, CH * 8));                           /* also reused for lc / SA / text below */
        CU(cudaMalloc(&d_stage2, (o_per_chunk + 4) * sizeof(uint64_t)));
        for (uint64_t row0 = 0; row0 < N; row0 += CH) {
            uint64_t nrows = (N - row0 < CH) ? N - row0 : CH;
            uint64_t o_first = 4 * (row0 / (uint64_t)v->o_ratio);
            uint64_t o_cnt = 4 * ((nrows - 1) / (uint64_t)v->o_ratio + 1);
            STEP(upload_chunk(d_stage, v->bwt + row0, nrows, s));
            STEP(upload_chunk(d_stage2, v->O + o_first, o_cnt * sizeof(uint64_t), s));
            uint64_t cb = (nrows + 63) >> 6;
            int grid = (int)((cb + 255) / 256 < 148 * 16 ? (cb + 255) / 256 : 148 * 16);
            k_build_fm_blocks<<<grid, 256, 0, s>>>((const uint8_t *)d_stage, row0, nrows, (const uint64_t *)d_stage2,
                                                   v->o_ratio, (const uint64_t *)h->d_super, (FmBlock *)h->d_fm, d_primary);
            CU(cudaGetLastError());
            CU(cudaStreamSynchronize(s));
        }
        unsigned long long prim = 0;
        CU(cudaMemcpy(&prim, d_primary, sizeof prim, cudaMemcpyDeviceToHost));
        if (prim == ~0ULL) {
            as_set_error("as_gpu_open: BWT holds no '$' row");
            rc = AS_EINVAL;
            goto out;
        }
        ix.primary = prim;
        ix.fm = (const FmBlock *)h->d_fm;
        ix.super = (const uint64_t *)h->d_super;

        /* ---- lc table (absent in a lookup-only view: fmi_aln / csa_access still work) ---- */
        uint64_t ncodes = v->lc ? 1ULL << (2 * v->hlen) : 0;
        if (v->lc) CU(cudaMalloc(&h->d_lc, ncodes * sizeof(ulonglong2)));
        uint64_t per = (CH * 8) / sizeof(ulonglong2);
        for (uint64_t c0 = 0; c0 < ncodes; c0 += per) {
            uint64_t nc = (ncodes - c0 < per) ? ncodes - c0 : per;
            STEP(upload_chunk(d_stage, v->lc + 2 * c0, nc * sizeof(ulonglong2), s));
            int grid = (int)((nc + 255) / 256 < 148 * 16 ? (nc + 255) / 256 : 148 * 16);
            k_permute_lc<<<grid, 256, 0, s>>>((const ulonglong2 *)d_stage, c0, nc, v->hlen, (ulonglong2 *)h->d_lc);
            CU(cudaGetLastError());
            CU(cudaStreamSynchronize(s));
        }
        ix.lc = (const ulonglong2 *)h->d_lc;
        if (!v->lc) ix.hlen = 0;

        /* ---- suffix array (may be absent) ---- */
        if (v->sa) CU(cudaMalloc(&h->d_sa_lo, N * sizeof(uint32_t)));
        if (v->sa && N > (1ULL << 32)) CU(cudaMalloc(&h->d_sa_hi, N));
        for (uint64_t r0 = 0; v->sa && r0 < N; r0 += CH) {
            uint64_t nr = (N - r0 < CH) ? N - r0 : CH;
            STEP(upload_chunk(d_stage, (const uint8_t *)v->sa + r0 * (uint64_t)v->sa_entry_bytes,
                              nr * (size_t)v->sa_entry_bytes, s));
            int grid = (int)((nr + 255) / 256 < 148 * 16 ? (nr + 255) / 256 : 148 * 16);
            k_split_sa<<<grid, 256, 0, s>>>((const uint8_t *)d_stage, v->sa_entry_bytes, nr, (uint32_t *)h->d_sa_lo + r0,
                                            h->d_sa_hi ? (uint8_t *)h->d_sa_hi + r0 : nullptr);
            CU(cudaGetLastError());
            CU(cudaStreamSynchronize(s));
        }
        ix.sa_lo = (const uint32_t *)h->d_sa_lo;
        ix.sa_hi = (const uint8_t *)h->d_sa_hi;

        /* ---- text (may be absent) ---- */
        uint64_t nwords = ((v->con_len + 15) >> 4) + 4;          /* slack for the funnel-shift readers */
        if (v->content) {
            CU(cudaMalloc(&h->d_text2, nwords * sizeof(uint32_t)));
            CU(cudaMemset(h->d_text2, 0, nwords * sizeof(uint32_t)));
        }
        for (uint64_t b0 = 0; v->content && b0 < v->con_len; b0 += CH) {
            uint64_t nb = (v->con_len - b0 < CH) ? v->con_len - b0 : CH;
            STEP(upload_chunk(d_stage, v->content + b0, nb, s));
            uint64_t nw = (nb + 15) >> 4;
            int grid = (int)((nw + 255) / 256 < 148 * 16 ? (nw + 255) / 256 : 148 * 16);
            k_pack_text<<<grid, 256, 0, s>>>((const uint8_t *)d_stage, nb, (uint32_t *)h->d_text2 + (b0 >> 4));
            CU(cudaGetLastError());
            CU(cudaStreamSynchronize(s));
        }
        ix.text2 = (const uint32_t *)h->d_text2;

        /* ---- mta ---- */
        size_t mb = (size_t)(v->mta_len > 0 ? v->mta_len : 1) * sizeof(as_mta);
        CU(cudaMalloc(&h->d_mta, mb));
        if (v->mta_len > 0) CU(cudaMemcpy(h->d_mta, v->mta, (size_t)v->mta_len * sizeof(as_mta), cudaMemcpyHostToDevice));
        ix.mta = (const as_mta *)h->d_mta;
        ix.mta_len = v->mta_len;
    }
out:
    if (d_stage) cudaFree(d_stage);
    if (d_stage2) cudaFree(d_stage2);
    if (d_primary) cudaFree(d_primary);
#undef STEP
#undef CU
    return rc;
}

/* device checks, stream, events, stats buffer; NULL (with as_last_error set) when the device is unusable */
as_gpu *as_handle_new(int device) {
    int ndev = as_device_count();
    if (ndev <= 0 || device < 0 || device >= ndev) {
        as_set_error("as_gpu_open: no usable CUDA device %d (found %d); there is no CPU fallback", device, ndev);
        return nullptr;
    }
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) {
        as_set_error("as_gpu_open: cudaGetDeviceProperties(%d) failed", device);
        return nullptr;
    }
    if (prop.major != 10) {
        as_set_error("as_gpu_open: device %d is sm_%d%d; this library is built for sm_100a only", device, prop.major, prop.minor);
        return nullptr;
    }
    if (cudaSetDevice(device) != cudaSuccess) {
        as_set_error("as_gpu_open: cudaSetDevice(%d) failed", device);
        return nullptr;
    }
    as_gpu *h = new (std::nothrow) as_gpu();
    if (!h) return nullptr;
    h->device = device;
    h->sm_count = prop.multiProcessorCount;
    h->l2_bytes = prop.l2CacheSize > 0 ? (size_t)prop.l2CacheSize : h->l2_bytes;
    cudaError_t e = cudaStreamCreateWithFlags(&h->own_stream, cudaStreamNonBlocking);
    if (e != cudaSuccess) {
        as_set_error("cudaStreamCreate: %s", cudaGetErrorString(e));
        delete h;
        return nullptr;
    }
    h->stream = h->own_stream;
    for (cudaEvent_t &ev : h->ev) cudaEventCreate(&ev);
    if (cudaMalloc((void **)&h->d_stats, 24 * sizeof(unsigned long long)) != cudaSuccess ||
        cudaMemset(h->d_stats, 0, 24 * sizeof(unsigned long long)) != cudaSuccess) {
        as_set_error("as_gpu_open: cudaMalloc(stats) failed");
        as_gpu_close(h);
        return nullptr;
    }
    return h;
}

int as_launch_split_sa(as_gpu *h, const uint8_t *d_in, int entry_bytes, uint64_t n, uint32_t *d_lo, uint8_t *d_hi) {
    int grid = (int)((n + 255) / 256 < 148 * 16 ? (n + 255) / 256 : 148 * 16);
    k_split_sa<<<grid ? grid : 1, 256, 0, h->stream>>>(d_in, entry_bytes, n, d_lo, d_hi);
    h->launches++;
    AS_CUDA(cudaGetLastError());
    return AS_OK;
}

/* a handle without an index: the aligners on explicit pairs (as_cigar_align_batch, as_edit_distance_batch) need none */
extern "C" int as_gpu_open_aligner(as_gpu **out, int device) {
    if (!out) return AS_EINVAL;
    *out = as_handle_new(device);
    return *out ? AS_OK : AS_ENODEV;
}

int as_launch_pack_text(as_gpu *h, const uint8_t *d_in, uint64_t n, uint32_t *d_out) {
    uint64_t nw = (n + 15) >> 4;
    int grid = (int)((nw + 255) / 256 < 148 * 16 ? (nw + 255) / 256 : 148 * 16);
    k_pack_text<<<grid ? grid : 1, 256, 0, h->stream>>>(d_in, n, d_out);
    h->launches++;
    AS_CUDA(cudaGetLastError());
    return AS_OK;
}

extern "C" int as_gpu_open(as_gpu **out, int device, const as_index_view *v) {
    return as_gpu_open_opts(out, device, v, nullptr);
}

extern "C" int as_gpu_open_opts(as_gpu **out, int device, const as_index_view *v, const as_open_opts *opts) {
    if (!out || !v) return AS_EINVAL;
    *out = nullptr;
    /* C, O and the BWT are the index; lc / sa / content may be NULL (a lookup-only handle: the scalar wrappers of
     * libaccseq_compat open one from a bare dna_fmi) -- the entries that need a missing table refuse to run */
    if (!v->C || !v->O || !v->bwt || v->length < 2 || v->mta_len < 0 || (v->mta_len > 0 && !v->mta)) {
        as_set_error("as_gpu_open: incomplete index view");
        return AS_EINVAL;
    }
    if (v->o_ratio <= 0 || (v->o_ratio & (v->o_ratio - 1)) != 0 || v->o_ratio > (1 << 20)) {
        as_set_error("as_gpu_open: o_ratio %d must be a power of two <= 2^20 (the reference builds with 32, asindex.c:138)", v->o_ratio);
        return AS_EINVAL;
    }
    if (v->lc && (v->hlen < 1 || v->hlen > 14 || v->lc_len != (2ULL << (2 * v->hlen)))) {
        as_set_error("as_gpu_open: hlen %d / lc_len %llu inconsistent", v->hlen, (unsigned long long)v->lc_len);
        return AS_EINVAL;
    }
    if (v->sa && v->sa_entry_bytes != 4 && v->sa_entry_bytes != 5 && v->sa_entry_bytes != 8) {
        as_set_error("as_gpu_open: sa_entry_bytes must be 4, 5 or 8");
        return AS_EINVAL;
    }
    if (v->length >= (1ULL << 40) || (v->content && v->con_len < v->length - 1)) {
        as_set_error("as_gpu_open: bad length/con_len");
        return AS_EINVAL;
    }
    /* seq_lookup here is a binary search (the reference scans linearly and takes the first match, alnmain.c:151-176):
     * the entries must be sorted by offset and tile the text the way create_meta lays them out (asindex.c:89-93) */
    for (int i = 0; i < v->mta_len; i++) {
        const uint64_t want = i == 0 ? 0 : v->mta[i - 1].offset + 2 * v->mta[i - 1].seq_len;
        if (v->mta[i].offset != want || v->mta[i].offset + 2 * v->mta[i].seq_len > v->length) {
            as_set_error("as_gpu_open: mta entry %d (offset %llu, seq_len %llu) does not tile the .cat text", i,
                         (unsigned long long)v->mta[i].offset, (unsigned long long)v->mta[i].seq_len);
            return AS_EINVAL;
        }
    }
    as_gpu *h = as_handle_new(device);
    if (!h) return AS_ENODEV;
    if (opts) h->opts = *opts;
    int rc = build_index(h, v);
    if (rc == AS_OK) rc = as_build_lc2(h);
    if (rc != AS_OK) {
        as_gpu_close(h);
        return rc;
    }
    *out = h;
    return AS_OK;
}

extern "C" int as_gpu_set_stream(as_gpu *h, void *cuda_stream) {
    if (!h) return AS_EINVAL;
    AS_CUDA(cudaSetDevice(h->device));
    AS_CUDA(cudaStreamSynchronize(h->stream));
    h->stream = cuda_stream ? (cudaStream_t)cuda_stream : h->own_stream;
    return AS_OK;
}

extern "C" void *as_host_alloc(size_t bytes) {
    void *p = nullptr;
    if (cudaMallocHost(&p, bytes ? bytes : 1) != cudaSuccess) {
        cudaGetLastError();
        as_set_error("as_host_alloc(%zu): pinned allocation failed", bytes);
        return nullptr;
    }
    return p;
}

extern "C" void as_host_free(void *p) {
    if (p) cudaFreeHost(p);
}
