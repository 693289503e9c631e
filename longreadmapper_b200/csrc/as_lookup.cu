/*
 * as_lookup.cu -- read packing and the batched forms of the reference's lookup primitives:
 *   lc_aln   lchash.c:89-104  (+ lc_access :12-16, _num_from_seq :36-49)
 *   fmi_aln  fmidx.c:295-313  (+ _occ_access :277-293)
 *   sa_access fmidx.c:18-33
 *   csa_access fmidx.c:315-331 (+ _build_csa :152-163)
 * One thread per query; every lookup is a chain of dependent 32-byte sector gathers, so the kernels are
 * sized for memory-level parallelism (many resident warps), not arithmetic.
 */
#include "as_common.cuh"

/* ---------------------------------------------------------------------------------------------- */
/* 2-bit packing of a read batch (refactor_reads_seq layout in, alnmain.c:87-103)                  */
/* ---------------------------------------------------------------------------------------------- */
__global__ void __launch_bounds__(256) k_pack_reads(const char *__restrict__ ascii, size_t stride,
                                                    const uint32_t *__restrict__ lens, size_t n,
                                                    uint32_t *__restrict__ pack, size_t wstride) {
    /* one thread per output word; consecutive threads write consecutive words of a read.  The 16 bytes of a word are
     * fetched as aligned 32-bit loads (rows start at any byte: stride = max_len+1) and shifted into place. */
    uint64_t total = (uint64_t)n * wstride;
    for (uint64_t g = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; g < total; g += (uint64_t)gridDim.x * blockDim.x) {
        uint64_t r = g / wstride;
        uint32_t w = (uint32_t)(g - r * wstride);
        uint32_t len = lens[r];
        uint32_t b0 = w << 4;
        uint32_t v = 0;
        if (b0 < len) {
            const uint32_t lim = len - b0 < 16 ? len - b0 : 16;
            v = as_pack16((const uint8_t *)ascii + r * stride + b0, lim);
        }
        pack[g] = v;
    }
}

int as_launch_pack(as_gpu *h, const char *d_ascii, size_t stride, const uint32_t *d_lens, size_t n,
                   uint32_t *d_pack, size_t wstride) {
    if (n == 0) return AS_OK;
    uint64_t total = (uint64_t)n * wstride;
    uint64_t blocks = (total + 255) / 256;
    uint64_t cap = (uint64_t)h->sm_count * 32;
    k_pack_reads<<<(unsigned)(blocks < cap ? blocks : cap), 256, 0, h->stream>>>(d_ascii, stride, d_lens, n, d_pack, wstride);
    h->launches++;
    AS_CUDA(cudaGetLastError());
    return AS_OK;
}

/* ---------------------------------------------------------------------------------------------- */
/* batched primitives on ASCII queries                                                             */
/* ---------------------------------------------------------------------------------------------- */
__global__ void __launch_bounds__(128) k_lc_aln_ascii(DevIndex ix, const char *__restrict__ q, int qlen, size_t n,
                                                      int use_table, uint64_t *__restrict__ ko, uint64_t *__restrict__ lo,
                                                      uint64_t *__restrict__ ro) {
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
        const uint8_t *s = (const uint8_t *)q + i * (uint64_t)qlen;
        uint64_t k = 1, l = ix.N - 1;                                   /* alnmain.c:354 */
        if (use_table == 2) {                                           /* fmi_aln from the caller's interval (fmidx.c:303) */
            k = ko[i];
            l = lo[i];
        }
        int left = qlen;
        bool absent = false;
        if (use_table == 1 && qlen >= ix.hlen) {                        /* lchash.c:93-96 */
            left = qlen - ix.hlen;
            uint64_t code = 0;
            for (int t = 0; t < ix.hlen; t++) code |= (uint64_t)as_base_code(s[left + t]) << (2 * t);
            ulonglong2 kl = __ldg(ix.lc + code);
            k = kl.x;
            l = kl.y;
            absent = (k == 0 && l == 0);                                /* lchash.c:101 */
        } else if (use_table == 1) {
            left = qlen - ix.hlen;                                      /* negative: no steps (lchash.c:90,103) */
        }
        uint64_t r = 0;
        if (!absent) {
            for (int t = left - 1; t >= 0; t--) {                       /* fmidx.c:300-310 */
                as_bs_step(ix, as_base_code(s[t]), k, l);
                if (k > l) break;
            }
            r = k > l ? 0 : l - k + 1;
        }
        ko[i] = k;
        lo[i] = l;
        ro[i] = r;
    }
}

__global__ void __launch_bounds__(256) k_sa_access(DevIndex ix, const uint64_t *__restrict__ rows, size_t n,
                                                   uint64_t *__restrict__ pos) {
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
        uint64_t r = rows[i];
        pos[i] = r < ix.N ? as_locate(ix, r) : 0;                           /* fmidx.c:20-23: out of scope -> 0 */
    }
}

static int aln_batch(as_gpu *h, const char *q, int qlen, size_t n, int use_table, uint64_t *k, uint64_t *l, uint64_t *r) {
    if (!h || !q || !k || !l || !r || qlen < 0) return AS_EINVAL;
    if (n == 0) return AS_OK;
    if (use_table == 1 && !h->ix.lc) {
        as_set_error("as_lc_aln_batch: the handle was opened without an lc table");
        return AS_EINVAL;
    }
    AS_CUDA(cudaSetDevice(h->device));
    size_t qb = n * (size_t)qlen;
    AS_TRY(as_scratch(h, &h->misc[0], qb + 16));
    AS_TRY(as_scratch(h, &h->misc[1], n * 3 * sizeof(uint64_t)));
    uint64_t *d_out = (uint64_t *)h->misc[1].p;
    AS_CUDA(cudaMemcpyAsync(h->misc[0].p, q, qb, cudaMemcpyHostToDevice, h->stream));
    if (use_table == 2) {
        AS_CUDA(cudaMemcpyAsync(d_out, k, n * sizeof(uint64_t), cudaMemcpyHostToDevice, h->stream));
        AS_CUDA(cudaMemcpyAsync(d_out + n, l, n * sizeof(uint64_t), cudaMemcpyHostToDevice, h->stream));
    }
    uint64_t blocks = (n + 127) / 128, cap = (uint64_t)h->sm_count * 16;
    k_lc_aln_ascii<<<(unsigned)(blocks < cap ? blocks : cap), 128, 0, h->stream>>>(h->ix, (const char *)h->misc[0].p, qlen, n,
                                                                                 use_table, d_out, d_out + n, d_out + 2 * n);
    h->launches++;
    AS_CUDA(cudaGetLastError());
    AS_CUDA(cudaMemcpyAsync(k, d_out, n * sizeof(uint64_t), cudaMemcpyDeviceToHost, h->stream));
    AS_CUDA(cudaMemcpyAsync(l, d_out + n, n * sizeof(uint64_t), cudaMemcpyDeviceToHost, h->stream));
    AS_CUDA(cudaMemcpyAsync(r, d_out + 2 * n, n * sizeof(uint64_t), cudaMemcpyDeviceToHost, h->stream));
    AS_CUDA(cudaStreamSynchronize(h->stream));
    return AS_OK;
}

extern "C" int as_lc_aln_batch(as_gpu *h, const char *seeds, int qlen, size_t n, uint64_t *k, uint64_t *l, uint64_t *r) {
    return aln_batch(h, seeds, qlen, n, 1, k, l, r);
}

extern "C" int as_fmi_aln_batch(as_gpu *h, const char *qry, int qlen, size_t n, uint64_t *k, uint64_t *l, uint64_t *r) {
    return aln_batch(h, qry, qlen, n, 0, k, l, r);
}

extern "C" int as_fmi_aln_from_batch(as_gpu *h, const char *qry, int qlen, size_t n, uint64_t *k, uint64_t *l, uint64_t *r) {
    return aln_batch(h, qry, qlen, n, 2, k, l, r);
}

/* lc_build (lchash.c:52-73) as the reference does it -- one backward search of hlen steps per k-mer from (1, N-1),
 * (0,0) when the k-mer is absent -- but 4^hlen searches at once, a thread each; needs nothing but the FM index.
 * Entry i of the output is the k-mer whose MSB-first code is i (_seq_from_num, lchash.c:19-33). */
__global__ void __launch_bounds__(256) k_lc_build_bs(DevIndex ix, int hlen, uint64_t code0, uint64_t n, ulonglong2 *__restrict__ out) {
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
        const uint64_t code = code0 + i;
        uint64_t k = 1, l = ix.N - 1;
        for (int t = 0; t < hlen; t++) {                                /* last base of the k-mer first: the low bits */
            as_bs_step(ix, (uint32_t)(code >> (2 * t)) & 3u, k, l);
            if (k > l) break;
        }
        out[i] = k > l ? make_ulonglong2(0, 0) : make_ulonglong2(k, l);
    }
}

extern "C" int as_lc_build(as_gpu *h, int hlen, uint64_t *lc) {
    if (!h || !lc || hlen < 1 || hlen > 14) return AS_EINVAL;
    AS_CUDA(cudaSetDevice(h->device));
    const uint64_t ncodes = 1ULL << (2 * hlen), CH = 1ULL << 22;
    AS_TRY(as_scratch(h, &h->misc[1], CH * sizeof(ulonglong2)));
    for (uint64_t c0 = 0; c0 < ncodes; c0 += CH) {
        const uint64_t n = ncodes - c0 < CH ? ncodes - c0 : CH;
        uint64_t blocks = (n + 255) / 256, cap = (uint64_t)h->sm_count * 16;
        k_lc_build_bs<<<(unsigned)(blocks < cap ? blocks : cap), 256, 0, h->stream>>>(h->ix, hlen, c0, n, (ulonglong2 *)h->misc[1].p);
        h->launches++;
        AS_CUDA(cudaGetLastError());
        AS_CUDA(cudaMemcpyAsync(lc + 2 * c0, h->misc[1].p, n * sizeof(ulonglong2), cudaMemcpyDeviceToHost, h->stream));
        AS_CUDA(cudaStreamSynchronize(h->stream));
    }
    return AS_OK;
}

extern "C" int as_sa_access_batch(as_gpu *h, const uint64_t *rows, size_t n, uint64_t *pos) {
    if (!h || !rows || !pos) return AS_EINVAL;
    if (n == 0) return AS_OK;
    if (!h->ix.sa_lo && !h->ix.csa_lo) {
        as_set_error("as_sa_access_batch: the handle has no suffix array");
        return AS_EINVAL;
    }
    AS_CUDA(cudaSetDevice(h->device));
    AS_TRY(as_scratch(h, &h->misc[0], n * sizeof(uint64_t)));
    AS_TRY(as_scratch(h, &h->misc[1], n * sizeof(uint64_t)));
    AS_CUDA(cudaMemcpyAsync(h->misc[0].p, rows, n * sizeof(uint64_t), cudaMemcpyHostToDevice, h->stream));
    uint64_t blocks = (n + 255) / 256, cap = (uint64_t)h->sm_count * 16;
    k_sa_access<<<(unsigned)(blocks < cap ? blocks : cap), 256, 0, h->stream>>>(h->ix, (const uint64_t *)h->misc[0].p, n,
                                                                              (uint64_t *)h->misc[1].p);
    h->launches++;
    AS_CUDA(cudaGetLastError());
    AS_CUDA(cudaMemcpyAsync(pos, h->misc[1].p, n * sizeof(uint64_t), cudaMemcpyDeviceToHost, h->stream));
    AS_CUDA(cudaStreamSynchronize(h->stream));
    return AS_OK;
}

/* ---------------------------------------------------------------------------------------------- */
/* csa_access (fmidx.c:315-331): locate through the sampled suffix array                           */
/* ---------------------------------------------------------------------------------------------- */
/* _build_csa (fmidx.c:152-163): csa[i] = SA[ratio * i], i < N / ratio + 1 (the last entry lies past the SA when
 * ratio divides N: the reference reads sa_access' out-of-scope 0 there, fmidx.c:20-23). */
__global__ void __launch_bounds__(256) k_csa_sample(DevIndex ix, uint64_t ratio, uint64_t len, uint32_t *__restrict__ lo,
                                                    uint8_t *__restrict__ hi) {
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < len; i += (uint64_t)gridDim.x * blockDim.x) {
        uint64_t row = i * ratio;
        uint64_t v = row < ix.N ? as_sa(ix, row) : 0;
        lo[i] = (uint32_t)v;
        if (hi) hi[i] = (uint8_t)(v >> 32);
    }
}

/* One thread per row.  A step reads ONE FM sector: the row's BWT code comes out of the two planes of the block whose
 * counts give occ(code, row), so the walk is `ratio/2` dependent 32-byte gathers on average plus one sampled-SA
 * gather.  literal: the reference's statement as written (LF - 1, give up after 5*ratio steps). */
__global__ void __launch_bounds__(256) k_csa_access(DevIndex ix, const uint32_t *__restrict__ csa_lo,
                                                    const uint8_t *__restrict__ csa_hi, uint64_t ratio, int literal,
                                                    const uint64_t *__restrict__ rows, size_t n, uint64_t *__restrict__ pos) {
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
        uint64_t loc = rows[i], counter = 0, out = 0;
        bool done = loc >= ix.N;                                        /* out of scope -> 0, as sa_access */
        while (!done && loc % ratio != 0) {
            if (loc == ix.primary) {                                    /* bwt[loc] == '$' (fmidx.c:320-322) */
                out = counter;
                done = true;
                break;
            }
            uint4 cnt;
            uint64_t lo, hi;
            as_ld_block(ix.fm + (loc >> 6), cnt, lo, hi);
            const uint32_t r = (uint32_t)loc & 63u;
            const uint32_t c = (uint32_t)((lo >> r) & 1u) | ((uint32_t)((hi >> r) & 1u) << 1);
            uint64_t m = ((c & 1u) ? lo : ~lo) & ((c & 2u) ? hi : ~hi);
            m &= (r == 63u) ? ~0ULL : ((1ULL << (r + 1)) - 1ULL);
            const uint32_t base = c == 0 ? cnt.x : c == 1 ? cnt.y : c == 2 ? cnt.z : cnt.w;
            uint64_t occ = __ldg(ix.super + 4 * (loc >> AS_SB_SHIFT) + c) + base + (uint64_t)__popcll(m);
            if (c == 0 && (ix.primary >> 6) == (loc >> 6) && ((uint32_t)ix.primary & 63u) <= r) occ -= 1;   /* '$' is stored as A */
            const uint64_t cc = c == 0 ? ix.C[0] : c == 1 ? ix.C[1] : c == 2 ? ix.C[2] : ix.C[3];
            loc = cc + occ - (literal ? 1u : 0u);                       /* fmidx.c:323 */
            counter++;
            if (literal && counter > 5 * ratio) done = true;            /* fmidx.c:325-327: returns 0 */
        }
        if (!done) {
            uint64_t j = loc / ratio;
            uint64_t v = __ldg(csa_lo + j);
            if (csa_hi) v |= (uint64_t)__ldg(csa_hi + j) << 32;
            out = v + counter;                                          /* fmidx.c:330 */
        }
        pos[i] = out;
    }
}

/* SA[csa_ratio * i] from the device SA into the handle (replaces a previous sample) */
int as_csa_sample(as_gpu *h, int csa_ratio) {
    if (h->d_csa_lo) cudaFree(h->d_csa_lo);
    if (h->d_csa_hi) cudaFree(h->d_csa_hi);
    h->d_csa_lo = h->d_csa_hi = nullptr;
    h->csa_ratio = 0;
    h->ix.csa_lo = nullptr;
    h->ix.csa_hi = nullptr;
    h->ix.csa_ratio = 0;
    const uint64_t len = h->ix.N / (uint64_t)csa_ratio + 1;
    AS_CUDA(cudaMalloc(&h->d_csa_lo, len * sizeof(uint32_t)));
    if (h->ix.sa_hi) AS_CUDA(cudaMalloc(&h->d_csa_hi, len));
    uint64_t blocks = (len + 255) / 256, cap = (uint64_t)h->sm_count * 16;
    k_csa_sample<<<(unsigned)(blocks < cap ? blocks : cap), 256, 0, h->stream>>>(h->ix, (uint64_t)csa_ratio, len,
                                                                                (uint32_t *)h->d_csa_lo, (uint8_t *)h->d_csa_hi);
    h->launches++;
    AS_CUDA(cudaGetLastError());
    h->csa_ratio = csa_ratio;
    h->ix.csa_lo = (const uint32_t *)h->d_csa_lo;
    h->ix.csa_hi = (const uint8_t *)h->d_csa_hi;
    h->ix.csa_ratio = (uint32_t)csa_ratio;
    return AS_OK;
}

/* the caller's sampled array (dna_fmi.csa as fmi_read loaded it, fmidx.c:266-271) instead of one sampled here */
extern "C" int as_gpu_attach_csa(as_gpu *h, const uint64_t *csa, uint64_t csa_len, int csa_ratio) {
    if (!h || !csa || csa_ratio <= 0) return AS_EINVAL;
    AS_CUDA(cudaSetDevice(h->device));
    if (csa_len < (h->ix.N + (uint64_t)csa_ratio - 1) / (uint64_t)csa_ratio) {
        as_set_error("as_gpu_attach_csa: %llu samples do not cover %llu rows at ratio %d", (unsigned long long)csa_len,
                     (unsigned long long)h->ix.N, csa_ratio);
        return AS_EINVAL;
    }
    AS_CUDA(cudaStreamSynchronize(h->stream));
    if (h->d_csa_lo) cudaFree(h->d_csa_lo);
    if (h->d_csa_hi) cudaFree(h->d_csa_hi);
    h->d_csa_lo = h->d_csa_hi = nullptr;
    h->csa_ratio = 0;
    uint32_t *lo = (uint32_t *)malloc(csa_len * sizeof(uint32_t));
    uint8_t *hi = (uint8_t *)malloc(csa_len);
    if (!lo || !hi) {
        free(lo);
        free(hi);
        return AS_ENOMEM;
    }
    bool wide = false;
    for (uint64_t i = 0; i < csa_len; i++) {
        lo[i] = (uint32_t)csa[i];
        hi[i] = (uint8_t)(csa[i] >> 32);
        wide |= hi[i] != 0;
    }
    cudaError_t e = cudaMalloc(&h->d_csa_lo, csa_len * sizeof(uint32_t));
    if (e == cudaSuccess) e = cudaMemcpy(h->d_csa_lo, lo, csa_len * sizeof(uint32_t), cudaMemcpyHostToDevice);
    if (e == cudaSuccess && wide) e = cudaMalloc(&h->d_csa_hi, csa_len);
    if (e == cudaSuccess && wide) e = cudaMemcpy(h->d_csa_hi, hi, csa_len, cudaMemcpyHostToDevice);
    free(lo);
    free(hi);
    AS_CUDA(e);
    h->csa_ratio = csa_ratio;
    h->ix.csa_lo = (const uint32_t *)h->d_csa_lo;
    h->ix.csa_hi = (const uint8_t *)h->d_csa_hi;
    h->ix.csa_ratio = (uint32_t)csa_ratio;
    return AS_OK;
}

/* a suffix array for a handle that was opened without one (or an aligner-only handle: then sa_access is all it serves) */
extern "C" int as_gpu_attach_sa(as_gpu *h, const void *sa, uint64_t n, int entry_bytes) {
    if (!h || !sa || n == 0 || (entry_bytes != 4 && entry_bytes != 5 && entry_bytes != 8)) return AS_EINVAL;
    if (h->ix.N != 0 && n > h->ix.N) return AS_EINVAL;
    AS_CUDA(cudaSetDevice(h->device));
    AS_CUDA(cudaStreamSynchronize(h->stream));
    if (h->d_sa_lo) cudaFree(h->d_sa_lo);
    if (h->d_sa_hi) cudaFree(h->d_sa_hi);
    h->d_sa_lo = h->d_sa_hi = nullptr;
    h->ix.sa_lo = nullptr;
    h->ix.sa_hi = nullptr;
    const uint64_t rows = h->ix.N ? h->ix.N : n;
    AS_CUDA(cudaMalloc(&h->d_sa_lo, rows * sizeof(uint32_t)));
    AS_CUDA(cudaMemset(h->d_sa_lo, 0, rows * sizeof(uint32_t)));
    if (entry_bytes > 4 && rows > (1ULL << 32)) {
        AS_CUDA(cudaMalloc(&h->d_sa_hi, rows));
        AS_CUDA(cudaMemset(h->d_sa_hi, 0, rows));
    }
    const uint64_t CH = 1ULL << 24;
    AS_TRY(as_scratch(h, &h->misc[0], CH * 8));
    for (uint64_t r0 = 0; r0 < n; r0 += CH) {
        const uint64_t nr = n - r0 < CH ? n - r0 : CH;
        AS_CUDA(cudaMemcpyAsync(h->misc[0].p, (const uint8_t *)sa + r0 * (uint64_t)entry_bytes, nr * (size_t)entry_bytes,
                                cudaMemcpyHostToDevice, h->stream));
        AS_TRY(as_launch_split_sa(h, (const uint8_t *)h->misc[0].p, entry_bytes, nr, (uint32_t *)h->d_sa_lo + r0,
                                  h->d_sa_hi ? (uint8_t *)h->d_sa_hi + r0 : nullptr));
        AS_CUDA(cudaStreamSynchronize(h->stream));
    }
    h->ix.sa_lo = (const uint32_t *)h->d_sa_lo;
    h->ix.sa_hi = (const uint8_t *)h->d_sa_hi;
    if (h->ix.N == 0) h->ix.N = n;
    return AS_OK;
}

extern "C" int as_csa_access_batch(as_gpu *h, int csa_ratio, int literal, const uint64_t *rows, size_t n, uint64_t *pos) {
    if (!h || !rows || !pos || csa_ratio <= 0) return AS_EINVAL;
    if (n == 0) return AS_OK;
    AS_CUDA(cudaSetDevice(h->device));
    if (h->csa_ratio != csa_ratio) {                                    /* (re)sample: _build_csa */
        if (!h->ix.sa_lo) {
            as_set_error("as_csa_access_batch: the handle has neither a suffix array to sample nor an attached csa of ratio %d", csa_ratio);
            return AS_EINVAL;
        }
        AS_TRY(as_csa_sample(h, csa_ratio));
    }
    AS_TRY(as_scratch(h, &h->misc[0], n * sizeof(uint64_t)));
    AS_TRY(as_scratch(h, &h->misc[1], n * sizeof(uint64_t)));
    AS_CUDA(cudaMemcpyAsync(h->misc[0].p, rows, n * sizeof(uint64_t), cudaMemcpyHostToDevice, h->stream));
    uint64_t blocks = (n + 255) / 256, cap = (uint64_t)h->sm_count * 16;
    k_csa_access<<<(unsigned)(blocks < cap ? blocks : cap), 256, 0, h->stream>>>(
        h->ix, (const uint32_t *)h->d_csa_lo, (const uint8_t *)h->d_csa_hi, (uint64_t)csa_ratio, literal,
        (const uint64_t *)h->misc[0].p, n, (uint64_t *)h->misc[1].p);
    h->launches++;
    AS_CUDA(cudaGetLastError());
    AS_CUDA(cudaMemcpyAsync(pos, h->misc[1].p, n * sizeof(uint64_t), cudaMemcpyDeviceToHost, h->stream));
    AS_CUDA(cudaStreamSynchronize(h->stream));
    return AS_OK;
}
