/*
 * as_build.cu -- index construction on the GPU, straight into the device layouts of as_common.cuh, and the
 * export of those layouts back into the reference's formats.
 *
 * Replaces:
 *   sa_build            psascan/sa_use.cc:8-18 (pSAscan)         -> prefix-doubling suffix sort (radix sort)
 *   _bwt_from_sa5       fmidx.c:76-98                            -> bit planes of the FM blocks
 *   _build_c_table      fmidx.c:101-125, _build_o_dna_table :128-150 -> block counts + prefix scan
 *   _build_csa          fmidx.c:153-163                          -> produced on export (SA[4i])
 *   lc_build            lchash.c:52-73 (4^12 backward searches)  -> one pass over the SA: run boundaries of the
 *                                                                   hlen-mer that starts every suffix
 * The suffix array of a text is unique, so every exported file is byte-identical to what accidx writes
 * (tests/test_index_build.py diffs them against the oracle and the compiled reference).
 * This file's sort takes N < 2^32 - 1 rows (rank pairs are sorted as one 64-bit key): configs 1-3 of BASELINE.json.
 * Larger texts (configs 4-5) go through the partitioned builder in as_build_big.cu.
 */
#include "as_common.cuh"

#include <cub/cub.cuh>
#include <cuda/functional>
#include <new>
#include <stdlib.h>

#define AS_INIT_CHARS 21   /* 3 bits per character: 21 characters in one 63-bit key */

/* ---------------------------------------------------------------------------------------------- */
/* suffix sort                                                                                     */
/* ---------------------------------------------------------------------------------------------- */
/* byte order of the alphabet: '$'(0x24) < 'A' < 'C' < 'G' < 'T'; 0 = past the end (never compared: '$' is unique) */
__device__ __forceinline__ uint32_t code3(uint32_t ch) { return ch == '$' ? 1u : 2u + as_base_code(ch); }

__global__ void __launch_bounds__(256) k_init_keys(const uint8_t *__restrict__ text, uint64_t N,
                                                   unsigned long long *__restrict__ keys, uint32_t *__restrict__ sa) {
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < N; i += (uint64_t)gridDim.x * blockDim.x) {
        unsigned long long k = 0;
#pragma unroll
        for (int t = 0; t < AS_INIT_CHARS; t++) {
            uint32_t c = (i + t < N) ? code3(text[i + t]) : 0u;
            k = (k << 3) | c;
        }
        keys[i] = k;
        sa[i] = (uint32_t)i;
    }
}

/* head[i] = i when sorted position i opens a new group of equal keys, else 0 (max-scan -> group start) */
__global__ void __launch_bounds__(256) k_group_heads(const unsigned long long *__restrict__ keys, uint64_t N,
                                                     uint32_t *__restrict__ head, unsigned long long *n_groups) {
    unsigned long long local = 0;
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < N; i += (uint64_t)gridDim.x * blockDim.x) {
        bool h = (i == 0) || keys[i] != keys[i - 1];
        head[i] = h ? (uint32_t)i : 0u;
        local += h;
    }
#pragma unroll
    for (int d = 16; d >= 1; d >>= 1) local += __shfl_xor_sync(0xFFFFFFFFu, local, d);
    if ((threadIdx.x & 31) == 0 && local) atomicAdd(n_groups, local);
}

__global__ void __launch_bounds__(256) k_scatter_rank(const uint32_t *__restrict__ sa, const uint32_t *__restrict__ grp,
                                                      uint64_t N, uint32_t *__restrict__ rank) {
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < N; i += (uint64_t)gridDim.x * blockDim.x)
        rank[sa[i]] = grp[i];
}

/* key of suffix sa[i] for the doubling round: (rank of its first h characters, rank of the next h) */
__global__ void __launch_bounds__(256) k_pair_keys(const uint32_t *__restrict__ sa, const uint32_t *__restrict__ rank,
                                                   uint64_t N, uint64_t h, unsigned long long *__restrict__ keys) {
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < N; i += (uint64_t)gridDim.x * blockDim.x) {
        const uint64_t p = sa[i];
        const unsigned long long hi = rank[p];
        /* a suffix that ends within h characters contains '$' and is already unique: its low half is never decisive */
        const unsigned long long lo = (p + h < N) ? (unsigned long long)rank[p + h] + 1ULL : 0ULL;
        keys[i] = (hi << 32) | lo;
    }
}

/* ---------------------------------------------------------------------------------------------- */
/* FM blocks, C table                                                                              */
/* ---------------------------------------------------------------------------------------------- */
__global__ void __launch_bounds__(128) k_fm_planes(const uint8_t *__restrict__ text, const uint32_t *__restrict__ sa,
                                                   uint64_t N, FmBlock *__restrict__ fm, unsigned long long *__restrict__ cnt,
                                                   uint64_t nblk, unsigned long long *primary) {
    for (uint64_t b = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; b < nblk; b += (uint64_t)gridDim.x * blockDim.x) {
        const uint64_t r0 = b << 6;
        const uint32_t lim = (uint32_t)((N - r0 < 64) ? N - r0 : 64);
        uint64_t lo = 0, hi = 0;
        uint32_t c4[4] = {0, 0, 0, 0};
        for (uint32_t r = 0; r < lim; r++) {
            const uint32_t p = sa[r0 + r];
            uint32_t code = 0;
            if (p == 0) {
                *primary = r0 + r;                               /* BWT char '$' (fmidx.c:88-93), stored as code 0 */
            } else {
                const uint32_t ch = text[p - 1];
                if (ch != 'A' && ch != 'C' && ch != 'G' && ch != 'T') primary[1] = ch;   /* not a .cat image */
                code = as_base_code(ch);
                c4[code]++;
            }
            lo |= (uint64_t)(code & 1u) << r;
            hi |= (uint64_t)(code >> 1) << r;
        }
        FmBlock blk;
        blk.cnt[0] = blk.cnt[1] = blk.cnt[2] = blk.cnt[3] = 0;
        blk.lo = lo;
        blk.hi = hi;
        fm[b] = blk;
#pragma unroll
        for (int c = 0; c < 4; c++) cnt[(uint64_t)c * nblk + b] = c4[c];
    }
}

/* scan[c*nblk + b] = # of c in rows [0, 64b) -> block counts relative to the super block, super block absolutes */
__global__ void __launch_bounds__(256) k_fm_counts(const unsigned long long *__restrict__ scan, uint64_t nblk,
                                                   FmBlock *__restrict__ fm, uint64_t *__restrict__ super) {
    const uint64_t blk_per_sb = 1ULL << (AS_SB_SHIFT - 6);
    for (uint64_t b = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; b < nblk; b += (uint64_t)gridDim.x * blockDim.x) {
        const uint64_t sb = b / blk_per_sb, b0 = sb * blk_per_sb;
#pragma unroll
        for (int c = 0; c < 4; c++) {
            const unsigned long long base = scan[(uint64_t)c * nblk + b0];
            fm[b].cnt[c] = (uint32_t)(scan[(uint64_t)c * nblk + b] - base);
            if (b == b0) super[4 * sb + c] = base;
        }
    }
}

/* ---------------------------------------------------------------------------------------------- */
/* lc table (lchash.c:52-73): SA interval of every hlen-mer                                        */
/* ---------------------------------------------------------------------------------------------- */
#define AS_NO_CODE 0xFFFFFFFFu

/* LSB-first code (the order the packed reads give, see k_permute_lc) of the hlen-mer starting suffix sa[r].
 * Rows whose suffix reaches '$' within hlen characters hold no hlen-mer; the occurrence that ends exactly on
 * the last base is excluded too: the reference's search starts from k = 1 (lchash.c:57, alnmain.c:354), so
 * occ(c, k-1) counts bwt[0] and that row -- always the first of its run -- is never inside the interval. */
__global__ void __launch_bounds__(256) k_lc_codes(const uint8_t *__restrict__ text, const uint32_t *__restrict__ sa,
                                                  uint64_t N, int hlen, uint32_t *__restrict__ codes) {
    for (uint64_t r = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; r < N; r += (uint64_t)gridDim.x * blockDim.x) {
        const uint64_t p = sa[r];
        uint32_t code = AS_NO_CODE;
        if (r >= 1 && p + (uint64_t)hlen < N - 1) {
            code = 0;
            for (int t = 0; t < hlen; t++) code |= as_base_code(text[p + t]) << (2 * t);
        }
        codes[r] = code;
    }
}

__global__ void __launch_bounds__(256) k_lc_runs(const uint32_t *__restrict__ codes, uint64_t N, ulonglong2 *__restrict__ lc) {
    for (uint64_t r = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; r < N; r += (uint64_t)gridDim.x * blockDim.x) {
        const uint32_t c = codes[r];
        if (c == AS_NO_CODE) continue;
        if (r == 0 || codes[r - 1] != c) lc[c].x = r;
        if (r + 1 == N || codes[r + 1] != c) lc[c].y = r;
    }
}


static inline unsigned grid_for(uint64_t n, int threads, int sm) {
    uint64_t b = (n + threads - 1) / threads, cap = (uint64_t)sm * 16;
    return (unsigned)(b < cap ? (b ? b : 1) : cap);
}

#define CU(call)                                                                                    \
    do {                                                                                            \
        cudaError_t e_ = (call);                                                                    \
        if (e_ != cudaSuccess) {                                                                    \
            as_set_error("%s:%d: %s failed: %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); \
            rc = (e_ == cudaErrorMemoryAllocation) ? AS_ENOMEM : AS_ECUDA;                          \
            goto out;                                                                               \
        }                                                                                           \
    } while (0)

static int build_on_device(as_gpu *h, const char *content, uint64_t N, const as_mta *mta, int mta_len, int hlen) {
    cudaStream_t s = h->stream;
    const int sm = h->sm_count;
    DevIndex &ix = h->ix;
    int rc = AS_OK;
    uint8_t *d_text = nullptr;
    unsigned long long *d_keys[2] = {nullptr, nullptr}, *d_ngroups = nullptr, *d_cnt = nullptr, *d_scan = nullptr;
    uint32_t *d_sa[2] = {nullptr, nullptr}, *d_rank = nullptr, *d_head = nullptr;
    void *d_tmp = nullptr;
    size_t tmp_bytes = 0;
    const uint64_t nblk = (N + 63) >> 6;
    const uint64_t nsuper = ((N - 1) >> AS_SB_SHIFT) + 1;
    unsigned long long totals[4] = {0, 0, 0, 0};
    {
        CU(cudaMalloc(&d_text, N + 64));
        CU(cudaMemsetAsync(d_text + N, 0, 64, s));
        CU(cudaMemcpyAsync(d_text, content, N, cudaMemcpyHostToDevice, s));
        for (int i = 0; i < 2; i++) {
            CU(cudaMalloc(&d_keys[i], (N > 4 * nblk ? N : 4 * nblk) * sizeof(unsigned long long)));
            CU(cudaMalloc(&d_sa[i], N * sizeof(uint32_t)));
        }
        CU(cudaMalloc(&d_rank, N * sizeof(uint32_t)));
        CU(cudaMalloc(&d_head, N * sizeof(uint32_t)));
        CU(cudaMalloc(&d_ngroups, 2 * sizeof(unsigned long long)));
        /* temp storage: the largest of the cub calls below */
        {
            size_t a = 0, b = 0, c = 0;
            cub::DoubleBuffer<unsigned long long> kb(d_keys[0], d_keys[1]);
            cub::DoubleBuffer<uint32_t> vb(d_sa[0], d_sa[1]);
            CU(cub::DeviceRadixSort::SortPairs(nullptr, a, kb, vb, (int64_t)N, 0, 64, s));
            CU(cub::DeviceScan::InclusiveScan(nullptr, b, d_head, d_head, ::cuda::maximum<>{}, (int64_t)N, s));
            CU(cub::DeviceScan::ExclusiveSum(nullptr, c, (unsigned long long *)nullptr, (unsigned long long *)nullptr,
                                             (int64_t)nblk, s));
            tmp_bytes = a > b ? a : b;
            if (c > tmp_bytes) tmp_bytes = c;
            CU(cudaMalloc(&d_tmp, tmp_bytes + 256));
        }
        /* ---- suffix sort: 21-character keys, then prefix doubling while groups of equal rank remain ---- */
        cub::DoubleBuffer<unsigned long long> kb(d_keys[0], d_keys[1]);
        cub::DoubleBuffer<uint32_t> vb(d_sa[0], d_sa[1]);
        k_init_keys<<<grid_for(N, 256, sm), 256, 0, s>>>(d_text, N, kb.Current(), vb.Current());
        h->launches++;
        CU(cudaGetLastError());
        uint64_t hlen_sorted = AS_INIT_CHARS;
        int end_bit = 3 * AS_INIT_CHARS;
        for (int round = 0;; round++) {
            size_t tb = tmp_bytes;
            CU(cub::DeviceRadixSort::SortPairs(d_tmp, tb, kb, vb, (int64_t)N, 0, end_bit, s));
            CU(cudaMemsetAsync(d_ngroups, 0, sizeof(unsigned long long), s));
            k_group_heads<<<grid_for(N, 256, sm), 256, 0, s>>>(kb.Current(), N, d_head, d_ngroups);
            h->launches++;
            CU(cudaGetLastError());
            unsigned long long ng = 0;
            CU(cudaMemcpyAsync(&ng, d_ngroups, sizeof ng, cudaMemcpyDeviceToHost, s));
            CU(cudaStreamSynchronize(s));
            if (ng == N) break;
            if (round > 40) {
                as_set_error("as_gpu_build: suffix sort did not converge (text without a unique final '$'?)");
                rc = AS_EINVAL;
                goto out;
            }
            tb = tmp_bytes;
            CU(cub::DeviceScan::InclusiveScan(d_tmp, tb, d_head, d_head, ::cuda::maximum<>{}, (int64_t)N, s));
            k_scatter_rank<<<grid_for(N, 256, sm), 256, 0, s>>>(vb.Current(), d_head, N, d_rank);
            k_pair_keys<<<grid_for(N, 256, sm), 256, 0, s>>>(vb.Current(), d_rank, N, hlen_sorted, kb.Current());
            h->launches += 2;
            CU(cudaGetLastError());
            hlen_sorted *= 2;
            end_bit = 64;
        }
        uint32_t *d_sa_final = vb.Current();

        /* ---- FM blocks ---- */
        unsigned long long *d_primary = d_ngroups;
        CU(cudaMemsetAsync(d_primary, 0xff, sizeof(unsigned long long), s));
        CU(cudaMemsetAsync(d_primary + 1, 0, sizeof(unsigned long long), s));
        CU(cudaMalloc(&h->d_fm, nblk * sizeof(FmBlock)));
        CU(cudaMalloc(&h->d_super, nsuper * 4 * sizeof(uint64_t)));
        /* the key buffers are free now: 4*nblk u64 counts and their scans fit (nblk = N/64) */
        d_cnt = d_keys[0];
        d_scan = d_keys[1];
        k_fm_planes<<<grid_for(nblk, 128, sm), 128, 0, s>>>(d_text, d_sa_final, N, (FmBlock *)h->d_fm, d_cnt, nblk, d_primary);
        h->launches++;
        CU(cudaGetLastError());
        for (int c = 0; c < 4; c++) {
            size_t tb = tmp_bytes;
            CU(cub::DeviceScan::ExclusiveSum(d_tmp, tb, d_cnt + (uint64_t)c * nblk, d_scan + (uint64_t)c * nblk, (int64_t)nblk, s));
        }
        k_fm_counts<<<grid_for(nblk, 256, sm), 256, 0, s>>>(d_scan, nblk, (FmBlock *)h->d_fm, (uint64_t *)h->d_super);
        h->launches++;
        CU(cudaGetLastError());
        unsigned long long prim[2] = {0, 0}, last_scan[4], last_cnt[4];
        CU(cudaMemcpyAsync(prim, d_primary, sizeof prim, cudaMemcpyDeviceToHost, s));
        for (int c = 0; c < 4; c++) {
            CU(cudaMemcpyAsync(&last_scan[c], d_scan + (uint64_t)c * nblk + (nblk - 1), 8, cudaMemcpyDeviceToHost, s));
            CU(cudaMemcpyAsync(&last_cnt[c], d_cnt + (uint64_t)c * nblk + (nblk - 1), 8, cudaMemcpyDeviceToHost, s));
        }
        CU(cudaStreamSynchronize(s));
        if (prim[0] == ~0ULL) {
            as_set_error("as_gpu_build: no suffix starts at 0?");
            rc = AS_EINVAL;
            goto out;
        }
        if (prim[1] != 0) {
            as_set_error("as_gpu_build: the text holds byte 0x%02x; a .cat image is upper-case ACGT before the final '$' "
                         "(asindex.c:97-110 replaces N and upper-cases)", (unsigned)prim[1]);
            rc = AS_EINVAL;
            goto out;
        }
        for (int c = 0; c < 4; c++) totals[c] = last_scan[c] + last_cnt[c];
        if (totals[0] + totals[1] + totals[2] + totals[3] != N - 1) {
            as_set_error("as_gpu_build: text holds characters other than ACGT before the final '$'");
            rc = AS_EINVAL;
            goto out;
        }
        ix.N = N;
        ix.primary = prim[0];
        ix.hlen = hlen;
        ix.C[0] = 0;                                  /* exclusive prefix counts (fmidx.c:101-125); '$' not counted */
        ix.C[1] = totals[0];
        ix.C[2] = totals[0] + totals[1];
        ix.C[3] = totals[0] + totals[1] + totals[2];
        ix.fm = (const FmBlock *)h->d_fm;
        ix.super = (const uint64_t *)h->d_super;

        /* ---- lc table ---- */
        const uint64_t ncodes = 1ULL << (2 * hlen);
        CU(cudaMalloc(&h->d_lc, ncodes * sizeof(ulonglong2)));
        CU(cudaMemsetAsync(h->d_lc, 0, ncodes * sizeof(ulonglong2), s));      /* absent k-mer: (0,0), lchash.c:64-66 */
        k_lc_codes<<<grid_for(N, 256, sm), 256, 0, s>>>(d_text, d_sa_final, N, hlen, d_rank);
        k_lc_runs<<<grid_for(N, 256, sm), 256, 0, s>>>(d_rank, N, (ulonglong2 *)h->d_lc);
        h->launches += 2;
        CU(cudaGetLastError());
        ix.lc = (const ulonglong2 *)h->d_lc;

        /* ---- SA: keep the sorted buffer ---- */
        CU(cudaMalloc(&h->d_sa_lo, N * sizeof(uint32_t)));
        CU(cudaMemcpyAsync(h->d_sa_lo, d_sa_final, N * sizeof(uint32_t), cudaMemcpyDeviceToDevice, s));
        ix.sa_lo = (const uint32_t *)h->d_sa_lo;
        ix.sa_hi = nullptr;

        /* ---- text, 2 bits per base ---- */
        const uint64_t nwords = ((N + 15) >> 4) + 4;
        CU(cudaMalloc(&h->d_text2, nwords * sizeof(uint32_t)));
        CU(cudaMemsetAsync(h->d_text2, 0, nwords * sizeof(uint32_t), s));
        rc = as_launch_pack_text(h, d_text, N, (uint32_t *)h->d_text2);
        if (rc != AS_OK) goto out;
        ix.text2 = (const uint32_t *)h->d_text2;

        /* ---- mta ---- */
        CU(cudaMalloc(&h->d_mta, (size_t)(mta_len > 0 ? mta_len : 1) * sizeof(as_mta)));
        if (mta_len > 0) CU(cudaMemcpyAsync(h->d_mta, mta, (size_t)mta_len * sizeof(as_mta), cudaMemcpyHostToDevice, s));
        ix.mta = (const as_mta *)h->d_mta;
        ix.mta_len = mta_len;
        CU(cudaStreamSynchronize(s));
    }
out:
    cudaFree(d_text);
    cudaFree(d_keys[0]);
    cudaFree(d_keys[1]);
    cudaFree(d_sa[0]);
    cudaFree(d_sa[1]);
    cudaFree(d_rank);
    cudaFree(d_head);
    cudaFree(d_ngroups);
    cudaFree(d_tmp);
    return rc;
}

extern "C" int as_gpu_build(as_gpu **out, int device, const char *content, uint64_t con_len, const as_mta *mta,
                            int mta_len, int hlen) {
    return as_gpu_build_opts(out, device, content, con_len, mta, mta_len, hlen, nullptr);
}

extern "C" int as_gpu_build_opts(as_gpu **out, int device, const char *content, uint64_t con_len, const as_mta *mta,
                                 int mta_len, int hlen, const as_open_opts *opts) {
    if (!out || !content || mta_len < 0 || (mta_len > 0 && !mta)) return AS_EINVAL;
    *out = nullptr;
    if (con_len < 2 || content[con_len - 1] != '$') {
        as_set_error("as_gpu_build: the text must be a .cat image ending in '$' (asindex.c:108-110)");
        return AS_EINVAL;
    }
    if (hlen < 1 || hlen > 14) {
        as_set_error("as_gpu_build: hlen %d out of range [1,14]", hlen);
        return AS_EINVAL;
    }
    as_gpu *h = as_handle_new(device);
    if (!h) return AS_ENODEV;
    cudaGetLastError();
    if (opts) h->opts = *opts;
    /* texts the single prefix-doubling sort cannot hold go through the partitioned builder (as_build_big.cu);
     * as_open_opts.build_part_max = <suffixes per partition> forces that path on small texts (tests) */
    const uint64_t part_max = h->opts.build_part_max;
    int rc = (con_len >= 0xFFFFFFFFull - 1 || part_max > 0) ? as_build_big(h, content, con_len, mta, mta_len, hlen, part_max)
                                                            : build_on_device(h, content, con_len, mta, mta_len, hlen);
    if (rc == AS_OK) rc = as_build_lc2(h);
    if (rc != AS_OK) {
        as_gpu_close(h);
        return rc;
    }
    *out = h;
    return AS_OK;
}

/* ---------------------------------------------------------------------------------------------- */
/* export to the reference's layouts (Notes.txt:6-25)                                              */
/* ---------------------------------------------------------------------------------------------- */
__global__ void __launch_bounds__(256) k_export_bwt(DevIndex ix, uint64_t row0, uint64_t n, char *__restrict__ out) {
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
        const uint64_t r = row0 + i;
        const FmBlock *b = ix.fm + (r >> 6);
        const uint32_t bit = (uint32_t)r & 63u;
        const uint32_t code = (uint32_t)((b->lo >> bit) & 1u) | ((uint32_t)((b->hi >> bit) & 1u) << 1);
        out[i] = r == ix.primary ? '$' : "ACGT"[code];
    }
}

/* O[4*j + c] = # of c in bwt[0, o_ratio*j)  (fmidx.c:128-150) */
__global__ void __launch_bounds__(256) k_export_o(DevIndex ix, int o_ratio, uint64_t j0, uint64_t n, uint64_t *__restrict__ out) {
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
        const uint64_t row = (j0 + i) * (uint64_t)o_ratio;       /* counts over rows < row */
#pragma unroll
        for (uint32_t c = 0; c < 4; c++) out[4 * i + c] = row == 0 ? 0 : as_occ(ix, c, row - 1);
    }
}

__global__ void __launch_bounds__(256) k_export_lc(DevIndex ix, uint64_t code0, uint64_t n, ulonglong2 *__restrict__ out) {
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
        uint64_t code = code0 + i, rev = 0, c = code;             /* MSB-first file order (lchash.c:36-49) */
        for (int t = 0; t < ix.hlen; t++) {
            rev = (rev << 2) | (c & 3u);
            c >>= 2;
        }
        out[i] = ix.lc[rev];
    }
}

__global__ void __launch_bounds__(256) k_export_sa(DevIndex ix, uint64_t row0, uint64_t n, uint64_t stride, int entry_bytes,
                                                   uint8_t *__restrict__ out) {
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
        const uint64_t v = as_locate(ix, row0 + i * stride);
        uint8_t *p = out + i * (uint64_t)entry_bytes;
        for (int b = 0; b < entry_bytes; b++) p[b] = (uint8_t)(v >> (8 * b));
    }
}

extern "C" int as_gpu_export(as_gpu *h, as_index_out *o) {
    if (!h || !o) return AS_EINVAL;
    AS_CUDA(cudaSetDevice(h->device));
    cudaStream_t s = h->stream;
    const DevIndex &ix = h->ix;
    const uint64_t N = ix.N;
    const size_t CH = (size_t)1 << 24;                     /* elements per staged chunk */
    int rc = AS_OK;
    void *d_stage = nullptr;
    {
        CU(cudaMalloc(&d_stage, CH * 32));
        if (o->C) {                                        /* exclusive prefix counts by ASCII code */
            for (int i = 0; i < 256; i++)
                o->C[i] = i <= 'A' ? 0 : i <= 'C' ? ix.C[1] : i <= 'G' ? ix.C[2] : i <= 'T' ? ix.C[3] : N - 1;
        }
        if (o->bwt) {
            for (uint64_t r0 = 0; r0 < N; r0 += CH) {
                const uint64_t n = N - r0 < CH ? N - r0 : CH;
                k_export_bwt<<<grid_for(n, 256, h->sm_count), 256, 0, s>>>(ix, r0, n, (char *)d_stage);
                h->launches++;
                CU(cudaGetLastError());
                CU(cudaMemcpyAsync(o->bwt + r0, d_stage, n, cudaMemcpyDeviceToHost, s));
                CU(cudaStreamSynchronize(s));
            }
        }
        if (o->O) {
            if (o->o_ratio <= 0) {
                as_set_error("as_gpu_export: o_ratio %d must be positive", o->o_ratio);
                rc = AS_EINVAL;
                goto out;
            }
            /* the reference never writes the block at row N when N %% o_ratio == 0 (heap garbage in its file);
             * it is written here, with the true counts */
            const uint64_t nj = N / (uint64_t)o->o_ratio + 1;
            for (uint64_t j0 = 0; j0 < nj; j0 += CH) {
                const uint64_t n = nj - j0 < CH ? nj - j0 : CH;
                k_export_o<<<grid_for(n, 256, h->sm_count), 256, 0, s>>>(ix, o->o_ratio, j0, n, (uint64_t *)d_stage);
                h->launches++;
                CU(cudaGetLastError());
                CU(cudaMemcpyAsync(o->O + 4 * j0, d_stage, n * 32, cudaMemcpyDeviceToHost, s));
                CU(cudaStreamSynchronize(s));
            }
        }
        if (o->lc) {
            const uint64_t ncodes = 1ULL << (2 * ix.hlen);
            for (uint64_t c0 = 0; c0 < ncodes; c0 += CH) {
                const uint64_t n = ncodes - c0 < CH ? ncodes - c0 : CH;
                k_export_lc<<<grid_for(n, 256, h->sm_count), 256, 0, s>>>(ix, c0, n, (ulonglong2 *)d_stage);
                h->launches++;
                CU(cudaGetLastError());
                CU(cudaMemcpyAsync(o->lc + 2 * c0, d_stage, n * 16, cudaMemcpyDeviceToHost, s));
                CU(cudaStreamSynchronize(s));
            }
        }
        if (o->sa) {
            if (o->sa_entry_bytes != 5 && o->sa_entry_bytes != 8) {
                as_set_error("as_gpu_export: sa_entry_bytes must be 5 (.sa5) or 8");
                rc = AS_EINVAL;
                goto out;
            }
            for (uint64_t r0 = 0; r0 < N; r0 += CH) {
                const uint64_t n = N - r0 < CH ? N - r0 : CH;
                k_export_sa<<<grid_for(n, 256, h->sm_count), 256, 0, s>>>(ix, r0, n, 1, o->sa_entry_bytes, (uint8_t *)d_stage);
                h->launches++;
                CU(cudaGetLastError());
                CU(cudaMemcpyAsync((uint8_t *)o->sa + r0 * (uint64_t)o->sa_entry_bytes, d_stage, n * (size_t)o->sa_entry_bytes,
                                   cudaMemcpyDeviceToHost, s));
                CU(cudaStreamSynchronize(s));
            }
        }
        if (o->csa) {                                       /* fmidx.c:153-163: SA[csa_ratio * i] */
            if (o->csa_ratio <= 0) { rc = AS_EINVAL; goto out; }
            const uint64_t nc = (N + (uint64_t)o->csa_ratio - 1) / (uint64_t)o->csa_ratio;
            for (uint64_t i0 = 0; i0 < nc; i0 += CH) {
                const uint64_t n = nc - i0 < CH ? nc - i0 : CH;
                k_export_sa<<<grid_for(n, 256, h->sm_count), 256, 0, s>>>(ix, i0 * (uint64_t)o->csa_ratio, n,
                                                                        (uint64_t)o->csa_ratio, 8, (uint8_t *)d_stage);
                h->launches++;
                CU(cudaGetLastError());
                CU(cudaMemcpyAsync(o->csa + i0, d_stage, n * 8, cudaMemcpyDeviceToHost, s));
                CU(cudaStreamSynchronize(s));
            }
        }
    }
out:
    cudaFree(d_stage);
    return rc;
}
