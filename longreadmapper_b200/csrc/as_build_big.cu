/*
 * as_build_big.cu -- index construction for texts of 2^32-1 rows and more (BASELINE.json configs 4-5: a 3.1 Gbp
 * genome gives a 6.2 G-row .cat), where the rank pairs of the prefix-doubling sort in as_build.cu no longer fit
 * one 64-bit key and the whole problem no longer fits the radix sort's double buffers.
 *
 * Replaces sa_build (psascan/sa_use.cc:8-18: pSAscan, external memory) for that size.  Method: the suffixes are
 * partitioned by their first AS_BIN_CHARS characters (a histogram over the text, consecutive bins grouped into
 * partitions of at most part_max suffixes); each partition is collected, radix-sorted by the next 21 characters as
 * one 63-bit key, and groups that still tie are refined 21 characters at a time (sort by the next key, stable sort
 * by group) until every group is a single suffix -- '$' is unique, so that always ends.  Partitions are in bin order,
 * so their concatenation is the suffix array, written straight into the split 40-bit device layout (sa_lo + sa_hi).
 * The FM blocks, C table, super blocks and the lc table follow from the SA exactly as in as_build.cu, with 40-bit
 * rows.  A suffix array is unique, so the result is the index accidx would write (asindex.c:129-153), and
 * tests/test_index_build.py checks the partitioned path against the single-sort path on small texts.
 */
#include "as_common.cuh"

#include <cub/cub.cuh>
#include <thrust/iterator/counting_iterator.h>
#include <cuda/functional>
#include <stdlib.h>
#include <vector>

#define AS_KEY_CHARS 21    /* 3 bits per character: 21 characters in one 63-bit key */
#define AS_BIN_CHARS 4     /* partition bins: first 4 characters, 12 bits */
#define AS_NBINS (1 << (3 * AS_BIN_CHARS))

__device__ __forceinline__ uint32_t big_code3(uint32_t ch) { return ch == '$' ? 1u : 2u + as_base_code(ch); }

/* 21-character key of the suffix at p; characters past the end read as 0 (below '$') */
__device__ __forceinline__ unsigned long long big_key(const uint8_t *__restrict__ text, uint64_t N, uint64_t p) {
    unsigned long long k = 0;
#pragma unroll
    for (int t = 0; t < AS_KEY_CHARS; t++) {
        const uint32_t c = (p + t < N) ? big_code3(text[p + t]) : 0u;
        k = (k << 3) | c;
    }
    return k;
}
__device__ __forceinline__ uint32_t big_bin(const uint8_t *__restrict__ text, uint64_t N, uint64_t p) {
    uint32_t k = 0;
#pragma unroll
    for (int t = 0; t < AS_BIN_CHARS; t++) {
        const uint32_t c = (p + t < N) ? big_code3(text[p + t]) : 0u;
        k = (k << 3) | c;
    }
    return k;
}

__global__ void __launch_bounds__(256) k_big_hist(const uint8_t *__restrict__ text, uint64_t N, unsigned long long *__restrict__ hist) {
    __shared__ unsigned int sh[AS_NBINS];
    for (int i = threadIdx.x; i < AS_NBINS; i += blockDim.x) sh[i] = 0;
    __syncthreads();
    /* blocks take contiguous slabs so that a block's shared counters stay below 2^32 */
    const uint64_t per = (N + gridDim.x - 1) / gridDim.x, lo = per * blockIdx.x, hi = lo + per < N ? lo + per : N;
    for (uint64_t i = lo + threadIdx.x; i < hi; i += blockDim.x) atomicAdd(&sh[big_bin(text, N, i)], 1u);
    __syncthreads();
    for (int i = threadIdx.x; i < AS_NBINS; i += blockDim.x)
        if (sh[i]) atomicAdd(&hist[i], (unsigned long long)sh[i]);
}

/* suffixes whose bin is in [b0, b1): position and first key, in any order (they are sorted next) */
__global__ void __launch_bounds__(256) k_big_collect(const uint8_t *__restrict__ text, uint64_t N, uint32_t b0, uint32_t b1,
                                                     unsigned long long *__restrict__ keys, unsigned long long *__restrict__ pos,
                                                     unsigned long long *__restrict__ counter) {
    const uint32_t lane = threadIdx.x & 31u;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    const uint64_t rounds = (N + stride - 1) / stride;
    uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
    for (uint64_t r = 0; r < rounds; r++, i += stride) {
        bool take = false;
        if (i < N) {
            const uint32_t b = big_bin(text, N, i);
            take = b >= b0 && b < b1;
        }
        const uint32_t m = __ballot_sync(0xFFFFFFFFu, take);
        if (m) {
            unsigned long long base = 0;
            if (lane == (uint32_t)(__ffs((int)m) - 1)) base = atomicAdd(counter, (unsigned long long)__popc(m));
            base = __shfl_sync(0xFFFFFFFFu, base, __ffs((int)m) - 1);
            if (take) {
                const unsigned long long o = base + (unsigned long long)__popc(m & ((1u << lane) - 1u));
                keys[o] = big_key(text, N, i);
                pos[o] = i;
            }
        }
    }
}

/* tie[i] = 1 when the sorted key at i equals a neighbour's */
__global__ void __launch_bounds__(256) k_big_tie_flags(const unsigned long long *__restrict__ keys, uint64_t n, uint8_t *__restrict__ tie) {
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x)
        tie[i] = ((i > 0 && keys[i] == keys[i - 1]) || (i + 1 < n && keys[i] == keys[i + 1])) ? 1 : 0;
}
/* tied suffixes: position, slot in the partition, head flag of their run (the slots arrive in sorted order) */
__global__ void __launch_bounds__(256) k_big_tie_init(const unsigned long long *__restrict__ keys, const unsigned long long *__restrict__ pos,
                                                      const uint32_t *__restrict__ slot, uint64_t nt,
                                                      unsigned long long *__restrict__ tpos, uint32_t *__restrict__ head) {
    for (uint64_t j = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; j < nt; j += (uint64_t)gridDim.x * blockDim.x) {
        const uint32_t s = slot[j];
        tpos[j] = pos[s];
        head[j] = (j == 0 || keys[s] != keys[slot[j - 1]]) ? (uint32_t)j : 0u;
    }
}
__global__ void __launch_bounds__(256) k_big_round_keys(const uint8_t *__restrict__ text, uint64_t N,
                                                        const unsigned long long *__restrict__ tpos, uint64_t off, uint64_t nt,
                                                        unsigned long long *__restrict__ keys, uint32_t *__restrict__ idx) {
    for (uint64_t j = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; j < nt; j += (uint64_t)gridDim.x * blockDim.x) {
        keys[j] = big_key(text, N, tpos[j] + off);
        idx[j] = (uint32_t)j;
    }
}
__global__ void __launch_bounds__(256) k_big_gather_grp(const uint32_t *__restrict__ grp, const uint32_t *__restrict__ idx, uint64_t nt,
                                                        uint32_t *__restrict__ out) {
    for (uint64_t j = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; j < nt; j += (uint64_t)gridDim.x * blockDim.x) out[j] = grp[idx[j]];
}
/* after the (group, key) sort: new order of the positions, sub-group heads */
__global__ void __launch_bounds__(256) k_big_regroup(const uint8_t *__restrict__ text, uint64_t N,
                                                     const unsigned long long *__restrict__ tpos, const uint32_t *__restrict__ idx,
                                                     const uint32_t *__restrict__ grp, uint64_t off, uint64_t nt,
                                                     unsigned long long *__restrict__ npos, uint32_t *__restrict__ head) {
    for (uint64_t j = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; j < nt; j += (uint64_t)gridDim.x * blockDim.x) {
        const unsigned long long p = tpos[idx[j]];
        npos[j] = p;
        bool h = j == 0 || grp[j] != grp[j - 1];
        if (!h) h = big_key(text, N, p + off) != big_key(text, N, tpos[idx[j - 1]] + off);
        head[j] = h ? (uint32_t)j : 0u;
    }
}
/* write the refined order back; flag what still ties (head[] holds the max-scanned group start here) */
__global__ void __launch_bounds__(256) k_big_writeback(const unsigned long long *__restrict__ npos, const uint32_t *__restrict__ slot,
                                                       const uint32_t *__restrict__ grp, uint64_t nt, unsigned long long *__restrict__ pos,
                                                       uint8_t *__restrict__ tie) {
    for (uint64_t j = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; j < nt; j += (uint64_t)gridDim.x * blockDim.x) {
        pos[slot[j]] = npos[j];
        const bool single = (j == 0 || grp[j] != grp[j - 1]) && (j + 1 == nt || grp[j + 1] != grp[j]);
        tie[j] = single ? 0 : 1;
    }
}
__global__ void __launch_bounds__(256) k_big_compact(const uint32_t *__restrict__ sel, uint64_t n2, const unsigned long long *__restrict__ npos,
                                                     const uint32_t *__restrict__ slot, const uint32_t *__restrict__ grp,
                                                     unsigned long long *__restrict__ tpos2, uint32_t *__restrict__ slot2, uint32_t *__restrict__ grp2) {
    for (uint64_t j = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; j < n2; j += (uint64_t)gridDim.x * blockDim.x) {
        const uint32_t s = sel[j];
        tpos2[j] = npos[s];
        slot2[j] = slot[s];
        grp2[j] = grp[s];
    }
}
__global__ void __launch_bounds__(256) k_big_store_sa(const unsigned long long *__restrict__ pos, uint64_t n, uint64_t row0,
                                                      uint32_t *__restrict__ sa_lo, uint8_t *__restrict__ sa_hi) {
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
        const unsigned long long p = pos[i];
        sa_lo[row0 + i] = (uint32_t)p;
        if (sa_hi) sa_hi[row0 + i] = (uint8_t)(p >> 32);
    }
}

/* ---------------------------------------------------------------------------------------------- */
/* FM blocks and lc table from a 40-bit SA (same results as k_fm_planes / k_lc_codes in as_build.cu) */
/* ---------------------------------------------------------------------------------------------- */
__device__ __forceinline__ uint64_t sa40(const uint32_t *__restrict__ lo, const uint8_t *__restrict__ hi, uint64_t r) {
    uint64_t v = lo[r];
    if (hi) v |= (uint64_t)hi[r] << 32;
    return v;
}

__global__ void __launch_bounds__(128) k_big_fm_planes(const uint8_t *__restrict__ text, const uint32_t *__restrict__ sa_lo,
                                                       const uint8_t *__restrict__ sa_hi, uint64_t N, FmBlock *__restrict__ fm,
                                                       unsigned long long *__restrict__ cnt, uint64_t nblk,
                                                       unsigned long long *primary) {
    for (uint64_t b = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; b < nblk; b += (uint64_t)gridDim.x * blockDim.x) {
        const uint64_t r0 = b << 6;
        const uint32_t lim = (uint32_t)((N - r0 < 64) ? N - r0 : 64);
        uint64_t lo = 0, hi = 0;
        uint32_t c4[4] = {0, 0, 0, 0};
        for (uint32_t r = 0; r < lim; r++) {
            const uint64_t p = sa40(sa_lo, sa_hi, r0 + r);
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

__global__ void __launch_bounds__(256) k_big_fm_counts(const unsigned long long *__restrict__ scan, uint64_t nblk,
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

#define AS_BIG_NO_CODE 0xFFFFFFFFu
/* LSB-first code of the hlen-mer that starts suffix SA[r]; see k_lc_codes in as_build.cu for the two exclusions */
__device__ __forceinline__ uint32_t big_lc_code(const uint8_t *__restrict__ text, const uint32_t *__restrict__ sa_lo,
                                                const uint8_t *__restrict__ sa_hi, uint64_t N, int hlen, uint64_t r) {
    const uint64_t p = sa40(sa_lo, sa_hi, r);
    if (r < 1 || p + (uint64_t)hlen >= N - 1) return AS_BIG_NO_CODE;
    uint32_t code = 0;
    for (int t = 0; t < hlen; t++) code |= as_base_code(text[p + t]) << (2 * t);
    return code;
}
/* run boundaries of equal codes down the SA = the SA interval of every hlen-mer (lchash.c:52-73) */
__global__ void __launch_bounds__(256) k_big_lc(const uint8_t *__restrict__ text, const uint32_t *__restrict__ sa_lo,
                                                const uint8_t *__restrict__ sa_hi, uint64_t N, int hlen, ulonglong2 *__restrict__ lc) {
    /* a thread walks a short run of rows so that each code is computed once, plus the two neighbours of the run */
    const uint64_t RUN = 16;
    const uint64_t nruns = (N + RUN - 1) / RUN;
    for (uint64_t q = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; q < nruns; q += (uint64_t)gridDim.x * blockDim.x) {
        const uint64_t r0 = q * RUN, r1 = r0 + RUN < N ? r0 + RUN : N;
        uint32_t prev = r0 > 0 ? big_lc_code(text, sa_lo, sa_hi, N, hlen, r0 - 1) : AS_BIG_NO_CODE;
        uint32_t cur = big_lc_code(text, sa_lo, sa_hi, N, hlen, r0);
        for (uint64_t r = r0; r < r1; r++) {
            const uint32_t next = r + 1 < N ? big_lc_code(text, sa_lo, sa_hi, N, hlen, r + 1) : AS_BIG_NO_CODE;
            if (cur != AS_BIG_NO_CODE) {
                if (r == 0 || prev != cur) lc[cur].x = r;
                if (r + 1 == N || next != cur) lc[cur].y = r;
            }
            prev = cur;
            cur = next;
        }
    }
}

/* ---------------------------------------------------------------------------------------------- */
/* host driver                                                                                      */
/* ---------------------------------------------------------------------------------------------- */
static inline unsigned big_grid(uint64_t n, int threads, int sm) {
    uint64_t b = (n + threads - 1) / threads, cap = (uint64_t)sm * 16;
    return (unsigned)(b < cap ? (b ? b : 1) : cap);
}

#define BCU(call)                                                                                   \
    do {                                                                                            \
        cudaError_t e_ = (call);                                                                    \
        if (e_ != cudaSuccess) {                                                                    \
            as_set_error("%s:%d: %s failed: %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); \
            rc = (e_ == cudaErrorMemoryAllocation) ? AS_ENOMEM : AS_ECUDA;                          \
            goto out;                                                                               \
        }                                                                                           \
    } while (0)

struct BigTmp {                     /* cub temporary storage, grown on demand */
    void *p = nullptr;
    size_t cap = 0;
    cudaError_t need(size_t bytes) {
        if (bytes <= cap) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
        cudaError_t e = cudaMalloc(&p, bytes + 256);
        if (e == cudaSuccess) cap = bytes;
        return e;
    }
};

/* refine the groups of equal 21-character keys of one sorted partition, 21 characters per round */
static int big_refine(as_gpu *h, const uint8_t *d_text, uint64_t N, unsigned long long *d_keys, unsigned long long *d_pos,
                      uint64_t n, BigTmp &tmp, uint64_t *rounds_out) {
    cudaStream_t s = h->stream;
    const int sm = h->sm_count;
    int rc = AS_OK;
    uint8_t *d_tie = nullptr;
    uint32_t *d_slot = nullptr, *d_grp = nullptr, *d_idx[2] = {nullptr, nullptr}, *d_g1[2] = {nullptr, nullptr}, *d_head = nullptr,
             *d_sel = nullptr, *d_slot2 = nullptr, *d_grp2 = nullptr;
    unsigned long long *d_tpos = nullptr, *d_npos = nullptr, *d_k[2] = {nullptr, nullptr}, *d_cnt = nullptr;
    unsigned long long nt = 0;
    size_t tb = 0;
    uint64_t round = 0;
    {
        BCU(cudaMalloc(&d_tie, n));
        BCU(cudaMalloc(&d_cnt, sizeof(unsigned long long)));
        k_big_tie_flags<<<big_grid(n, 256, sm), 256, 0, s>>>(d_keys, n, d_tie);
        h->launches++;
        BCU(cudaGetLastError());
        /* slots of the tied suffixes, in sorted order */
        BCU(cudaMalloc(&d_slot, (n ? n : 1) * sizeof(uint32_t)));
        {
            thrust::counting_iterator<uint32_t> it(0);
            BCU(cub::DeviceSelect::Flagged(nullptr, tb, it, d_tie, d_slot, d_cnt, (int64_t)n, s));
            BCU(tmp.need(tb));
            tb = tmp.cap;
            BCU(cub::DeviceSelect::Flagged(tmp.p, tb, it, d_tie, d_slot, d_cnt, (int64_t)n, s));
        }
        BCU(cudaMemcpyAsync(&nt, d_cnt, sizeof nt, cudaMemcpyDeviceToHost, s));
        BCU(cudaStreamSynchronize(s));
        if (nt == 0) goto out;
        BCU(cudaMalloc(&d_tpos, nt * 8));
        BCU(cudaMalloc(&d_npos, nt * 8));
        BCU(cudaMalloc(&d_k[0], nt * 8));
        BCU(cudaMalloc(&d_k[1], nt * 8));
        BCU(cudaMalloc(&d_grp, nt * 4));
        BCU(cudaMalloc(&d_head, nt * 4));
        BCU(cudaMalloc(&d_sel, nt * 4));
        BCU(cudaMalloc(&d_slot2, nt * 4));
        BCU(cudaMalloc(&d_grp2, nt * 4));
        for (int i = 0; i < 2; i++) {
            BCU(cudaMalloc(&d_idx[i], nt * 4));
            BCU(cudaMalloc(&d_g1[i], nt * 4));
        }
        k_big_tie_init<<<big_grid(nt, 256, sm), 256, 0, s>>>(d_keys, d_pos, d_slot, nt, d_tpos, d_head);
        h->launches++;
        BCU(cudaGetLastError());
        BCU(cub::DeviceScan::InclusiveScan(nullptr, tb, d_head, d_grp, ::cuda::maximum<>{}, (int64_t)nt, s));
        BCU(tmp.need(tb));
        tb = tmp.cap;
        BCU(cub::DeviceScan::InclusiveScan(tmp.p, tb, d_head, d_grp, ::cuda::maximum<>{}, (int64_t)nt, s));
        while (nt > 0) {
            round++;
            if (round > 200000) {
                as_set_error("as_gpu_build: suffix groups did not separate (text without a unique final '$'?)");
                rc = AS_EINVAL;
                goto out;
            }
            const uint64_t off = (uint64_t)AS_KEY_CHARS * round;
            /* order inside every group by the next 21 characters: sort by key, then stable sort by group */
            k_big_round_keys<<<big_grid(nt, 256, sm), 256, 0, s>>>(d_text, N, d_tpos, off, nt, d_k[0], d_idx[0]);
            h->launches++;
            BCU(cudaGetLastError());
            cub::DoubleBuffer<unsigned long long> kb(d_k[0], d_k[1]);
            cub::DoubleBuffer<uint32_t> ib(d_idx[0], d_idx[1]);
            BCU(cub::DeviceRadixSort::SortPairs(nullptr, tb, kb, ib, (int64_t)nt, 0, 64, s));
            BCU(tmp.need(tb));
            tb = tmp.cap;
            BCU(cub::DeviceRadixSort::SortPairs(tmp.p, tb, kb, ib, (int64_t)nt, 0, 64, s));
            uint32_t *idx_sorted = ib.Current(), *idx_other = ib.Alternate();
            k_big_gather_grp<<<big_grid(nt, 256, sm), 256, 0, s>>>(d_grp, idx_sorted, nt, d_g1[0]);
            h->launches++;
            BCU(cudaGetLastError());
            cub::DoubleBuffer<uint32_t> gb(d_g1[0], d_g1[1]);
            cub::DoubleBuffer<uint32_t> ib2(idx_sorted, idx_other);
            BCU(cub::DeviceRadixSort::SortPairs(nullptr, tb, gb, ib2, (int64_t)nt, 0, 32, s));
            BCU(tmp.need(tb));
            tb = tmp.cap;
            BCU(cub::DeviceRadixSort::SortPairs(tmp.p, tb, gb, ib2, (int64_t)nt, 0, 32, s));
            /* groups keep their slots, so the sorted group column equals d_grp */
            k_big_regroup<<<big_grid(nt, 256, sm), 256, 0, s>>>(d_text, N, d_tpos, ib2.Current(), d_grp, off, nt, d_npos, d_head);
            h->launches++;
            BCU(cudaGetLastError());
            BCU(cub::DeviceScan::InclusiveScan(nullptr, tb, d_head, d_grp2, ::cuda::maximum<>{}, (int64_t)nt, s));
            BCU(tmp.need(tb));
            tb = tmp.cap;
            BCU(cub::DeviceScan::InclusiveScan(tmp.p, tb, d_head, d_grp2, ::cuda::maximum<>{}, (int64_t)nt, s));
            k_big_writeback<<<big_grid(nt, 256, sm), 256, 0, s>>>(d_npos, d_slot, d_grp2, nt, d_pos, d_tie);
            h->launches++;
            BCU(cudaGetLastError());
            {
                thrust::counting_iterator<uint32_t> it(0);
                BCU(cub::DeviceSelect::Flagged(nullptr, tb, it, d_tie, d_sel, d_cnt, (int64_t)nt, s));
                BCU(tmp.need(tb));
                tb = tmp.cap;
                BCU(cub::DeviceSelect::Flagged(tmp.p, tb, it, d_tie, d_sel, d_cnt, (int64_t)nt, s));
            }
            unsigned long long n2 = 0;
            BCU(cudaMemcpyAsync(&n2, d_cnt, sizeof n2, cudaMemcpyDeviceToHost, s));
            BCU(cudaStreamSynchronize(s));
            if (n2 > 0) {
                k_big_compact<<<big_grid(n2, 256, sm), 256, 0, s>>>(d_sel, n2, d_npos, d_slot, d_grp2, d_tpos, d_slot2, d_head);
                h->launches++;
                BCU(cudaGetLastError());
                /* d_head now holds the compacted group ids; swap the roles for the next round */
                uint32_t *t = d_slot; d_slot = d_slot2; d_slot2 = t;
                t = d_grp; d_grp = d_head; d_head = t;
            }
            nt = n2;
        }
    }
out:
    if (rounds_out) *rounds_out = round;
    cudaFree(d_tie); cudaFree(d_slot); cudaFree(d_grp); cudaFree(d_idx[0]); cudaFree(d_idx[1]); cudaFree(d_g1[0]); cudaFree(d_g1[1]);
    cudaFree(d_head); cudaFree(d_sel); cudaFree(d_slot2); cudaFree(d_grp2); cudaFree(d_tpos); cudaFree(d_npos); cudaFree(d_k[0]);
    cudaFree(d_k[1]); cudaFree(d_cnt);
    return rc;
}

/*
 * part_max: most suffixes sorted at once (0 = default 2^29).  Tests pass small values to run the partitioned path on
 * small texts; sa_hi is allocated when N > 2^32 (or when wide_sa forces it).
 */
int as_build_big(as_gpu *h, const char *content, uint64_t N, const as_mta *mta, int mta_len, int hlen, uint64_t part_max) {
    cudaStream_t s = h->stream;
    const int sm = h->sm_count;
    DevIndex &ix = h->ix;
    int rc = AS_OK;
    uint8_t *d_text = nullptr;
    unsigned long long *d_hist = nullptr, *d_keys[2] = {nullptr, nullptr}, *d_pos[2] = {nullptr, nullptr}, *d_counter = nullptr,
                       *d_cnt = nullptr, *d_scan = nullptr;
    BigTmp tmp;
    const uint64_t nblk = (N + 63) >> 6;
    const uint64_t nsuper = ((N - 1) >> AS_SB_SHIFT) + 1;
    unsigned long long totals[4] = {0, 0, 0, 0};
    std::vector<unsigned long long> hist(AS_NBINS);
    std::vector<std::pair<uint32_t, uint32_t>> parts;      /* [b0, b1) */
    std::vector<uint64_t> part_n;
    uint64_t pmax = 0, total_rounds = 0;
    if (part_max == 0) part_max = 1ULL << 29;
    {
        BCU(cudaMalloc(&d_text, N + 64));
        BCU(cudaMemsetAsync(d_text + N, 0, 64, s));
        /* the text arrives from pageable host memory; cudaMemcpy stages it */
        BCU(cudaMemcpy(d_text, content, N, cudaMemcpyHostToDevice));
        BCU(cudaMalloc(&d_hist, AS_NBINS * sizeof(unsigned long long)));
        BCU(cudaMemsetAsync(d_hist, 0, AS_NBINS * sizeof(unsigned long long), s));
        BCU(cudaMalloc(&d_counter, 2 * sizeof(unsigned long long)));
        k_big_hist<<<(unsigned)(sm * 8), 256, 0, s>>>(d_text, N, d_hist);
        h->launches++;
        BCU(cudaGetLastError());
        BCU(cudaMemcpyAsync(hist.data(), d_hist, AS_NBINS * sizeof(unsigned long long), cudaMemcpyDeviceToHost, s));
        BCU(cudaStreamSynchronize(s));
        /* consecutive bins -> partitions of at most part_max suffixes (a single larger bin stands alone) */
        {
            uint32_t b0 = 0;
            uint64_t acc = 0;
            for (uint32_t b = 0; b < AS_NBINS; b++) {
                if (acc > 0 && acc + hist[b] > part_max) {
                    parts.push_back({b0, b});
                    part_n.push_back(acc);
                    b0 = b;
                    acc = 0;
                }
                acc += hist[b];
            }
            if (acc > 0) {
                parts.push_back({b0, (uint32_t)AS_NBINS});
                part_n.push_back(acc);
            }
            for (uint64_t v : part_n) pmax = v > pmax ? v : pmax;
            if (pmax >= (1ULL << 31)) {
                as_set_error("as_gpu_build: %llu suffixes share their first %d characters; the partitioned sort takes < 2^31 per partition",
                             (unsigned long long)pmax, AS_BIN_CHARS);
                rc = AS_ERANGE;
                goto out;
            }
        }
        for (int i = 0; i < 2; i++) {
            BCU(cudaMalloc(&d_keys[i], pmax * 8));
            BCU(cudaMalloc(&d_pos[i], pmax * 8));
        }
        BCU(cudaMalloc(&h->d_sa_lo, N * sizeof(uint32_t)));
        if (N > (1ULL << 32)) BCU(cudaMalloc(&h->d_sa_hi, N));
        /* ---- suffix sort, one partition at a time ---- */
        {
            uint64_t row0 = 0;
            for (size_t pi = 0; pi < parts.size(); pi++) {
                const uint64_t n = part_n[pi];
                BCU(cudaMemsetAsync(d_counter, 0, sizeof(unsigned long long), s));
                k_big_collect<<<big_grid(N, 256, sm), 256, 0, s>>>(d_text, N, parts[pi].first, parts[pi].second, d_keys[0], d_pos[0], d_counter);
                h->launches++;
                BCU(cudaGetLastError());
                cub::DoubleBuffer<unsigned long long> kb(d_keys[0], d_keys[1]);
                cub::DoubleBuffer<unsigned long long> vb(d_pos[0], d_pos[1]);
                size_t tb = 0;
                BCU(cub::DeviceRadixSort::SortPairs(nullptr, tb, kb, vb, (int64_t)n, 0, 3 * AS_KEY_CHARS, s));
                BCU(tmp.need(tb));
                tb = tmp.cap;
                BCU(cub::DeviceRadixSort::SortPairs(tmp.p, tb, kb, vb, (int64_t)n, 0, 3 * AS_KEY_CHARS, s));
                uint64_t rounds = 0;
                rc = big_refine(h, d_text, N, kb.Current(), vb.Current(), n, tmp, &rounds);
                if (rc != AS_OK) goto out;
                total_rounds += rounds;
                k_big_store_sa<<<big_grid(n, 256, sm), 256, 0, s>>>(vb.Current(), n, row0, (uint32_t *)h->d_sa_lo, (uint8_t *)h->d_sa_hi);
                h->launches++;
                BCU(cudaGetLastError());
                BCU(cudaStreamSynchronize(s));
                row0 += n;
            }
            if (row0 != N) {
                as_set_error("as_gpu_build: partitions hold %llu of %llu suffixes", (unsigned long long)row0, (unsigned long long)N);
                rc = AS_ECUDA;
                goto out;
            }
        }
        for (int i = 0; i < 2; i++) {
            cudaFree(d_keys[i]); d_keys[i] = nullptr;
            cudaFree(d_pos[i]); d_pos[i] = nullptr;
        }
        ix.sa_lo = (const uint32_t *)h->d_sa_lo;
        ix.sa_hi = (const uint8_t *)h->d_sa_hi;

        /* ---- FM blocks ---- */
        unsigned long long *d_primary = d_counter;
        BCU(cudaMemsetAsync(d_primary, 0xff, sizeof(unsigned long long), s));
        BCU(cudaMemsetAsync(d_primary + 1, 0, sizeof(unsigned long long), s));
        BCU(cudaMalloc(&h->d_fm, nblk * sizeof(FmBlock)));
        BCU(cudaMalloc(&h->d_super, nsuper * 4 * sizeof(uint64_t)));
        BCU(cudaMalloc(&d_cnt, 4 * nblk * sizeof(unsigned long long)));
        BCU(cudaMalloc(&d_scan, 4 * nblk * sizeof(unsigned long long)));
        k_big_fm_planes<<<big_grid(nblk, 128, sm), 128, 0, s>>>(d_text, ix.sa_lo, ix.sa_hi, N, (FmBlock *)h->d_fm, d_cnt, nblk, d_primary);
        h->launches++;
        BCU(cudaGetLastError());
        for (int c = 0; c < 4; c++) {
            size_t tb = 0;
            BCU(cub::DeviceScan::ExclusiveSum(nullptr, tb, d_cnt + (uint64_t)c * nblk, d_scan + (uint64_t)c * nblk, (int64_t)nblk, s));
            BCU(tmp.need(tb));
            tb = tmp.cap;
            BCU(cub::DeviceScan::ExclusiveSum(tmp.p, tb, d_cnt + (uint64_t)c * nblk, d_scan + (uint64_t)c * nblk, (int64_t)nblk, s));
        }
        k_big_fm_counts<<<big_grid(nblk, 256, sm), 256, 0, s>>>(d_scan, nblk, (FmBlock *)h->d_fm, (uint64_t *)h->d_super);
        h->launches++;
        BCU(cudaGetLastError());
        unsigned long long prim[2] = {0, 0}, last_scan[4], last_cnt[4];
        BCU(cudaMemcpyAsync(prim, d_primary, sizeof prim, cudaMemcpyDeviceToHost, s));
        for (int c = 0; c < 4; c++) {
            BCU(cudaMemcpyAsync(&last_scan[c], d_scan + (uint64_t)c * nblk + (nblk - 1), 8, cudaMemcpyDeviceToHost, s));
            BCU(cudaMemcpyAsync(&last_cnt[c], d_cnt + (uint64_t)c * nblk + (nblk - 1), 8, cudaMemcpyDeviceToHost, s));
        }
        BCU(cudaStreamSynchronize(s));
        if (prim[0] == ~0ULL || prim[1] != 0) {
            as_set_error("as_gpu_build: the text is not a .cat image (upper-case ACGT before one final '$', asindex.c:97-110)");
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
        cudaFree(d_cnt); d_cnt = nullptr;
        cudaFree(d_scan); d_scan = nullptr;

        /* ---- lc table ---- */
        const uint64_t ncodes = 1ULL << (2 * hlen);
        BCU(cudaMalloc(&h->d_lc, ncodes * sizeof(ulonglong2)));
        BCU(cudaMemsetAsync(h->d_lc, 0, ncodes * sizeof(ulonglong2), s));     /* absent k-mer: (0,0), lchash.c:64-66 */
        k_big_lc<<<big_grid((N + 15) / 16, 256, sm), 256, 0, s>>>(d_text, ix.sa_lo, ix.sa_hi, N, hlen, (ulonglong2 *)h->d_lc);
        h->launches++;
        BCU(cudaGetLastError());
        ix.lc = (const ulonglong2 *)h->d_lc;

        /* ---- text, 2 bits per base ---- */
        const uint64_t nwords = ((N + 15) >> 4) + 4;
        BCU(cudaMalloc(&h->d_text2, nwords * sizeof(uint32_t)));
        BCU(cudaMemsetAsync(h->d_text2, 0, nwords * sizeof(uint32_t), s));
        rc = as_launch_pack_text(h, d_text, N, (uint32_t *)h->d_text2);
        if (rc != AS_OK) goto out;
        ix.text2 = (const uint32_t *)h->d_text2;

        /* ---- mta ---- */
        BCU(cudaMalloc(&h->d_mta, (size_t)(mta_len > 0 ? mta_len : 1) * sizeof(as_mta)));
        if (mta_len > 0) BCU(cudaMemcpyAsync(h->d_mta, mta, (size_t)mta_len * sizeof(as_mta), cudaMemcpyHostToDevice, s));
        ix.mta = (const as_mta *)h->d_mta;
        ix.mta_len = mta_len;
        BCU(cudaStreamSynchronize(s));
        (void)total_rounds;
    }
out:
    cudaFree(d_text);
    cudaFree(d_hist);
    cudaFree(d_counter);
    cudaFree(d_cnt);
    cudaFree(d_scan);
    for (int i = 0; i < 2; i++) {
        cudaFree(d_keys[i]);
        cudaFree(d_pos[i]);
    }
    cudaFree(tmp.p);
    return rc;
}
