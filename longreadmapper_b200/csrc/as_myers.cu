/*
 * as_myers.cu -- edlib-equivalent global (NW) edit distance, the verification path beside GACT.
 * Semantics of edlibAlign(mode NW, task DISTANCE): edlib.cpp:141-260 -> myersCalcEditDistanceNW
 * (edlib.cpp:694-895): unit costs, result = distance, or -1 when k >= 0 and the distance exceeds k.
 * Arithmetic: Myers 1999 / Hyyro 2001 bit-vector column update as in calculateBlock (edlib.cpp:373-408),
 * 64 query rows per word.
 *
 *   k_myers_short  one thread per pair, query <= 256 bases: all words in registers, no band.
 *   k_myers_band   one warp per pair, Ukkonen band |i-j| <= k (edlib.cpp:730-790 keeps the same cells
 *                  alive block-wise): word b of the band is owned by lane b%32 and runs column s-b in
 *                  step s, so the horizontal carry of word b-1 reaches word b by one rotate-shuffle per
 *                  step.  Cells outside the band are bounded from above by +1 deltas exactly as edlib does
 *                  for blocks entering its band, so every distance <= k is exact.
 */
#include "as_common.cuh"

__device__ __forceinline__ uint32_t compress_even(uint32_t x) {
    x &= 0x55555555u;
    x = (x | (x >> 1)) & 0x33333333u;
    x = (x | (x >> 2)) & 0x0F0F0F0Fu;
    x = (x | (x >> 4)) & 0x00FF00FFu;
    x = (x | (x >> 8)) & 0x0000FFFFu;
    return x;
}

/* bit planes of query rows [64b, 64b+64): p0 bit r = low bit of base 64b+r, p1 = high bit; vm = rows < qlen */
__device__ __forceinline__ void query_planes(const uint32_t *qp, int b, int qlen, uint64_t &p0, uint64_t &p1, uint64_t &vm) {
    p0 = 0;
    p1 = 0;
#pragma unroll
    for (int w = 0; w < 4; w++) {
        uint32_t x = __ldg(qp + 4 * b + w);
        p0 |= (uint64_t)compress_even(x) << (16 * w);
        p1 |= (uint64_t)compress_even(x >> 1) << (16 * w);
    }
    int valid = qlen - 64 * b;
    vm = valid >= 64 ? ~0ULL : (valid <= 0 ? 0ULL : ((1ULL << valid) - 1ULL));
}

__device__ __forceinline__ uint64_t eq_mask(uint32_t c, uint64_t p0, uint64_t p1, uint64_t vm) {
    return ((c & 1u) ? p0 : ~p0) & ((c & 2u) ? p1 : ~p1) & vm;
}

/* one word, one column (edlib.cpp:373-408); returns hout in {-1,0,1} taken at bit 63 */
__device__ __forceinline__ int myers_block(uint64_t Eq, uint64_t &Pv, uint64_t &Mv, int hin) {
    const uint64_t hneg = hin < 0 ? 1ULL : 0ULL;
    const uint64_t Xv = Eq | Mv;
    Eq |= hneg;
    const uint64_t Xh = (((Eq & Pv) + Pv) ^ Pv) | Eq;
    uint64_t Ph = Mv | ~(Xh | Pv);
    uint64_t Mh = Pv & Xh;
    const int hout = (int)(Ph >> 63) - (int)(Mh >> 63);
    Ph <<= 1;
    Mh <<= 1;
    Mh |= hneg;
    if (hin > 0) Ph |= 1ULL;
    Pv = Mh | ~(Xv | Ph);
    Mv = Ph & Xv;
    return hout;
}

struct MyersArgs {
    const uint32_t *qpack;
    size_t q_wstride;
    const uint32_t *qlens;
    const uint32_t *tbase;
    const uint64_t *toff;
    const uint32_t *tlens;       /* NULL: tlen = qlen */
    size_t n;
    const int32_t *ks;           /* NULL with k_from_len: k = (int)(0.05*len*2) (alnmain.c:428); NULL otherwise: unbounded */
    int k_from_len;
    int32_t *dist;               /* may be NULL */
    as_result *res;              /* may be NULL; writes .myers */
    unsigned int *work_counter;
};

__device__ __forceinline__ int pair_k(const MyersArgs &a, size_t i, int qlen) {
    if (a.ks) return a.ks[i];
    if (a.k_from_len) return (int)(0.05 * (double)qlen * 2);
    return -1;
}

__global__ void __launch_bounds__(128) k_myers_short(MyersArgs a) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < a.n; i += (size_t)gridDim.x * blockDim.x) {
        const uint64_t toff = a.toff ? a.toff[i] : 0;
        if (toff == 0xFFFFFFFFFFFFFFFFULL) continue;
        const int qlen = (int)a.qlens[i];
        const int tlen = a.tlens ? (int)a.tlens[i] : qlen;
        const int k = pair_k(a, i, qlen);
        const uint32_t *qp = a.qpack + i * a.q_wstride;
        int d;
        if (qlen == 0) d = tlen;
        else if (tlen == 0) d = qlen;
        else {
            const int nb = (qlen + 63) >> 6;
            uint64_t p0[4], p1[4], vm[4], Pv[4], Mv[4];
#pragma unroll
            for (int b = 0; b < 4; b++) {
                if (b < nb) query_planes(qp, b, qlen, p0[b], p1[b], vm[b]);
                else { p0[b] = p1[b] = vm[b] = 0; }
                Pv[b] = ~0ULL;
                Mv[b] = 0;
            }
            int score = 64 * nb;                         /* D at the last row of the last word, column 0 */
            uint32_t tw = 0;
            for (int j = 0; j < tlen; j++) {
                const uint64_t tp = toff + (uint64_t)j;
                if ((tp & 15u) == 0 || j == 0) tw = __ldg(a.tbase + (tp >> 4));
                const uint32_t c = (tw >> (2u * ((uint32_t)tp & 15u))) & 3u;
                int h = 1;                               /* D[0][j] - D[0][j-1] */
#pragma unroll
                for (int b = 0; b < 4; b++)
                    if (b < nb) h = myers_block(eq_mask(c, p0[b], p1[b], vm[b]), Pv[b], Mv[b], h);
                score += h;
            }
            /* walk up from the word's last row to row qlen */
            const int lastbit = (qlen - 1) & 63;
            const uint64_t above = lastbit == 63 ? 0ULL : (~0ULL << (lastbit + 1));
            uint64_t pv = Pv[0], mv = Mv[0];
#pragma unroll
            for (int b = 1; b < 4; b++)
                if (b == nb - 1) { pv = Pv[b]; mv = Mv[b]; }
            d = score - (__popcll(pv & above) - __popcll(mv & above));
        }
        const int out = (k >= 0 && d > k) ? -1 : d;
        if (a.dist) a.dist[i] = out;
        if (a.res) a.res[i].myers = out;
    }
}

template <int W>
__global__ void __launch_bounds__(128) k_myers_band(MyersArgs a) {
    const int l = threadIdx.x & 31;
    for (;;) {
        uint32_t pi = 0;
        if (l == 0) pi = atomicAdd(a.work_counter, 1u);
        pi = __shfl_sync(0xFFFFFFFFu, pi, 0);
        if (pi >= a.n) break;
        const uint64_t toff = a.toff ? a.toff[pi] : 0;
        if (toff == 0xFFFFFFFFFFFFFFFFULL) continue;
        const int qlen = (int)a.qlens[pi];
        const int tlen = a.tlens ? (int)a.tlens[pi] : qlen;
        const int kin = pair_k(a, pi, qlen);
        const uint32_t *qp = a.qpack + (size_t)pi * a.q_wstride;
        int out;
        const int diff = qlen > tlen ? qlen - tlen : tlen - qlen;
        if (qlen == 0 || tlen == 0) {
            const int d = qlen + tlen;
            out = (kin >= 0 && d > kin) ? -1 : d;
        } else if (kin >= 0 && diff > kin) {
            out = -1;
        } else {
            const int k = kin >= 0 ? kin : (qlen > tlen ? qlen : tlen);
            const int nb = (qlen + 63) >> 6;
            const int B = nb - 1;
            uint64_t p0[W], p1[W], vm[W], Pv[W], Mv[W];
            int cur[W], score[W], hout[W];
#pragma unroll
            for (int w = 0; w < W; w++) {
                cur[w] = l + 32 * w;
                score[w] = 0;
                hout[w] = 0;
                p0[w] = p1[w] = vm[w] = Mv[w] = 0;
                Pv[w] = ~0ULL;
            }
            int d_final = 0;
            const int last_step = tlen + B;
            for (int s = 1; s <= last_step; s++) {
                /* carry of word b-1 (same column, previous step): rotate by one lane; lane 0 takes the previous slot */
                int rh[W];
#pragma unroll
                for (int w = 0; w < W; w++) rh[w] = __shfl_sync(0xFFFFFFFFu, (score[w] << 2) | (hout[w] + 1), (l + 31) & 31);
#pragma unroll
                for (int w = 0; w < W; w++) {
                    const int b = cur[w];
                    if (b > B) continue;
                    const int j = s - b;
                    const int jf = 64 * b + 1 - k < 1 ? 1 : 64 * b + 1 - k;
                    int jl = 64 * b + 64 + k;
                    if (jl > tlen || jl < 0) jl = tlen;
                    if (j < jf) continue;
                    if (j > jl) {                        /* this word left the band: hand the slot to word b + 32W */
                        cur[w] = b + 32 * W;
                        hout[w] = 0;
                        continue;
                    }
                    int rv = rh[w];
                    if (l == 0) rv = rh[(w + W - 1) % W];
                    if (j == jf) {
                        query_planes(qp, b, qlen, p0[w], p1[w], vm[w]);
                        Pv[w] = ~0ULL;
                        Mv[w] = 0;
                        /* D at this word's last row, column jf-1 (edlib.cpp:757-763) */
                        score[w] = (jf == 1) ? 64 * (b + 1) : (rv >> 2) - ((rv & 3) - 1) + 64;
                    }
                    const int top = j - k - 1;
                    const int fb = top <= 0 ? 0 : top >> 6;
                    const int hin = (b == fb) ? 1 : (rv & 3) - 1;
                    const uint32_t c = as_base_at(a.tbase, toff + (uint64_t)(j - 1));
                    const int ho = myers_block(eq_mask(c, p0[w], p1[w], vm[w]), Pv[w], Mv[w], hin);
                    hout[w] = ho;
                    score[w] += ho;
                    if (b == B && j == tlen) {
                        const int lastbit = (qlen - 1) & 63;
                        const uint64_t above = lastbit == 63 ? 0ULL : (~0ULL << (lastbit + 1));
                        d_final = score[w] - (__popcll(Pv[w] & above) - __popcll(Mv[w] & above));
                    }
                }
            }
            const int d = __shfl_sync(0xFFFFFFFFu, d_final, B & 31);
            out = (kin >= 0 && d > kin) ? -1 : d;
        }
        if (l == 0) {
            if (a.dist) a.dist[pi] = out;
            if (a.res) a.res[pi].myers = out;
        }
    }
}

int as_launch_myers(as_gpu *h, const uint32_t *d_qpack, size_t q_wstride, const uint32_t *d_qlens,
                    const uint32_t *d_tbase, const uint64_t *d_toff, const uint32_t *d_tlens, size_t n,
                    uint32_t max_qlen, const int32_t *d_ks, int32_t *d_dist, as_result *d_res) {
    /* d_ks == NULL && d_res != NULL: map mode, k from the read length.  max band half-width for sizing: */
    if (n == 0) return AS_OK;
    MyersArgs a;
    a.qpack = d_qpack; a.q_wstride = q_wstride; a.qlens = d_qlens;
    a.tbase = d_tbase; a.toff = d_toff; a.tlens = d_tlens; a.n = n;
    a.ks = d_ks; a.k_from_len = (d_ks == nullptr && d_res != nullptr);
    a.dist = d_dist; a.res = d_res;
    AS_TRY(as_scratch(h, &h->misc[4], 64));
    AS_CUDA(cudaMemsetAsync(h->misc[4].p, 0, 64, h->stream));
    a.work_counter = (unsigned int *)h->misc[4].p;
    if (max_qlen <= 256) {
        uint64_t blocks = (n + 127) / 128, cap = (uint64_t)h->sm_count * 16;
        k_myers_short<<<(unsigned)(blocks < cap ? blocks : cap), 128, 0, h->stream>>>(a);
    } else {
        /* band words: map mode k = 0.1*len; otherwise assume the full matrix (exact for any k) */
        uint64_t nb = ((uint64_t)max_qlen + 63) / 64;
        uint64_t band_words = nb;
        if (a.k_from_len) {
            uint64_t k = (uint64_t)(0.05 * (double)max_qlen * 2);
            uint64_t bw = (2 * k + 1) / 64 + 2;
            if (bw < band_words) band_words = bw;
        }
        int W = (int)((band_words + 31) / 32) + 1;
        uint64_t blocks = (n + 3) / 4, cap = (uint64_t)h->sm_count * 8;
        unsigned grid = (unsigned)(blocks < cap ? blocks : cap);
        if (W <= 2) k_myers_band<2><<<grid, 128, 0, h->stream>>>(a);
        else if (W <= 3) k_myers_band<3><<<grid, 128, 0, h->stream>>>(a);
        else if (W <= 4) k_myers_band<4><<<grid, 128, 0, h->stream>>>(a);
        else if (W <= 6) k_myers_band<6><<<grid, 128, 0, h->stream>>>(a);
        else if (W <= 9) k_myers_band<9><<<grid, 128, 0, h->stream>>>(a);
        else {
            as_set_error("edit distance: query of %u bases needs a %llu-word band; at most 256 words are supported",
                         max_qlen, (unsigned long long)band_words);
            return AS_ERANGE;
        }
    }
    h->launches++;
    AS_CUDA(cudaGetLastError());
    return AS_OK;
}

/* ------------------------------------------------------------------------------------------------------------------------
 * edlib's alignment path (EDLIB_MODE_NW, EDLIB_TASK_PATH): edlibAlign edlib.cpp:141-260 -> obtainAlignment :1110-1150 ->
 * obtainAlignmentTraceback :909-1108.  The traceback starts at the last cell and asks, in this order: is the cell above one
 * less (INSERT 1), is the cell to the left one less (DELETE 2), else the diagonal (MATCH 0 / MISMATCH 3).  "The cell above is
 * one less" is the vertical delta the column step leaves in Pv, "the cell to the left is one less" the horizontal delta Ph
 * it computes on the way (before the shift): the fill keeps both words of every (column, block), the walk reads two bits per
 * step and compares two bases for the diagonal's label.  One thread per pair, queries of at most 256 bases (4 words in
 * registers), planes [column][block][Pv | Ph][pair of the launch] so that a warp's stores are contiguous.
 * ---------------------------------------------------------------------------------------------------------------------- */
struct MyersPathArgs {
    MyersArgs m;
    size_t first, count;         /* pairs [first, first + count) of the batch */
    int nbmax;                   /* blocks of the longest query of the batch */
    uint64_t *planes;            /* [tlen_max][nbmax][2][count] */
    uint8_t *ops;                /* row i at ops + i * ops_stride, forward order */
    size_t ops_stride;
    int32_t *n_ops;
};

__global__ void __launch_bounds__(128) k_myers_path(MyersPathArgs pa) {
    const MyersArgs &a = pa.m;
    for (size_t li = blockIdx.x * (size_t)blockDim.x + threadIdx.x; li < pa.count; li += (size_t)gridDim.x * blockDim.x) {
        const size_t i = pa.first + li;
        const uint64_t toff = a.toff[i];
        const int qlen = (int)a.qlens[i], tlen = (int)a.tlens[i];
        const int k = a.ks ? a.ks[i] : -1;
        const uint32_t *qp = a.qpack + i * a.q_wstride;
        uint8_t *ops = pa.ops + i * pa.ops_stride;
        uint64_t *pl = pa.planes + li;
        const size_t cs = pa.count;                     /* stride between consecutive plane words of one pair */
        const int nb = (qlen + 63) >> 6;
        int d;
        if (qlen == 0) d = tlen;
        else if (tlen == 0) d = qlen;
        else {
            uint64_t p0[4], p1[4], vm[4], Pv[4], Mv[4];
#pragma unroll
            for (int b = 0; b < 4; b++) {
                if (b < nb) query_planes(qp, b, qlen, p0[b], p1[b], vm[b]);
                else { p0[b] = p1[b] = vm[b] = 0; }
                Pv[b] = ~0ULL;
                Mv[b] = 0;
            }
            int score = 64 * nb;
            uint32_t tw = 0;
            for (int j = 0; j < tlen; j++) {
                const uint64_t tp = toff + (uint64_t)j;
                if ((tp & 15u) == 0 || j == 0) tw = __ldg(a.tbase + (tp >> 4));
                const uint32_t c = (tw >> (2u * ((uint32_t)tp & 15u))) & 3u;
                int h = 1;
#pragma unroll
                for (int b = 0; b < 4; b++) {
                    if (b < nb) {
                        /* myers_block, with the horizontal +1 word kept before it is shifted */
                        uint64_t Eq = eq_mask(c, p0[b], p1[b], vm[b]);
                        const uint64_t hneg = h < 0 ? 1ULL : 0ULL;
                        const uint64_t Xv = Eq | Mv[b];
                        Eq |= hneg;
                        const uint64_t Xh = (((Eq & Pv[b]) + Pv[b]) ^ Pv[b]) | Eq;
                        uint64_t Ph = Mv[b] | ~(Xh | Pv[b]);
                        uint64_t Mh = Pv[b] & Xh;
                        const uint64_t ph_keep = Ph;
                        const int hout = (int)(Ph >> 63) - (int)(Mh >> 63);
                        Ph <<= 1;
                        Mh <<= 1;
                        Mh |= hneg;
                        if (h > 0) Ph |= 1ULL;
                        Pv[b] = Mh | ~(Xv | Ph);
                        Mv[b] = Ph & Xv;
                        h = hout;
                        uint64_t *w = pl + ((size_t)(j * pa.nbmax + b) * 2) * cs;
                        w[0] = Pv[b];
                        w[cs] = ph_keep;
                    }
                }
                score += h;
            }
            const int lastbit = (qlen - 1) & 63;
            const uint64_t above = lastbit == 63 ? 0ULL : (~0ULL << (lastbit + 1));
            uint64_t pv = Pv[0], mv = Mv[0];
#pragma unroll
            for (int b = 1; b < 4; b++)
                if (b == nb - 1) { pv = Pv[b]; mv = Mv[b]; }
            d = score - (__popcll(pv & above) - __popcll(mv & above));
        }
        const bool over = k >= 0 && d > k;
        if (a.dist) a.dist[i] = over ? -1 : d;
        int n = 0;
        if (!over) {
            /* two walks: the first counts the columns, the second writes them in forward order */
            for (int pass = 0; pass < 2; pass++) {
                int ci = qlen, cj = tlen, m = 0;
                while (ci > 0 || cj > 0) {
                    uint32_t op;
                    if (ci == 0) op = 2u;
                    else if (cj == 0) op = 1u;
                    else {
                        const int b = (ci - 1) >> 6, r = (ci - 1) & 63;
                        const uint64_t *w = pl + ((size_t)((cj - 1) * pa.nbmax + b) * 2) * cs;
                        if ((w[0] >> r) & 1ULL) op = 1u;                          /* edlib.cpp:975 */
                        else if ((w[cs] >> r) & 1ULL) op = 2u;                    /* :1005 */
                        else op = as_base_at(qp, (uint64_t)(ci - 1)) == as_base_at(a.tbase, toff + (uint64_t)(cj - 1)) ? 0u : 3u;   /* :1038 */
                    }
                    if (pass) ops[n - 1 - m] = (uint8_t)op;
                    m++;
                    ci -= (op != 2u);
                    cj -= (op != 1u);
                }
                n = m;
            }
        }
        pa.n_ops[i] = n;
    }
}

int as_launch_myers_path(as_gpu *h, const uint32_t *d_qpack, size_t q_wstride, const uint32_t *d_qlens, const uint32_t *d_tbase,
                         const uint64_t *d_toff, const uint32_t *d_tlens, size_t n, uint32_t max_qlen, uint32_t max_tlen,
                         const int32_t *d_ks, int32_t *d_dist, uint8_t *d_ops, size_t ops_stride, int32_t *d_nops) {
    if (n == 0) return AS_OK;
    MyersPathArgs pa;
    pa.m.qpack = d_qpack; pa.m.q_wstride = q_wstride; pa.m.qlens = d_qlens;
    pa.m.tbase = d_tbase; pa.m.toff = d_toff; pa.m.tlens = d_tlens; pa.m.n = n;
    pa.m.ks = d_ks; pa.m.k_from_len = 0; pa.m.dist = d_dist; pa.m.res = nullptr; pa.m.work_counter = nullptr;
    pa.nbmax = (int)((max_qlen + 63) / 64);
    if (pa.nbmax < 1) pa.nbmax = 1;
    pa.ops = d_ops; pa.ops_stride = ops_stride; pa.n_ops = d_nops;
    /* planes of one pair: max_tlen * nbmax * 16 bytes; the batch goes through in slices of at most 1 GB of them */
    const size_t per_pair = (size_t)(max_tlen ? max_tlen : 1) * (size_t)pa.nbmax * 16;
    size_t slice = ((size_t)1 << 30) / per_pair;
    if (slice < 128) slice = 128;
    if (slice > n) slice = n;
    AS_TRY(as_scratch(h, &h->path_planes, slice * per_pair));
    pa.planes = (uint64_t *)h->path_planes.p;
    for (size_t first = 0; first < n; first += slice) {
        pa.first = first;
        pa.count = n - first < slice ? n - first : slice;
        uint64_t blocks = (pa.count + 127) / 128, cap = (uint64_t)h->sm_count * 16;
        k_myers_path<<<(unsigned)(blocks < cap ? blocks : cap), 128, 0, h->stream>>>(pa);
        h->launches++;
        AS_CUDA(cudaGetLastError());
    }
    return AS_OK;
}

/* ------------------------------------------------------------------------------------------------------------------------
 * as_edit_distance_device_noalloc: k_myers_short for callers that cannot allocate (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * The scratch as_edit_distance_batch takes from the handle's growable buffers -- the 2-bit packed queries and
 * targets, the targets' base offsets, the checked lengths -- lives in ONE file-scope device array of fixed size, so the call
 * allocates nothing: AS_NOALLOC_MAX_PAIRS pairs of at most AS_NOALLOC_MAX_LEN bases each, ~5.8 MB held by every GPU that
 * loads the library until the process exits.  Every call shares the array (it keeps the last call's contents; nothing in it
 * has to start at zero, every word a call reads it has written first), so calls that overlap on the GPU race.
 * ---------------------------------------------------------------------------------------------------------------------- */
#define NA_WS ((AS_NOALLOC_MAX_LEN + 15) / 16 + 4)          /* packed words per sequence, with the readers' slack */
struct NoallocScratch {
    uint32_t qpack[(size_t)AS_NOALLOC_MAX_PAIRS * NA_WS];
    uint32_t tpack[(size_t)AS_NOALLOC_MAX_PAIRS * NA_WS];
    unsigned long long toff[AS_NOALLOC_MAX_PAIRS];
    uint32_t qlens[AS_NOALLOC_MAX_PAIRS];
    uint32_t tlens[AS_NOALLOC_MAX_PAIRS];
    uint32_t bad[AS_NOALLOC_MAX_PAIRS];                      /* 1: a length of the pair exceeds AS_NOALLOC_MAX_LEN */
};
static __device__ __align__(256) NoallocScratch g_noalloc_scratch;

/* thread per packed word: word w of pair i of the queries (z == 0) or the targets (z == 1) */
__global__ void __launch_bounds__(256) k_noalloc_pack(const char *__restrict__ q, size_t q_stride, const uint32_t *__restrict__ qlens,
                                                      const char *__restrict__ t, size_t t_stride, const uint32_t *__restrict__ tlens,
                                                      size_t n, NoallocScratch *s) {
    const size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (idx >= n * NA_WS) return;
    const size_t i = idx / NA_WS;
    const uint32_t w = (uint32_t)(idx - i * NA_WS);
    const uint32_t ql = qlens[i], tl = tlens[i];
    const bool bad = ql > AS_NOALLOC_MAX_LEN || tl > AS_NOALLOC_MAX_LEN;
    if (w == 0) {
        s->qlens[i] = bad ? 0u : ql;
        s->tlens[i] = bad ? 0u : tl;
        s->bad[i] = bad ? 1u : 0u;
        s->toff[i] = (unsigned long long)i * (16ull * NA_WS);
    }
    uint32_t vq = 0, vt = 0;
    if (!bad && 16u * w < ql) vq = as_pack16(reinterpret_cast<const uint8_t *>(q) + i * q_stride + 16u * w, min(16u, ql - 16u * w));
    if (!bad && 16u * w < tl) vt = as_pack16(reinterpret_cast<const uint8_t *>(t) + i * t_stride + 16u * w, min(16u, tl - 16u * w));
    s->qpack[idx] = vq;
    s->tpack[idx] = vt;
}

__global__ void __launch_bounds__(256) k_noalloc_flag(const NoallocScratch *s, size_t n, int32_t *__restrict__ dist) {
    const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i < n && s->bad[i]) dist[i] = AS_ERANGE;
}

extern "C" int as_edit_distance_device_noalloc(void *cuda_stream, const char *d_q, size_t q_stride, const uint32_t *d_qlens,
                                               const char *d_t, size_t t_stride, const uint32_t *d_tlens, size_t n,
                                               const int32_t *d_ks, int32_t *d_dist) {
    /* per device: the array's DEVICE address (its name is a host-side stand-in in host code) and the SM count */
    static NoallocScratch *s_ptr[64];
    static int s_sm[64];
    if (!d_q || !d_t || !d_qlens || !d_tlens || !d_dist) return AS_EINVAL;
    if (n > AS_NOALLOC_MAX_PAIRS) {
        as_set_error("as_edit_distance_device_noalloc: %zu pairs; the static scratch holds at most %d (of at most %d bases each)", n,
                     AS_NOALLOC_MAX_PAIRS, AS_NOALLOC_MAX_LEN);
        return AS_ERANGE;
    }
    if (q_stride < 1 || t_stride < 1) return AS_EINVAL;
    int dev = -1;
    AS_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return AS_EINVAL;
    if (!s_ptr[dev]) {                   /* set-up, once per device, with that device current */
        int major = 0, sm = 0;
        AS_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
        if (major != 10) {
            as_set_error("as_edit_distance_device_noalloc: device %d is not sm_100 (no CPU fallback)", dev);
            return AS_ENODEV;
        }
        AS_CUDA(cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, dev));
        NoallocScratch *p = nullptr;
        AS_CUDA(cudaGetSymbolAddress((void **)&p, g_noalloc_scratch));
        s_sm[dev] = sm;
        s_ptr[dev] = p;
    }
    if (n == 0) return AS_OK;
    NoallocScratch *s = s_ptr[dev];
    cudaStream_t st = (cudaStream_t)cuda_stream;
    const size_t words = n * NA_WS;
    k_noalloc_pack<<<(unsigned)((words + 255) / 256), 256, 0, st>>>(d_q, q_stride, d_qlens, d_t, t_stride, d_tlens, n, s);
    AS_CUDA(cudaGetLastError());
    MyersArgs a;
    a.qpack = s->qpack; a.q_wstride = NA_WS; a.qlens = s->qlens;
    a.tbase = s->tpack; a.toff = (const uint64_t *)s->toff; a.tlens = s->tlens; a.n = n;
    a.ks = d_ks; a.k_from_len = 0;
    a.dist = d_dist; a.res = nullptr;
    a.work_counter = nullptr;            /* k_myers_short takes a grid-stride loop, no work counter */
    const uint64_t blocks = (n + 127) / 128, cap = (uint64_t)s_sm[dev] * 16;
    k_myers_short<<<(unsigned)(blocks < cap ? blocks : cap), 128, 0, st>>>(a);
    AS_CUDA(cudaGetLastError());
    k_noalloc_flag<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(s, n, d_dist);
    AS_CUDA(cudaGetLastError());
    return AS_OK;
}
