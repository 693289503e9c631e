/*
 * as_gact_bp.cu -- bit-parallel GACT: the aligner behind cigar_align (mutils.c:94-105 -> simple_gact, the
 * un-vendored lisanhu/GACT; spec: DESIGN.md "GACT spec", oracle: oracle/oracle.c orc_gact).
 *
 * Same tiles, same DP matrix, same traceback priority (diag, up, left) as the warp-per-read kernel in
 * as_align.cu -- but the tile is filled with the Myers/Hyyro bit-vector recurrence (the one edlib's
 * calculateBlock uses, edlib.cpp:373-408): 32 DP rows per 32-bit word, ~25 integer instructions per word instead of
 * ~10 per CELL.  One read per LANE; a warp walks the tiles of its 32 reads in lock step.
 *
 * What makes the traceback exact: with dv = D[i][j]-D[i-1][j] (+1 <=> bit of Pv'), D0 <=> D[i][j]==D[i-1][j-1],
 *   diag taken  <=> D[i-1][j-1] + neq == D[i][j]  <=> Eq | ~D0
 *   up taken    <=> not diag and D[i-1][j]+1 == D[i][j] <=> Pv'
 *   left taken  otherwise
 * so the 2-bit op code of every cell ('=' 0, 'I'/up 1, 'D'/left 2, 'X' 3) is two LOP3s per word:
 *   lo = ~Eq & (Pv' | ~D0)      hi = ~Eq & (~Pv' | ~D0)
 * The two planes of every (column, word) go to a per-warp slab in HBM laid out [column][word][lane] as uint2, so a
 * warp's store is one 256-byte line and the traceback of lanes that sit on the same column reads one line.
 */
#include "as_common.cuh"
#include <stdlib.h>

#define BP_THREADS 128
#define BP_PREFETCH 12

/* even bits of x (bit 2r) -> bit r of the low word; odd bits -> high word */
__device__ __forceinline__ uint32_t compress_even(uint64_t x) {
    x &= 0x5555555555555555ULL;
    x = (x | (x >> 1)) & 0x3333333333333333ULL;
    x = (x | (x >> 2)) & 0x0F0F0F0F0F0F0F0FULL;
    x = (x | (x >> 4)) & 0x00FF00FF00FF00FFULL;
    x = (x | (x >> 8)) & 0x0000FFFF0000FFFFULL;
    x = (x | (x >> 16));
    return (uint32_t)x;
}

/* bytes [s0, s1) of `acc` to row offsets pos+s, skipping what falls outside the row */
__device__ __forceinline__ void put_bytes(uint8_t *ops, long pos, uint32_t acc, int s0, int s1, size_t stride) {
    for (int s = s0; s < s1; s++) {
        const long q = pos + s;
        if (q >= 0 && (size_t)q < stride) ops[q] = (uint8_t)(acc >> (8 * s));
    }
}

template <int W>
__global__ void __launch_bounds__(BP_THREADS) k_gact_bp(GactArgs a, uint2 *__restrict__ planes_all, size_t planes_per_warp,
                                                        uint32_t *__restrict__ tb_all, size_t tb_per_warp) {
    const uint32_t lane = threadIdx.x & 31u;
    const size_t gwarp = ((size_t)blockIdx.x * BP_THREADS + threadIdx.x) >> 5;
    uint2 *planes = planes_all + gwarp * planes_per_warp + lane;          /* [(col*W + w)*32] */
    uint32_t *tb = tb_all + gwarp * tb_per_warp + lane;                    /* [word*32]: 16 ops per word, walk order */
    const int T = a.T, inner = a.T - a.O;
    unsigned long long st_cells = 0;

    for (;;) {
        uint32_t base = 0;
        if (lane == 0) base = atomicAdd(a.work_counter, 32u);
        base = __shfl_sync(0xFFFFFFFFu, base, 0);
        if (base >= a.n) break;
        const size_t read = (size_t)base + lane;
        bool done = read >= a.n;
        int qlen = 0, tlen = 0;
        uint64_t toff = 0;
        if (!done) {
            qlen = (int)a.qlens[read];
            tlen = a.tlens ? (int)a.tlens[read] : qlen;
            toff = a.toff ? a.toff[read] : 0;
            /* seq_lookup failed: flag 4 was set by PART 2; reference behaviour is undefined (alnmain.c:409,430-446) */
            if (toff == AS_TOFF_NONE) done = true;
        }
        const bool skipped = done;
        const uint32_t *qp = a.qpack + (done ? 0 : read * a.q_wstride);
        uint8_t *ops = (a.ops && !done) ? a.ops + read * a.ops_stride : nullptr;
        int i0 = 0, j0 = 0, n = 0, score = 0;
        /* column ops leave as aligned 32-bit words: omis = misalignment of the row, oacc = the word being filled */
        const uint32_t omis = (uint32_t)(reinterpret_cast<uintptr_t>(ops) & 3u);
        uint32_t oacc = 0;
        if (!done && qlen == 0 && tlen == 0) done = true;

        while (__any_sync(0xFFFFFFFFu, !done)) {
            /* ---- tile geometry of this lane ---- */
            int Tq = 0, Tt = 0;
            bool q_done = true, t_done = true;
            if (!done) {
                Tq = qlen - i0 < T ? qlen - i0 : T;
                Tt = tlen - j0 < T ? tlen - j0 : T;
                q_done = (i0 + Tq == qlen);
                t_done = (j0 + Tt == tlen);
            }
            const bool last = q_done && t_done;
            /* ---- query bit planes of the tile rows ---- */
            uint32_t qlo[W], qhi[W], Pv[W], Mv[W];
#pragma unroll
            for (int w = 0; w < W; w++) {
                uint64_t x = 0;
                if (32 * w < Tq) x = as_bases16(qp, (uint64_t)(i0 + 32 * w));
                if (32 * w + 16 < Tq) x |= (uint64_t)as_bases16(qp, (uint64_t)(i0 + 32 * w + 16)) << 32;
                qlo[w] = compress_even(x);
                qhi[w] = compress_even(x >> 1);
                Pv[w] = 0xFFFFFFFFu;                    /* D[i][0] = i */
                Mv[w] = 0u;
            }
            const int wlast = Tq > 0 ? (Tq - 1) >> 5 : 0;
            const uint32_t blast = Tq > 0 ? (uint32_t)(Tq - 1) & 31u : 0u;
            int sc = Tq;                                            /* D[Tq][j] */
            int rowkey = (Tq << 12) | (Tt << 1);                    /* bottom-row candidates (edge 0), column 0 first */
            int maxTt = Tt;
#pragma unroll
            for (int d = 16; d >= 1; d >>= 1) maxTt = max(maxTt, __shfl_xor_sync(0xFFFFFFFFu, maxTt, d));
            /* the bottom row only matters to tiles that leave query rows behind (never for reads of one tile) */
            const bool track_row = __any_sync(0xFFFFFFFFu, !done && !q_done);

            /* ---- fill: one column per iteration, W words per column ---- */
            uint32_t tw = 0;
            uint2 *pcol = planes;
            for (int j = 1; j <= maxTt; j++, pcol += W * 32) {
                if (j <= Tt) {
                    if (((j - 1) & 15) == 0) tw = as_bases16(a.tbase, toff + (uint64_t)(j0 + j - 1));
                    const uint32_t tl = 0u - (tw & 1u), th = 0u - ((tw >> 1) & 1u);
                    tw >>= 2;
                    uint32_t hpos = 1u, hneg = 0u;                   /* D[0][j] - D[0][j-1] = +1 */
#pragma unroll
                    for (int w = 0; w < W; w++) {
                        const uint32_t eq = ~((qlo[w] ^ tl) | (qhi[w] ^ th));
                        const uint32_t pv = Pv[w], mv = Mv[w];
                        const uint32_t xv = eq | mv;
                        const uint32_t eq2 = eq | hneg;
                        const uint32_t xh = (((eq2 & pv) + pv) ^ pv) | eq2;
                        uint32_t ph = mv | ~(xh | pv);
                        uint32_t mh = pv & xh;
                        if (track_row) {
                            if (w == wlast && Tq > 0) sc += (int)((ph >> blast) & 1u) - (int)((mh >> blast) & 1u);
                        }
                        const uint32_t hp_out = ph >> 31, hn_out = mh >> 31;
                        ph = (ph << 1) | hpos;
                        mh = (mh << 1) | hneg;
                        const uint32_t npv = mh | ~(xv | ph);
                        Pv[w] = npv;
                        Mv[w] = ph & xv;
                        const uint32_t nd0 = ~(xh | mv);            /* ~D0 */
                        uint2 o;
                        o.x = ~eq & (npv | nd0);
                        o.y = ~eq & (~npv | nd0);
                        pcol[w * 32] = o;
                        hpos = hp_out;
                        hneg = hn_out;
                    }
                    if (track_row) rowkey = min(rowkey, (sc << 12) | ((Tt - j) << 1));
                }
            }
            __syncwarp();
            if (!done) st_cells += (unsigned long long)Tq * (unsigned long long)Tt;

            /* ---- end cell (non-last tiles): min (D, distance from the corner, edge) over the interior edges ---- */
            int ei = Tq, ej = Tt;
            if (!done && !last) {
                int best = 0x7FFFFFFF;
                if (!q_done) best = rowkey;
                if (!t_done) {
                    int v = Tt;                                      /* D[0][Tt] */
                    int ck = (v << 12) | (Tq << 1) | 1;
#pragma unroll
                    for (int w = 0; w < W; w++) {
                        if (32 * w < Tq) {
                            const uint32_t pv = Pv[w], mv = Mv[w];
                            const int lim = Tq - 32 * w < 32 ? Tq - 32 * w : 32;
                            for (int b = 0; b < lim; b++) {
                                v += (int)((pv >> b) & 1u) - (int)((mv >> b) & 1u);
                                const int i = 32 * w + b + 1;
                                ck = min(ck, (v << 12) | ((Tq - i) << 1) | 1);
                            }
                        }
                    }
                    best = min(best, ck);
                }
                const int dist = (best >> 1) & 0x7FF;
                if (best & 1) { ei = Tq - dist; ej = Tt; } else { ei = Tq; ej = Tt - dist; }
            }

            /* ---- traceback: ops of the walk, packed 16 per word, into the warp's slab ---- */
            int ci = done ? 0 : ei, cj = done ? 0 : ej, m = 0, ai = ei, aj = ej;
            bool anchored = false;
            uint32_t acc = 0;
            while (__any_sync(0xFFFFFFFFu, (ci | cj) != 0)) {
                if ((ci | cj) != 0) {
                    uint32_t op;
                    if (ci == 0) op = 2;
                    else if (cj == 0) op = 1;
                    else {
                        const int w = (ci - 1) >> 5, b = (ci - 1) & 31;
                        const uint2 v = planes[((size_t)(cj - 1) * W + w) * 32];
                        op = ((v.x >> b) & 1u) | (((v.y >> b) & 1u) << 1);
                    }
                    const int si = ci - (op != 2u), sj = cj - (op != 1u);
                    /* the walk is a chain of dependent loads of planes written long ago: pull the line it will need
                     * BP_PREFETCH columns further down the diagonal into L2 (lanes of a warp share that line) */
                    if (sj > BP_PREFETCH && si > BP_PREFETCH) {
                        const uint2 *pf = planes + ((size_t)(sj - 1 - BP_PREFETCH) * W + ((si - 1 - BP_PREFETCH) >> 5)) * 32;
                        asm volatile("prefetch.global.L2 [%0];" ::"l"(pf));
                    }
                    if (last || (si < inner && sj < inner)) {
                        if (!anchored) { ai = ci; aj = cj; anchored = true; }
                        acc |= op << (2 * (m & 15));
                        m++;
                        if ((m & 15) == 0) { tb[(size_t)((m >> 4) - 1) * 32] = acc; acc = 0; }
                        score += (op != 0u);
                    }
                    ci = si;
                    cj = sj;
                }
            }
            if (m & 15) tb[(size_t)(m >> 4) * 32] = acc;
            __syncwarp();
            /* ---- emit the committed ops in forward order ---- */
            int mym = 0;
            if (!done) {
                if (m == 0 && !last) { score = -1; done = true; }
                else mym = m;
            }
            int maxm = mym;
#pragma unroll
            for (int d = 16; d >= 1; d >>= 1) maxm = max(maxm, __shfl_xor_sync(0xFFFFFFFFu, maxm, d));
            if (a.ops) {
                uint32_t word = 0;
                int have = -1;
                for (int x = 0; x < maxm; x++) {
                    if (x < mym) {
                        const int p = mym - 1 - x;
                        if ((p >> 4) != have) { have = p >> 4; word = tb[(size_t)have * 32]; }
                        const uint32_t vp = (uint32_t)(n + x) + omis;             /* byte slot counted from the aligned base */
                        oacc |= ((word >> (2 * (p & 15))) & 3u) << (8 * (vp & 3u));
                        if ((vp & 3u) == 3u) {
                            const long pos = (long)(vp & ~3u) - (long)omis;     /* offset of the word inside the row */
                            if (pos >= 0 && (size_t)pos + 4 <= a.ops_stride) *reinterpret_cast<uint32_t *>(ops + pos) = oacc;
                            else put_bytes(ops, pos, oacc, 0, 4, a.ops_stride);
                            oacc = 0;
                        }
                    }
                }
            }
            if (!done) {
                n += m;
                i0 += ai;
                j0 += aj;
                if (last) done = true;
            }
            __syncwarp();
        }
        if (ops && (((uint32_t)n + omis) & 3u)) {                  /* the last, partial word */
            const uint32_t vp = (uint32_t)n + omis;
            put_bytes(ops, (long)(vp & ~3u) - (long)omis, oacc, 0, (int)(vp & 3u), a.ops_stride);
        }
        if (!skipped && read < a.n) {
            if (a.res) {                                        /* PART 3, alnmain.c:458-477 */
                as_result *o = a.res + read;
                o->ed = score;
                o->n_cigar_op = n;
                if (score == -1) { o->valid = 0; o->flag = 4; o->mapq = 0; }
            }
            if (a.n_ops) a.n_ops[read] = n;
            if (a.score) a.score[read] = score;
        }
    }
#pragma unroll
    for (int d = 16; d >= 1; d >>= 1) st_cells += __shfl_xor_sync(0xFFFFFFFFu, st_cells, d);
    if (lane == 0 && st_cells) atomicAdd(a.stats + 5, st_cells);
}

template <int W>
static int launch_bp(as_gpu *h, const GactArgs &a, size_t n, int blocks_per_sm) {
    const size_t planes_per_warp = (size_t)a.TS * W * 32;                 /* uint2 */
    const size_t tb_per_warp = (size_t)((a.TS + a.TQ + 15) / 16 + 1) * 32;  /* u32 */
    size_t warps_needed = (n + 31) / 32;
    size_t blocks = (warps_needed + BP_THREADS / 32 - 1) / (BP_THREADS / 32);
    if (blocks_per_sm <= 0) {        /* as many resident blocks as the register file allows: the traceback is latency bound */
        int occ = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_gact_bp<W>, BP_THREADS, 0) != cudaSuccess || occ < 1) occ = 4;
        blocks_per_sm = occ;
    }
    size_t cap = (size_t)h->sm_count * blocks_per_sm;
    if (blocks > cap) blocks = cap;
    const size_t warps = blocks * (BP_THREADS / 32);
    AS_TRY(as_scratch(h, &h->gact_planes, warps * planes_per_warp * sizeof(uint2)));
    AS_TRY(as_scratch(h, &h->gact_tb, warps * tb_per_warp * sizeof(uint32_t)));
    k_gact_bp<W><<<(unsigned)blocks, BP_THREADS, 0, h->stream>>>(a, (uint2 *)h->gact_planes.p, planes_per_warp,
                                                                  (uint32_t *)h->gact_tb.p, tb_per_warp);
    h->launches++;
    AS_CUDA(cudaGetLastError());
    return AS_OK;
}

int as_launch_gact_bp(as_gpu *h, GactArgs &a, size_t n, uint32_t max_qlen, uint32_t max_tlen) {
    (void)max_tlen;
    if (n >= 0xFFFFFFE0ull) return AS_ERANGE;
    const int rows = (int)(max_qlen < (uint32_t)a.T ? max_qlen : (uint32_t)a.T);
    const int words = rows > 0 ? (rows + 31) / 32 : 1;
    int bps = 0;
    if (const char *e = getenv("AS_GACT_BLOCKS_PER_SM")) {
        int v = atoi(e);
        if (v >= 1 && v <= 16) bps = v;
    }
    if (words <= 2) return launch_bp<2>(h, a, n, bps);
    if (words <= 4) return launch_bp<4>(h, a, n, bps);
    if (words <= 5) return launch_bp<5>(h, a, n, bps);
    if (words <= 8) return launch_bp<8>(h, a, n, bps);
    if (words <= 10) return launch_bp<10>(h, a, n, bps);
    if (words <= 12) return launch_bp<12>(h, a, n, bps);
    return launch_bp<16>(h, a, n, bps);
}
