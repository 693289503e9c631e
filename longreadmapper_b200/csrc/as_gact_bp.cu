/*
 * as_gact_bp.cu -- bit-parallel GACT: the aligner behind cigar_align (mutils.c:94-105 -> simple_gact, the
 * un-vendored lisanhu/GACT; spec: DESIGN.md "GACT spec", oracle: oracle/oracle.c orc_gact).
 *
 * Same tiles, same DP matrix, same traceback priority (diag, up, left) as the warp-per-read kernel in
 * as_align.cu -- but the tile is filled with the Myers/Hyyro bit-vector recurrence (the one edlib's
 * calculateBlock uses, edlib.cpp:373-408): 32 DP rows per 32-bit word, ~24 integer instructions per word instead of
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
 *
 * Sliding band (first attempt at a tile).  A cell of value v has every optimal path inside |i-j| <= v, so a tile
 * whose traceback costs at most KB is fully determined by the cells with |i-j| <= KB.  The first attempt therefore
 * keeps only NB words per column in registers -- the words that meet rows [j-KB, j+KB], KB = 16(NB-1) -- and slides
 * them down one word every 32 columns (the range is the same for all 32 lanes of the warp, because they sit on the
 * same column).  A word entering the band starts from Pv = ~0, Mv = 0 and the top word takes hin = +1, which is what
 * edlib's block band does (edlib.cpp:694-895): values computed that way are never below the true ones and are exact
 * wherever the true value is <= KB.  Hence: if the walk from the end cell stays in |i-j| <= KB and costs <= KB, every
 * value and every tie-break on it equals the full matrix's, and the ops are bit-identical; otherwise the warp fills
 * the same tile again with the whole matrix (lanes that had passed recompute what they had).  For 150-bp reads at
 * NB = 2 that is 2 of 5 words per column filled and stored.
 *
 * Traceback reads the planes through a per-lane window in shared memory: BP_WIN columns down the diagonal are fetched
 * with independent loads in one go, then walked from shared memory, so the chain of dependent global loads is one
 * round trip per window instead of one per op.  The committed ops leave as whole 32-bit words: four 2-bit codes of
 * the walk are turned into four op bytes, in forward order, by one lookup.
 */
#include "as_common.cuh"
#include <stdlib.h>
#include <cub/cub.cuh>

#define BP_THREADS 128
#define BP_WIN 16                     /* columns of traceback planes staged in shared memory per refill */

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

/* bit planes of query rows [32w, 32w+32) of the tile that starts at base i0 (rows at and past Tq read as 'A') */
__device__ __forceinline__ void load_qword(const uint32_t *qp, int i0, int w, int Tq, uint32_t &lo, uint32_t &hi) {
    uint64_t x = 0;
    if (32 * w < Tq) x = as_bases16(qp, (uint64_t)(i0 + 32 * w));
    if (32 * w + 16 < Tq) x |= (uint64_t)as_bases16(qp, (uint64_t)(i0 + 32 * w + 16)) << 32;
    lo = compress_even(x);
    hi = compress_even(x >> 1);
}

struct TileGeom {
    const uint32_t *qp;
    const uint32_t *tbase;
    uint64_t toff;
    int i0, j0, Tq, Tt, maxTt, T, inner, KB;
    bool done, q_done, t_done, last, track_row, track_col;
};
struct TileOut {
    int m, ai, aj, tscore, walkcost;
    bool bad;
};

/*
 * One attempt at one tile: fill (NBW words per column; SLIDE: the words slide down the diagonal, else they are the
 * whole tile), end cell, traceback into the warp's slab `tb` (16 ops per word, walk order).
 */
#define BP_FW 16                      /* diagonal steps looked at per round of the run-skipping walk */
/*
 * Run-skipping traceback for tiles that are their read's only one (every op is committed, the walk runs from the far corner
 * to the origin).  Reads at a few percent error are mostly diagonal runs of '=': a round fetches the plane words of the
 * next BP_FW cells down the diagonal with independent loads, keeps one bit of each (op != '='), and one FFS gives the
 * length of the run; only the cell that ends a run is decoded.  The ops are not materialised: a '=' column is the zero the
 * row already holds (the rows are cleared before the launch), every other op leaves as an event (walk position, code) in
 * the warp's slab, and the caller writes those few bytes once the column count is known.  Same cells, same codes, same
 * order as the cell-by-cell walk: the planes decide, not the walker.
 */
template <int NBW, bool SLIDE>
__device__ __forceinline__ void fast_walk(const TileGeom &g, const uint2 *__restrict__ planes, uint32_t *__restrict__ tb,
                                          int KB, TileOut &out, int &n_events) {
    int ci = g.done ? 0 : g.Tq, cj = g.done ? 0 : g.Tt;
    int pos = 0, ne = 0, walkcost = 0;
    bool oob = false;
    while (__any_sync(0xFFFFFFFFu, ci > 0 && cj > 0)) {
        if (SLIDE && ci > 0 && cj > 0 && abs(ci - cj) > KB) {   /* left the band: planes there were never written */
            oob = true;
            ci = cj = 0;
        }
        const int K = (ci > 0 && cj > 0) ? min(min(ci, cj), BP_FW) : 0;
        uint32_t mask = 0;
#pragma unroll
        for (int k = 0; k < BP_FW; k++) {
            if (k < K) {
                const int col = cj - 1 - k, row = ci - 1 - k;
                const int cwlo = SLIDE ? (col - KB > 0 ? (col - KB) >> 5 : 0) : 0;      /* band of column j = col+1 */
                int sl = (row >> 5) - cwlo;
                if (SLIDE) sl = min(max(sl, 0), NBW - 1);
                const uint2 v = planes[((size_t)col * NBW + sl) * 32];
                mask |= (((v.x | v.y) >> (row & 31)) & 1u) << k;
            }
        }
        mask |= 1u << K;                                         /* sentinel: the run ends with the round at the latest */
        const int run = __ffs((int)mask) - 1;
        pos += run;
        ci -= run;
        cj -= run;
        if (run < K) {                                           /* the cell that ended the run: its code (never '=') */
            const int col = cj - 1, row = ci - 1;
            const int cwlo = SLIDE ? (col - KB > 0 ? (col - KB) >> 5 : 0) : 0;
            int sl = (row >> 5) - cwlo;
            if (SLIDE) sl = min(max(sl, 0), NBW - 1);
            const uint2 v = planes[((size_t)col * NBW + sl) * 32];
            const uint32_t op = ((v.x >> (row & 31)) & 1u) | (((v.y >> (row & 31)) & 1u) << 1);
            tb[(size_t)ne * 32] = ((uint32_t)pos << 2) | op;
            ne++;
            pos++;
            walkcost++;
            ci -= (op != 2u);
            cj -= (op != 1u);
        }
    }
    /* the rest of a border: row 0 is all 'D' (left), column 0 all 'I' (up) */
    while (__any_sync(0xFFFFFFFFu, (ci | cj) != 0)) {
        if ((ci | cj) != 0) {
            const uint32_t op = ci == 0 ? 2u : 1u;
            tb[(size_t)ne * 32] = ((uint32_t)pos << 2) | op;
            ne++;
            pos++;
            walkcost++;
            ci -= (op == 1u);
            cj -= (op == 2u);
        }
    }
    __syncwarp();
    n_events = ne;
    out.m = pos;
    out.ai = g.Tq;
    out.aj = g.Tt;
    out.tscore = walkcost;
    out.walkcost = walkcost;
    out.bad = SLIDE && !g.done && (oob || walkcost > KB);
}

#define BP_DC 16                      /* one-word diagonal band: rows [j - BP_DC, j - BP_DC + 31] of column j */
/*
 * One-word DIAGONAL band for tiles that are their read's only one (first attempt of reads at a few percent error).  The
 * 32-row window of column j starts at row top(j) = max(1, j - BP_DC): it stands still for the first BP_DC + 1 columns (its
 * top neighbour is then the matrix's own row 0, whose horizontal delta is +1 exactly) and moves down ONE ROW per column
 * afterwards -- Pv, Mv shift right by a bit, the row that enters at the bottom starts from Pv = 1 (+1: an upper bound), the
 * row above the window is taken as +1 per column too, and the column step itself is the unchanged Myers/Hyyro step.  So
 * the argument of the word-sliding band applies bit for bit: computed values are never below the true ones and are exact
 * wherever the true value is <= KB = BP_DC - 1, every cell with |i-j| <= KB is in the window, and a walk that stays inside
 * |i-j| <= KB at a cost <= KB reads only exact values and exact tie-breaks.  Otherwise the tile goes to the wider bands.
 * Half the words of the two-word band, and bit b of a word is the DIAGONAL i - j + BP_DC of the sliding part, which is what
 * the run-skipping walk follows.
 *
 * Planes: only the 16 diagonals around the main one are KEPT -- bits [s(j), s(j) + 16) of Pv' (low half) and of ~D0 (high
 * half), s(j) = clamp(j - top(j) - 8, 0, 16), one 32-bit word per column and lane ([column][lane]): 19 KB per warp for
 * 150-bp reads, so the slabs of all resident warps (80 MB) stay in the 126 MB L2.  The op codes themselves are not formed in
 * the fill: the walk finds the '=' runs from the sequences and reads a word only at a mismatch, where (Pv', ~D0) is the op.
 * A walk that needs a diagonal outside the kept 16 (more than seven net indels) is flagged like one that leaves the band.
 */
__device__ __forceinline__ int diag1_top(int j) { return j - BP_DC > 1 ? j - BP_DC : 1; }
__device__ __forceinline__ int diag1_s(int j) {               /* first kept bit of column j's word */
    const int s = j - diag1_top(j) - 8;
    return s < 0 ? 0 : (s > 16 ? 16 : s);
}
__device__ __forceinline__ void tile_diag1(const TileGeom &g, uint2 *__restrict__ planes2, uint32_t *__restrict__ tb,
                                           unsigned long long &st_eval, TileOut &out, int &n_events) {
    /* [column][lane] of u32 in the warp's slab: planes2 = slab + lane in uint2 units = slab + 2 * lane in u32 units */
    uint32_t *planes = reinterpret_cast<uint32_t *>(planes2) - (threadIdx.x & 31u);
    const int Tq = g.Tq, Tt = g.Tt;
    constexpr int KB = BP_DC - 1;
    uint32_t qlo_a, qhi_a, qlo_b, qhi_b;                     /* query planes of words w0, w0 + 1 */
    load_qword(g.qp, g.i0, 0, Tq, qlo_a, qhi_a);
    load_qword(g.qp, g.i0, 1, Tq, qlo_b, qhi_b);
    uint32_t Pv = 0xFFFFFFFFu, Mv = 0u, tw = 0;
    /* one column: the Myers/Hyyro step on the window (ql, qh = the query planes of its 32 rows), the kept 16 diagonals of
     * Pv' and ~D0 stored as one word.  (x << 1) | 1 and x << 1 are written as multiply-adds: they go to the FMA
     * pipe, which idles while the ALU pipe carries the logic ops. */
#define DIAG1_STEP(QL, QH, SS, PDST) DIAG1_STEP_C(QL, QH, SS, PDST, 0u - (tw & 1u), 0u - ((tw >> 1) & 1u), tw >>= 2)
    /* column c of a target word that stays in its register (the unrolled chunks): both masks are a shift left (FMA pipe, as a
     * multiply) and an arithmetic shift right, the word is not consumed */
#define DIAG1_STEP_U(QL, QH, SS, PDST, C)                                                              \
    DIAG1_STEP_C(QL, QH, SS, PDST, (uint32_t)((int32_t)(tw * (1u << (31 - 2 * (C)))) >> 31),           \
                 (uint32_t)((int32_t)(tw * (1u << (30 - 2 * (C)))) >> 31), (void)0)
#define DIAG1_STEP_C(QL, QH, SS, PDST, TL, TH, NEXT)                                                   \
    do {                                                                                               \
        const uint32_t tl_ = (TL), th_ = (TH);                                                         \
        NEXT;                                                                                          \
        const uint32_t eq_ = ~(((QL) ^ tl_) | ((QH) ^ th_));                                          \
        const uint32_t xv_ = eq_ | Mv;                                                                 \
        const uint32_t xh_ = (((eq_ & Pv) + Pv) ^ Pv) | eq_;                                           \
        const uint32_t ph_ = (Mv | ~(xh_ | Pv)) * 2u + 1u;   /* the row above the window: +1 */       \
        const uint32_t mh_ = (Pv & xh_) * 2u;                                                          \
        const uint32_t nd0_ = ~(xh_ | Mv);                   /* ~D0 */                                 \
        const uint32_t npv_ = mh_ | ~(xv_ | ph_);                                                      \
        Mv = ph_ & xv_;                                                                                \
        Pv = npv_;                                                                                     \
        if (act) *(PDST) = (SS) == 8u ? __byte_perm(npv_, nd0_, 0x6521) : __byte_perm(npv_ >> (SS), nd0_ >> (SS), 0x5410); \
    } while (0)
    /* ---- columns 1 .. 16: the window stands on rows 1 .. 32 (column 17 too: it opens the first 16-column chunk below).
     * Every lane of the warp has them (done lanes ride along into their own slab words): fully unrolled, the kept
     * diagonals' shift is an immediate ---- */
    if (__all_sync(0xFFFFFFFFu, g.done || Tt >= 16)) {
        const bool act = true;
        if (!g.done) tw = as_bases16(g.tbase, g.toff + (uint64_t)g.j0);
#pragma unroll
        for (int c = 0; c < 16; c++) DIAG1_STEP_U(qlo_a, qhi_a, (uint32_t)(c > 8 ? c - 8 : 0), planes + 32 * c, c);
    } else {
        const int jA = g.maxTt < 16 ? g.maxTt : 16;
        for (int j = 1; j <= jA; j++) {
            const bool act = j <= Tt;
            if (act && j == 1) tw = as_bases16(g.tbase, g.toff + (uint64_t)g.j0);
            const uint32_t ss = (uint32_t)(j - 9 > 0 ? j - 9 : 0);        /* diag1_s(j) for top(j) = 1 */
            uint32_t pv0 = Pv, mv0 = Mv;
            DIAG1_STEP(qlo_a, qhi_a, ss, planes + (size_t)(j - 1) * 32);
            if (!act) { Pv = pv0; Mv = mv0; }
        }
    }
    /* ---- columns 17 ..: 16 columns per target word.  Column 17 still stands on rows 1 .. 32 (its kept bits start at 8, like
     * every later column's); from column 18 on the window moves down one row per column; the query words change every 32
     * columns, on a target-word boundary (column 49, 81, ...) ---- */
    for (int j0 = BP_DC + 1; j0 <= g.maxTt; j0 += 16) {
        const int jend = j0 + 15 < g.maxTt ? j0 + 15 : g.maxTt;              /* last column of this target word */
        const bool first = j0 == BP_DC + 1;
        if (j0 <= Tt) tw = as_bases16(g.tbase, g.toff + (uint64_t)(g.j0 + j0 - 1));
        uint32_t sh = (uint32_t)(j0 - (BP_DC + 1)) & 31u;              /* window start inside query word a, this column */
        if (sh == 0 && !first) {                                       /* crossed into the next query word */
            const int w0 = (j0 - (BP_DC + 1)) >> 5;
            qlo_a = qlo_b;
            qhi_a = qhi_b;
            load_qword(g.qp, g.i0, w0 + 1, Tq, qlo_b, qhi_b);
        }
        uint32_t *pc = planes + (size_t)(j0 - 1) * 32;
        /* a whole target word of 16 columns that every lane of the warp still needs (all but the last chunk of a warp of
         * equally long reads): fully unrolled, the window's position inside the query words is a constant (0 or 16 plus the
         * column), no per-column length test, store offsets are immediates */
        if (jend - j0 == 15 && __all_sync(0xFFFFFFFFu, g.done || jend <= Tt)) {
            const bool act = true;
            if (sh == 0) {
#pragma unroll
                for (int c = 0; c < 16; c++) {
                    if (c || !first) {
                        Pv = __funnelshift_r(Pv, 1u, 1u);
                        Mv >>= 1;
                    }
                    const uint32_t ql = c ? __funnelshift_r(qlo_a, qlo_b, c) : qlo_a, qh = c ? __funnelshift_r(qhi_a, qhi_b, c) : qhi_a;
                    DIAG1_STEP_U(ql, qh, 8u, pc + 32 * c, c);
                }
            } else {
#pragma unroll
                for (int c = 0; c < 16; c++) {
                    Pv = __funnelshift_r(Pv, 1u, 1u);
                    Mv >>= 1;
                    const uint32_t ql = __funnelshift_r(qlo_a, qlo_b, 16 + c), qh = __funnelshift_r(qhi_a, qhi_b, 16 + c);
                    DIAG1_STEP_U(ql, qh, 8u, pc + 32 * c, c);
                }
            }
            continue;
        }
#pragma unroll 4
        for (int j = j0; j <= jend; j++, sh++, pc += 32) {
            const bool act = j <= Tt;
            if (act) {
                if (j > BP_DC + 1) {
                    Pv = __funnelshift_r(Pv, 1u, 1u);       /* (Pv >> 1) | 0x80000000: the entering row starts at +1 */
                    Mv >>= 1;
                }
                const uint32_t ql = __funnelshift_r(qlo_a, qlo_b, sh), qh = __funnelshift_r(qhi_a, qhi_b, sh);
                DIAG1_STEP(ql, qh, 8u, pc);
            }
        }
    }
#undef DIAG1_STEP
#undef DIAG1_STEP_U
#undef DIAG1_STEP_C
    __syncwarp();
    if (!g.done) st_eval += 32ull * (unsigned long long)Tt;

    /* ---- run-skipping walk.  A cell's op is '=' exactly where its two bases are equal (both op-code planes are masked by
     * ~Eq: a match always takes the diagonal at no cost), so the runs of '=' down a diagonal are the common suffixes of
     * q[0, ci) and t[0, cj): a round XORs the 16 packed bases that end at row ci with the 16 that end at column cj and one
     * CLZ gives the run.  The planes are read only at the cell that ends a run -- a mismatch, whose op the DP decides:
     * 'X' where D = D[i-1][j-1] + 1 (~D0), else 'I' (up) where Pv', else 'D' (left).  Cell (ci, cj) is bit b of column cj's
     * kept word, b = (ci - cj) + 8 from column 17 on (the window slides with the diagonal), b = ci - 1 - max(0, cj - 9)
     * before; the low half holds Pv', the high half ~D0. */
    int ci = g.done ? 0 : Tq, cj = g.done ? 0 : Tt;
    int pos = 0, ne = 0, walkcost = 0;
    bool oob = false;
    /* (this pass only sees tiles that start at the origin: i0 = j0 = 0) */
    const int tlow = g.toff < 16 ? (int)g.toff : 16;         /* target bases in front of the tile, as far as a round can reach */
    while (__any_sync(0xFFFFFFFFu, ci > 0 && cj > 0)) {
        if (ci > 0 && cj > 0 && abs(ci - cj) > 7) {          /* outside the 16 kept diagonals (a fortiori outside the band) */
            oob = true;
            ci = cj = 0;
        }
        if (ci > 0 && cj > 0) {
            const int K = min(min(ci, cj), 16);
            /* the 16 bases that end at row ci / column cj, the last one in the top two bits (a start before the first word
             * of the sequence is loaded from base 0 and shifted up; cells past K are cut off by the sentinel) */
            const int nq = max(16 - ci, 0), nt = max(16 - cj - tlow, 0);        /* bases missing in front of the sequence */
            const uint32_t q16 = as_bases16(g.qp, (uint64_t)(uint32_t)(ci + nq - 16)) << (2 * nq);
            const uint32_t t16 = as_bases16(g.tbase, g.toff + (uint64_t)(uint32_t)(cj + tlow + nt - 16) - (uint64_t)(uint32_t)tlow) << (2 * nt);
            const uint32_t x = q16 ^ t16;
            uint32_t m = (x | (x >> 1)) & 0x55555555u;       /* bit 30 - 2k: cell k of the round (k = 0: (ci, cj)) mismatches */
            if (K < 16) m |= 1u << (30 - 2 * K);             /* sentinel: the run ends with the round at the latest */
            const int run = __clz((int)m) >> 1;
            pos += run;
            ci -= run;
            cj -= run;
            if (run < K) {                                   /* the mismatch that ended the run */
                const int bb = cj >= 17 ? ci - cj + 8 : ci - 1 - max(0, cj - 9);
                const uint32_t u = planes[(size_t)(cj - 1) * 32] >> bb;
                const uint32_t op = (u & 0x10000u) ? 3u : ((u & 1u) ? 1u : 2u);
                tb[(size_t)ne * 32] = ((uint32_t)pos << 2) | op;
                ne++;
                pos++;
                walkcost++;
                ci -= (op != 2u);
                cj -= (op != 1u);
            }
        }
    }
    while (__any_sync(0xFFFFFFFFu, (ci | cj) != 0)) {
        if ((ci | cj) != 0) {
            const uint32_t op = ci == 0 ? 2u : 1u;
            tb[(size_t)ne * 32] = ((uint32_t)pos << 2) | op;
            ne++;
            pos++;
            walkcost++;
            ci -= (op == 1u);
            cj -= (op == 2u);
        }
    }
    __syncwarp();
    n_events = ne;
    out.m = pos;
    out.ai = Tq;
    out.aj = Tt;
    out.tscore = walkcost;
    out.walkcost = walkcost;
    out.bad = !g.done && (oob || walkcost > KB);
}

template <int NBW, bool SLIDE, bool TRACK, bool FAST = false>
__device__ __forceinline__ void tile_pass(const TileGeom &g, uint2 *__restrict__ planes, uint32_t *__restrict__ tb,
                                          uint2 (*win)[32], uint32_t lane, unsigned long long &st_eval, TileOut &out,
                                          int *n_events = nullptr) {
    const int Tq = g.Tq, Tt = g.Tt, KB = SLIDE ? g.KB : 0x3FFFFFFF;
    uint32_t qlo[NBW], qhi[NBW], Pv[NBW], Mv[NBW];
#pragma unroll
    for (int s = 0; s < NBW; s++) {
        load_qword(g.qp, g.i0, s, Tq, qlo[s], qhi[s]);
        Pv[s] = 0xFFFFFFFFu;                    /* D[i][0] = i */
        Mv[s] = 0u;
    }
    const int wlast = Tq > 0 ? (Tq - 1) >> 5 : 0;
    const uint32_t blast = Tq > 0 ? (uint32_t)(Tq - 1) & 31u : 0u;
    int sc = Tq;                                            /* D[Tq][j], valid while the bottom row is in the band */
    int rowkey = (Tq << 12) | (Tt << 1);                    /* bottom-row candidates (edge 0), column 0 first */
    /* sliding band: sb = value at the bottom row of the band's lowest word, in the column just filled */
    /* TRACK = some lane of the warp needs edge values (a tile that is not its read's last); reads of one tile never do,
     * and their instantiation carries none of the bookkeeping */
    const bool track_row = TRACK && g.track_row;
    const bool track_sb = TRACK && SLIDE && (g.track_row || g.track_col);
    int sb = 32 * NBW, wlo = 0;

    /* ---- fill: one column per iteration ---- */
    uint32_t tw = 0;
    uint2 *pcol = planes;
    for (int j = 1; j <= g.maxTt; j++, pcol += NBW * 32) {
        if (SLIDE) {
            const int nwlo = j - KB > 1 ? (j - KB - 1) >> 5 : 0;
            if (nwlo != wlo) {                               /* warp-uniform: the band moves down one word */
                wlo = nwlo;
#pragma unroll
                for (int s = 0; s + 1 < NBW; s++) {
                    Pv[s] = Pv[s + 1]; Mv[s] = Mv[s + 1]; qlo[s] = qlo[s + 1]; qhi[s] = qhi[s + 1];
                }
                Pv[NBW - 1] = 0xFFFFFFFFu;                   /* the entering word's column j-1 is +1 per row */
                Mv[NBW - 1] = 0u;
                load_qword(g.qp, g.i0, wlo + NBW - 1, Tq, qlo[NBW - 1], qhi[NBW - 1]);
                if (track_sb && j <= Tt) {
                    sb += 32;
                    if (wlo + NBW - 1 == wlast) sc = sb - 31 + (int)blast;
                }
            }
        }
        if (j <= Tt) {
            if (((j - 1) & 15) == 0) tw = as_bases16(g.tbase, g.toff + (uint64_t)(g.j0 + j - 1));
            const uint32_t tl = 0u - (tw & 1u), th = 0u - ((tw >> 1) & 1u);
            tw >>= 2;
            uint32_t hpos = 1u, hneg = 0u;                   /* D[0][j] - D[0][j-1] = +1 (also above a band) */
#pragma unroll
            for (int s = 0; s < NBW; s++) {
                const uint32_t eq = ~((qlo[s] ^ tl) | (qhi[s] ^ th));
                const uint32_t pv = Pv[s], mv = Mv[s];
                const uint32_t xv = eq | mv;
                const uint32_t eq2 = eq | hneg;
                const uint32_t xh = (((eq2 & pv) + pv) ^ pv) | eq2;
                uint32_t ph = mv | ~(xh | pv);
                uint32_t mh = pv & xh;
                if (track_row) {
                    if (wlo + s == wlast && Tq > 0) sc += (int)((ph >> blast) & 1u) - (int)((mh >> blast) & 1u);
                }
                const uint32_t hp_out = ph >> 31, hn_out = mh >> 31;
                ph = (ph << 1) | hpos;
                mh = (mh << 1) | hneg;
                const uint32_t npv = mh | ~(xv | ph);
                Pv[s] = npv;
                Mv[s] = ph & xv;
                const uint32_t nd0 = ~(xh | mv);            /* ~D0 */
                uint2 o;
                o.x = ~eq & (npv | nd0);
                o.y = ~eq & (~npv | nd0);
                pcol[s * 32] = o;
                hpos = hp_out;
                hneg = hn_out;
            }
            if (track_sb) sb += (int)hpos - (int)hneg;
            if (track_row && wlast >= wlo && wlast < wlo + NBW) rowkey = min(rowkey, (sc << 12) | ((Tt - j) << 1));
        }
    }
    __syncwarp();
    if (!g.done) st_eval += (unsigned long long)(32 * NBW) * (unsigned long long)Tt;
    if (FAST) {                                 /* every lane's tile is its read's only one */
        fast_walk<NBW, SLIDE>(g, planes, tb, KB, out, *n_events);
        return;
    }

    /* ---- end cell (non-last tiles): min (D, distance from the corner, edge) over the interior edges ---- */
    int ei = Tq, ej = Tt;
    if (!g.done && !g.last) {
        int best = 0x7FFFFFFF;
        if (!g.q_done) best = rowkey;
        if (!g.t_done) {                                     /* Tt == maxTt here, so wlo is this lane's last band */
            int v = Tt;                                      /* D[0][Tt] */
            if (SLIDE) {                                     /* value just above the band of the last column */
                v = sb;
#pragma unroll
                for (int s = 0; s < NBW; s++) v -= __popc(Pv[s]) - __popc(Mv[s]);
            }
            int ck = wlo == 0 ? ((v << 12) | (Tq << 1) | 1) : 0x7FFFFFFF;
#pragma unroll
            for (int s = 0; s < NBW; s++) {
                const int r0 = 32 * (wlo + s);
                if (r0 < Tq) {
                    const uint32_t pv = Pv[s], mv = Mv[s];
                    const int lim = Tq - r0 < 32 ? Tq - r0 : 32;
                    for (int b = 0; b < lim; b++) {
                        v += (int)((pv >> b) & 1u) - (int)((mv >> b) & 1u);
                        ck = min(ck, (v << 12) | ((Tq - (r0 + b + 1)) << 1) | 1);
                    }
                }
            }
            best = min(best, ck);
        }
        const int dist = (best >> 1) & 0x7FF;
        if (best & 1) { ei = Tq - dist; ej = Tt; } else { ei = Tq; ej = Tt - dist; }
    }

    /* ---- traceback: ops of the walk, packed 16 per word, into the warp's slab ---- */
    int ci = g.done ? 0 : ei, cj = g.done ? 0 : ej;
    bool anchored = false, oob = false;
    if (SLIDE && (ci < 0 || cj < 0)) {                      /* no edge cell inside the band */
        oob = true;
        ci = cj = 0;
    }
    uint32_t acc = 0;
    int m = 0, ai = ei, aj = ej, tscore = 0, walkcost = 0;
    int wc0 = -1, wi0 = 0;               /* window: entry k holds column wc0-k at the word of row wi0-1-k */
    while (__any_sync(0xFFFFFFFFu, (ci | cj) != 0)) {
        if (SLIDE && (ci | cj) != 0 && abs(ci - cj) > KB) {  /* left the band: planes there were never written */
            oob = true;
            ci = cj = 0;
        }
        const bool interior = ci > 0 && cj > 0;
        int k = wc0 - (cj - 1);
        const bool hit = interior && (unsigned)k < (unsigned)BP_WIN && ((wi0 - 1 - k) >> 5) == ((ci - 1) >> 5);
        if (__any_sync(0xFFFFFFFFu, interior && !hit)) {
            /* refill: BP_WIN independent loads down the diagonal from the current cell; the lines of the window
             * after this one are pulled into L2 meanwhile (lanes of a warp mostly share them) */
            if (interior) {
                wc0 = cj - 1;
                wi0 = ci;
#pragma unroll
                for (int kk = 0; kk < BP_WIN; kk++) {
                    const int col = wc0 - kk, row = wi0 - 1 - kk;
                    if (col >= 0 && row >= 0) {
                        const int cwlo = SLIDE ? (col - KB > 0 ? (col - KB) >> 5 : 0) : 0;    /* band of column j = col+1 */
                        int s = (row >> 5) - cwlo;
                        if (SLIDE) s = min(max(s, 0), NBW - 1);                               /* only off-band entries clamp */
                        win[kk][lane] = planes[((size_t)col * NBW + s) * 32];
                    }
                }
                if (wc0 >= 2 * BP_WIN && wi0 > 2 * BP_WIN) {
                    const int col = wc0 - BP_WIN - BP_WIN / 2, row = wi0 - 1 - BP_WIN - BP_WIN / 2;
                    const int cwlo = SLIDE ? (col - KB > 0 ? (col - KB) >> 5 : 0) : 0;
                    const int s = SLIDE ? min(max((row >> 5) - cwlo, 0), NBW - 1) : (row >> 5);
                    const uint2 *pf = planes + ((size_t)col * NBW + s) * 32;
#pragma unroll
                    for (int kk = -BP_WIN / 2; kk < BP_WIN / 2; kk += 2)
                        asm volatile("prefetch.global.L2 [%0];" ::"l"(pf + (ptrdiff_t)kk * NBW * 32));
                }
                k = 0;
            }
        }
        if ((ci | cj) != 0) {
            uint32_t op;
            if (ci == 0) op = 2;
            else if (cj == 0) op = 1;
            else {
                const int b = (ci - 1) & 31;
                const uint2 v = win[k][lane];
                op = ((v.x >> b) & 1u) | (((v.y >> b) & 1u) << 1);
            }
            const int si = ci - (op != 2u), sj = cj - (op != 1u);
            walkcost += (op != 0u);
            if (g.last || (si < g.inner && sj < g.inner)) {
                if (!anchored) { ai = ci; aj = cj; anchored = true; }
                acc |= op << (2 * (m & 15));
                m++;
                if ((m & 15) == 0) { tb[(size_t)((m >> 4) - 1) * 32] = acc; acc = 0; }
                tscore += (op != 0u);
            }
            ci = si;
            cj = sj;
        }
    }
    if (m & 15) tb[(size_t)(m >> 4) * 32] = acc;
    __syncwarp();
    out.m = m; out.ai = ai; out.aj = aj; out.tscore = tscore; out.walkcost = walkcost;
    out.bad = SLIDE && !g.done && (oob || walkcost > KB);
}

/*
 * W = words of a whole tile column (full-matrix attempt); NB = words of the sliding band of the first attempt
 * (NB == W: no band, every tile is filled once with the whole matrix).
 */
template <int W, int NB>
__global__ void __launch_bounds__(BP_THREADS) k_gact_bp(GactArgs a, uint2 *__restrict__ planes_all, size_t planes_per_warp,
                                                        uint32_t *__restrict__ tb_all, size_t tb_per_warp) {
    constexpr bool BANDED = NB < W;
    __shared__ uint2 s_win[BP_THREADS / 32][BP_WIN][32];
    __shared__ uint32_t s_lut[256];      /* 4 walk-order op codes (one byte of the slab) -> 4 op bytes in forward order */
    /* second launch over a retry queue that the warp-per-read kernel takes (short queue): nothing to do, leave at once */
    if (a.retry_mode == 2 && a.n_dev && *a.n_dev <= AS_GACT_RETRY_WARP_MAX) return;
    for (uint32_t e = threadIdx.x; e < 256; e += BP_THREADS)
        s_lut[e] = ((e >> 6) & 3u) | (((e >> 4) & 3u) << 8) | (((e >> 2) & 3u) << 16) | ((e & 3u) << 24);
    __syncthreads();
    const uint32_t lane = threadIdx.x & 31u;
    const size_t gwarp = ((size_t)blockIdx.x * BP_THREADS + threadIdx.x) >> 5;
    uint2 *planes = planes_all + gwarp * planes_per_warp + lane;          /* [(col*words + slot)*32] */
    uint32_t *tb = tb_all + gwarp * tb_per_warp + lane;                    /* [word*32]: 16 ops per word, walk order */
    uint2 (*win)[32] = s_win[threadIdx.x >> 5];
    unsigned long long st_cells = 0, st_eval = 0;
    unsigned int st_retry = 0, st_tiles = 0, st_wtiles = 0, st_half = 0, st_diag = 0, st_diag_retry = 0;
    TileGeom g;
    g.T = a.T;
    g.inner = a.T - a.O;
    g.KB = BANDED ? 16 * (NB - 1) : 256;
    g.tbase = a.tbase;

    for (;;) {
        uint32_t base = 0;
        if (lane == 0) base = atomicAdd(a.work_counter, 32u);
        base = __shfl_sync(0xFFFFFFFFu, base, 0);
        const size_t n_slots = a.n_dev ? (size_t)*a.n_dev : a.n;
        if (base >= n_slots || (a.retry_mode == 2 && n_slots <= AS_GACT_RETRY_WARP_MAX)) break;
        const size_t slot = (size_t)base + lane;
        bool done = slot >= n_slots;
        const size_t read = done ? 0 : (a.order ? (size_t)a.order[slot] : slot);
        int qlen = 0, tlen = 0;
        g.toff = 0;
        if (!done) {
            qlen = (int)a.qlens[read];
            tlen = a.tlens ? (int)a.tlens[read] : qlen;
            g.toff = a.toff ? a.toff[read] : 0;
            /* seq_lookup failed: flag 4 was set by PART 2; reference behaviour is undefined (alnmain.c:409,430-446) */
            if (g.toff == AS_TOFF_NONE) done = true;
        }
        bool skipped = done;
        g.qp = a.qpack + (done ? 0 : read * a.q_wstride);
        uint8_t *ops = (a.ops && !done) ? a.ops + read * a.ops_stride : nullptr;
        int n = 0, score = 0;
        g.i0 = 0;
        g.j0 = 0;
        /* column ops leave as aligned 32-bit words: omis = misalignment of the row, oacc = the word being filled */
        const uint32_t omis = (uint32_t)(reinterpret_cast<uintptr_t>(ops) & 3u);
        uint32_t oacc = 0;
        bool by_events = false;          /* the read's ops were placed as events: no partial word is pending */
        int n_runs_out = -1;             /* emit_runs: run bytes written | AS_NOPS_RUNS */
        if (!done && qlen == 0 && tlen == 0) done = true;

        while (__any_sync(0xFFFFFFFFu, !done)) {
            /* ---- tile geometry of this lane ---- */
            g.Tq = 0;
            g.Tt = 0;
            g.q_done = g.t_done = true;
            if (!done) {
                g.Tq = qlen - g.i0 < g.T ? qlen - g.i0 : g.T;
                g.Tt = tlen - g.j0 < g.T ? tlen - g.j0 : g.T;
                g.q_done = (g.i0 + g.Tq == qlen);
                g.t_done = (g.j0 + g.Tt == tlen);
            }
            g.done = done;
            g.last = g.q_done && g.t_done;
            g.maxTt = __reduce_max_sync(0xFFFFFFFFu, g.Tt);
            /* the bottom row only matters to tiles that leave query rows behind (never for reads of one tile) */
            g.track_row = __any_sync(0xFFFFFFFFu, !done && !g.q_done);
            g.track_col = __any_sync(0xFFFFFFFFu, !done && !g.t_done);

            TileOut o;
            bool full = !BANDED;
            /* reads of one tile (all 32 of the warp): the run-skipping walk, ops left as events */
            /* (compiled into the short-read kernels only: in the wide ones the extra code costs registers -- 96 -> 124 and
             * 128 -> 157 per thread -- and their tiles are rarely a read's only one) */
            const bool fast = W <= 5 && a.ops_zeroed && __all_sync(0xFFFFFFFFu, done || (g.last && g.i0 == 0 && g.j0 == 0));
            int n_events = 0;
            if (W <= 5 && fast) {
                bool wider = true;
                if (BANDED && NB == 2) {                     /* lowest band level: the one-word diagonal band first */
                    tile_diag1(g, planes, tb, st_eval, o, n_events);
                    if (a.retry_list) {
                        /* the reads this band cannot prove exact leave for the second launch (wider bands); the warp goes on */
                        if (!done && o.bad) {
                            a.retry_list[atomicAdd(a.retry_count, 1u)] = (uint32_t)read;
                            done = true;
                            skipped = true;
                            g.done = true;
                            o.m = 0;
                            o.tscore = 0;
                            n_events = 0;
                        }
                        wider = false;
                    } else {
                        wider = __any_sync(0xFFFFFFFFu, o.bad);
                    }
                    st_diag += (lane == 0);
                    st_diag_retry += (wider && lane == 0);
                }
                if (BANDED && wider) {
                    tile_pass<NB, true, false, true>(g, planes, tb, win, lane, st_eval, o, &n_events);
                    full = __any_sync(0xFFFFFFFFu, o.bad);
                    st_retry += (full && lane == 0);
                }
                if (full) tile_pass<W, false, false, true>(g, planes, tb, win, lane, st_eval, o, &n_events);
            } else {
            if (BANDED) {
                /* the lean instantiation pays off for short reads only: in the wide kernels a third copy of the tile code
                 * costs more registers (occupancy) than the bookkeeping it saves */
                if (W > 5 || g.track_row || g.track_col) tile_pass<NB, true, true>(g, planes, tb, win, lane, st_eval, o);
                else tile_pass<NB, true, (W > 5)>(g, planes, tb, win, lane, st_eval, o);
                /* some lane's walk cannot be proven exact: the warp fills the tile again with the whole matrix */
                full = __any_sync(0xFFFFFFFFu, o.bad);
                st_retry += (full && lane == 0);
            }
            if (full) tile_pass<W, false, true>(g, planes, tb, win, lane, st_eval, o);
            }
            if (!done) {
                st_cells += (unsigned long long)g.Tq * (unsigned long long)g.Tt;
                st_tiles++;
                st_half += (2 * o.walkcost <= g.KB);
            }
            st_wtiles += (lane == 0);
            score += o.tscore;
            /* ---- emit the committed ops in forward order ---- */
            int mym = 0;
            if (!done) {
                if (o.m == 0 && !g.last) { score = -1; done = true; }
                else mym = o.m;
            }
            if (fast) by_events = true;
            if (W <= 5 && a.ops && fast && a.emit_runs) {
                /* the caller wants run bytes (k_rle_ops' format: (len << 2) | class, class 0 M for '=' and 'X', 1 I, 2 D, runs
                 * of more than 63 continue in the next byte): the events ARE the run boundaries, so the lane writes the few
                 * bytes of its read directly -- no cleared row, no op bytes, no second pass over them.  Events are in walk
                 * order (last column first): taken from the back they run forward along the row. */
                if (mym > 0) {
                    int out = 0, cur = 0, e = n_events - 1;
                    auto emit = [&](int len, uint32_t cls) {
                        while (len > 0) {
                            const int l = len > 63 ? 63 : len;
                            if ((size_t)out < a.ops_stride) ops[out] = (uint8_t)(((uint32_t)l << 2) | cls);
                            out++;
                            len -= l;
                        }
                    };
                    while (e >= 0) {
                        const uint32_t wd = tb[(size_t)e * 32];
                        const uint32_t op = wd & 3u;
                        if (op == 3u) { e--; continue; }                       /* 'X' is part of an M run */
                        const int col = mym - 1 - (int)(wd >> 2);
                        emit(col - cur, 0u);
                        int len = 1;
                        e--;
                        while (e >= 0) {                                        /* the indel goes on while the class and the columns do */
                            const uint32_t w2 = tb[(size_t)e * 32];
                            if ((w2 & 3u) != op || mym - 1 - (int)(w2 >> 2) != col + len) break;
                            len++;
                            e--;
                        }
                        emit(len, op);
                        cur = col + len;
                    }
                    emit(mym - cur, 0u);
                    n_runs_out = out | AS_NOPS_RUNS;
                }
            } else if (W <= 5 && a.ops && fast) {
                /* the row holds '=' everywhere already: place the few other ops, walk position p -> column mym-1-p */
                const int maxe = __reduce_max_sync(0xFFFFFFFFu, mym > 0 ? n_events : 0);
                for (int e = 0; e < maxe; e++) {
                    if (e < n_events && mym > 0) {
                        const uint32_t wd = tb[(size_t)e * 32];
                        const size_t col = (size_t)(mym - 1) - (size_t)(wd >> 2);
                        if (col < a.ops_stride) ops[col] = (uint8_t)(wd & 3u);
                    }
                }
            } else if (a.ops) {
                /* one op: walk index p = mym-1-x of the slab -> byte slot n+x of the row */
                auto put_one = [&](int x) {
                    const int p = mym - 1 - x;
                    const uint32_t wd = tb[(size_t)(p >> 4) * 32];
                    const uint32_t vp = (uint32_t)(n + x) + omis;              /* byte slot counted from the aligned base */
                    oacc |= ((wd >> (2 * (p & 15))) & 3u) << (8 * (vp & 3u));
                    if ((vp & 3u) == 3u) {
                        const long pos = (long)(vp & ~3u) - (long)omis;       /* offset of the word inside the row */
                        if (pos >= 0 && (size_t)pos + 4 <= a.ops_stride) *reinterpret_cast<uint32_t *>(ops + pos) = oacc;
                        else put_bytes(ops, pos, oacc, 0, 4, a.ops_stride);
                        oacc = 0;
                    }
                };
                int x = 0;
                while (x < mym && ((((uint32_t)(n + x) + omis) & 3u) != 0u)) { put_one(x); x++; }     /* up to 3 ops */
                /* whole words: 4 ops = 8 bits of the slab (walk order) -> 4 bytes (forward order) by one lookup */
                const int nb4 = (mym - x) >> 2;
                const int maxnb = __reduce_max_sync(0xFFFFFFFFu, nb4);
                const int wtop = mym > 0 ? (mym - 1) >> 4 : 0;
                uint32_t lo = 0, hi = 0;
                int have = -2;
                for (int it = 0; it < maxnb; it++) {
                    if (it < nb4) {
                        const int x0 = x + 4 * it;
                        const int q = mym - 4 - x0;                             /* lowest walk index of the four */
                        const int wq = q >> 4;
                        if (wq != have) {
                            hi = (have == wq + 1) ? lo : (wq + 1 <= wtop ? tb[(size_t)(wq + 1) * 32] : 0u);
                            lo = tb[(size_t)wq * 32];
                            have = wq;
                        }
                        const uint32_t e = __funnelshift_r(lo, hi, 2u * ((uint32_t)q & 15u)) & 0xFFu;
                        const uint32_t word = s_lut[e];
                        const long pos = (long)((uint32_t)(n + x0) + omis) - (long)omis;
                        if ((size_t)pos + 4 <= a.ops_stride) *reinterpret_cast<uint32_t *>(ops + pos) = word;
                        else put_bytes(ops, pos, word, 0, 4, a.ops_stride);
                    }
                }
                x += 4 * nb4;
                while (x < mym) { put_one(x); x++; }                            /* up to 3 ops, left in oacc */
            }
            if (!done) {
                n += o.m;
                g.i0 += o.ai;
                g.j0 += o.aj;
                if (g.last) done = true;
            }
            __syncwarp();
        }
        if (ops && !by_events && (((uint32_t)n + omis) & 3u)) {    /* the last, partial word */
            const uint32_t vp = (uint32_t)n + omis;
            put_bytes(ops, (long)(vp & ~3u) - (long)omis, oacc, 0, (int)(vp & 3u), a.ops_stride);
        }
        if (!skipped && read < a.n) {
            if (a.res) {                                        /* PART 3, alnmain.c:458-477 */
                as_result *r = a.res + read;
                r->ed = score;
                r->n_cigar_op = n_runs_out >= 0 ? n_runs_out : n;
                if (score == -1) { r->valid = 0; r->flag = 4; r->mapq = 0; }
            }
            if (a.n_ops) a.n_ops[read] = n;
            if (a.score) a.score[read] = score;
        }
    }
#pragma unroll
    for (int d = 16; d >= 1; d >>= 1) {
        st_cells += __shfl_xor_sync(0xFFFFFFFFu, st_cells, d);
        st_eval += __shfl_xor_sync(0xFFFFFFFFu, st_eval, d);
        st_tiles += __shfl_xor_sync(0xFFFFFFFFu, st_tiles, d);
        st_half += __shfl_xor_sync(0xFFFFFFFFu, st_half, d);
    }
    if (lane == 0) {
        if (st_cells) atomicAdd(a.stats + 5, st_cells);
        if (st_eval) atomicAdd(a.stats + 10, st_eval);
        if (st_retry) atomicAdd(a.stats + 11, (unsigned long long)st_retry);
        atomicAdd(a.fail_count + 0, st_retry);                  /* policy counters of this launch */
        atomicAdd(a.fail_count + 1, st_wtiles);
        atomicAdd(a.fail_count + 2, st_half);
        atomicAdd(a.fail_count + 3, st_tiles);
    }
}

/* Adaptive band.  cnt: warp-tiles filled twice, warp-tiles, lane-tiles half the band would serve, lane-tiles.
 * Widen when more than a quarter of the warp-tiles were filled twice; narrow when at most one lane-tile in 256 would
 * fail at half the band (about one warp-tile in eight refilled).  The verdict goes to pinned host memory, which the
 * launcher reads when it picks the next launch's instantiation; results never depend on it. */
__global__ void k_gact_policy(volatile unsigned int *host_band, const unsigned int *cnt, unsigned int cur,
                              const unsigned int *queued, unsigned int n_reads, unsigned long long *stats) {
    if (queued && *queued) atomicAdd(stats + 17, (unsigned long long)*queued);        /* as_stats.gact_queued */
    if (!host_band) return;                                                            /* fixed band: only the counter */
    const unsigned long long retry = cnt[0], wtiles = cnt[1], half = cnt[2], tiles = cnt[3];
    if (wtiles == 0) return;
    unsigned int kb = cur;
    /* one-tile reads at the lowest level queue their failures instead of refilling (retry stays 0): widen when more than
     * a quarter of the READS had to be queued -- the one-word pass is then mostly wasted */
    if (queued && 4ull * *queued > n_reads) kb = cur * 2;
    else if (4ull * retry > wtiles) kb = cur >= 128 ? 256 : cur * 2;
    else if (cur > 16 && 256ull * (tiles - half) <= tiles) kb = cur / 2;
    *host_band = kb;
}

template <int W, int NB>
static int launch_bp(as_gpu *h, GactArgs &a, size_t n, int blocks_per_sm, unsigned int cur_band) {
    const size_t planes_per_warp = (size_t)a.TS * W * 32;                 /* uint2; the band's slab is the smaller one */
    /* u32 per lane: 16 ops per word of the cell-by-cell walk; one word per EVENT of the run-skipping walk of one-tile reads,
     * which can be every op of the walk (two unrelated sequences) */
    const size_t tb_per_warp = (size_t)(W <= 5 && a.ops_zeroed ? a.TS + a.TQ + 2 : (a.TS + a.TQ + 15) / 16 + 2) * 32;
    size_t warps_needed = (n + 31) / 32;
    size_t blocks = (warps_needed + BP_THREADS / 32 - 1) / (BP_THREADS / 32);
    if (blocks_per_sm <= 0) {        /* reads are claimed from a work counter: the grid is the resident blocks */
        int occ = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_gact_bp<W, NB>, BP_THREADS, 0) != cudaSuccess || occ < 1) occ = 4;
        blocks_per_sm = occ;
    }
    size_t cap = (size_t)h->sm_count * blocks_per_sm;
    if (blocks > cap) blocks = cap;
    const size_t warps = blocks * (BP_THREADS / 32);
    AS_TRY(as_scratch(h, &h->gact_planes, warps * planes_per_warp * sizeof(uint2)));
    AS_TRY(as_scratch(h, &h->gact_tb, warps * tb_per_warp * sizeof(uint32_t)));
    /* counters (zeroed by as_launch_gact): [0] work counter, [1..4] policy counters of this launch */
    unsigned int *cnt = a.work_counter;
    a.fail_count = cnt + 1;
    k_gact_bp<W, NB><<<(unsigned)blocks, BP_THREADS, 0, h->stream>>>(a, (uint2 *)h->gact_planes.p, planes_per_warp,
                                                                      (uint32_t *)h->gact_tb.p, tb_per_warp);
    h->launches++;
    AS_CUDA(cudaGetLastError());
    if (a.band == 0 || a.retry_count) {
        k_gact_policy<<<1, 1, 0, h->stream>>>(a.band == 0 ? h->d_band : nullptr, cnt + 1, cur_band, a.retry_count, (unsigned int)n, a.stats);
        h->launches++;
        AS_CUDA(cudaGetLastError());
    }
    return AS_OK;
}

template <int W>
static int launch_bp_w(as_gpu *h, GactArgs &a, size_t n, int bps, unsigned int kb) {
    /* band half-width -> words of the sliding band, KB = 16 (NB - 1); a band as tall as the tile is no band */
    if (kb <= 16 && 2 < W) return launch_bp<W, 2>(h, a, n, bps, 16);
    if (kb <= 32 && 3 < W) return launch_bp<W, 3>(h, a, n, bps, 32);
    if (kb <= 64 && 5 < W) return launch_bp<W, 5>(h, a, n, bps, 64);
    if (kb <= 128 && 9 < W) return launch_bp<W, 9>(h, a, n, bps, 128);
    return launch_bp<W, W>(h, a, n, bps, 256);
}

/*
 * Reads of several tiles: a warp walks its 32 reads tile by tile until the longest is through, so lanes with short
 * reads idle (1-50 kbp reads: the mean is half the maximum).  Handing out the reads by decreasing work (query + target
 * bases; unmapped reads last) gives a warp 32 reads of nearly one length and puts the longest at the front of the
 * work queue.  One radix sort of n (key, index) pairs per launch; results do not depend on the order.
 */
__global__ void __launch_bounds__(256) k_gact_order_keys(GactArgs a, uint32_t maxw, uint32_t *__restrict__ keys,
                                                         uint32_t *__restrict__ idx) {
    const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= a.n) return;
    const uint32_t q = a.qlens[i], t = a.tlens ? a.tlens[i] : q;
    const bool skip = a.toff && a.toff[i] == AS_TOFF_NONE;
    const uint32_t w = q + t < maxw ? q + t : maxw;
    keys[i] = skip ? maxw + 1u : maxw - w;
    idx[i] = (uint32_t)i;
}

static int gact_order(as_gpu *h, GactArgs &a, size_t n, uint32_t max_qlen, uint32_t max_tlen) {
    const uint32_t maxw = max_qlen + max_tlen;
    int bits = 1;
    while (bits < 32 && ((maxw + 1u) >> bits)) bits++;
    size_t temp = 0;
    const size_t n4 = (n + 63) & ~(size_t)63;
    AS_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, temp, (const uint32_t *)nullptr, (uint32_t *)nullptr,
                                            (const uint32_t *)nullptr, (uint32_t *)nullptr, (int)n, 0, bits, h->stream));
    AS_TRY(as_scratch(h, &h->gact_order, 4 * n4 * sizeof(uint32_t) + temp + 256));
    uint32_t *k0 = (uint32_t *)h->gact_order.p, *i0 = k0 + n4, *k1 = i0 + n4, *i1 = k1 + n4;
    k_gact_order_keys<<<(unsigned)((n + 255) / 256), 256, 0, h->stream>>>(a, maxw, k0, i0);
    AS_CUDA(cudaGetLastError());
    AS_CUDA(cub::DeviceRadixSort::SortPairs((void *)(i1 + n4), temp, (const uint32_t *)k0, k1, (const uint32_t *)i0, i1, (int)n, 0,
                                            bits, h->stream));
    h->launches += 4;
    a.order = i1;
    return AS_OK;
}

/* as_map_batch sizes the order buffer for the whole batch before it enqueues chunks (growing a scratch buffer waits) */
int as_gact_order_reserve(as_gpu *h, size_t n) {
    return as_scratch(h, &h->gact_order, 4 * ((n + 63) & ~(size_t)63) * sizeof(uint32_t) + (n * 16 + (1u << 20)));
}

int as_launch_gact_bp(as_gpu *h, GactArgs &a, size_t n, uint32_t max_qlen, uint32_t max_tlen) {
    if (n >= 0x7FFFFFE0ull) return AS_ERANGE;
    {
        bool order = max_qlen > (uint32_t)a.T || max_tlen > (uint32_t)a.T;
        if (a.order_mode >= 0) order = a.order_mode != 0;
        if (order) AS_TRY(gact_order(h, a, n, max_qlen, max_tlen));
    }
    const int rows = (int)(max_qlen < (uint32_t)a.T ? max_qlen : (uint32_t)a.T);
    const int words = rows > 0 ? (rows + 31) / 32 : 1;
    const int bps = a.blocks_per_sm >= 1 && a.blocks_per_sm <= 16 ? a.blocks_per_sm : 0;
    if (!h->h_band) {          /* adaptive band: written by k_gact_policy, read here (mapped pinned memory) */
        AS_CUDA(cudaHostAlloc((void **)&h->h_band, sizeof(unsigned int), cudaHostAllocMapped));
        *h->h_band = AS_GACT_BAND0;
        AS_CUDA(cudaHostGetDevicePointer((void **)&h->d_band, (void *)h->h_band, 0));
    }
    unsigned int kb = a.band < 0 ? 0xFFFFu : a.band > 0 ? (unsigned int)a.band : *(volatile unsigned int *)h->h_band;
    if (kb == 0) kb = AS_GACT_BAND0;
    if (words <= 5 && kb <= 16 && a.ops_zeroed && max_qlen <= (uint32_t)a.T && max_tlen <= (uint32_t)a.T) {
        /* every read is one tile and the first attempt is the one-word diagonal band: reads it cannot prove exact (a few
         * per cent on real genomes: chimeric or mis-placed reads, diverged repeat copies) are queued and aligned by a second
         * launch over the queue with the 32-wide band / the whole matrix.  Before, one such read made its 32-read warp fill
         * the tile twice more, and a per-cent of them pushed the adaptive band up for everybody (c2r: GACT 0.75 -> 2.26 ms) */
        AS_TRY(as_scratch(h, &h->gact_retry, n * sizeof(uint32_t)));
        unsigned int *cnt = a.work_counter;                       /* 16 words, zeroed by as_launch_gact */
        a.retry_list = (uint32_t *)h->gact_retry.p;
        a.retry_count = cnt + 5;
        AS_TRY(launch_bp_w<5>(h, a, n, bps, kb));
        GactArgs b = a;
        b.retry_list = nullptr;
        b.retry_count = nullptr;
        b.order = (const uint32_t *)h->gact_retry.p;
        b.n_dev = cnt + 5;
        b.work_counter = cnt + 6;                                 /* its policy counters: cnt + 7 .. 10, not read */
        b.band = 32;
        /* a short queue (uniform genomes: a few reads per million) goes to the warp-per-read kernel: 32 lanes on a read
         * finish it in microseconds, where one lane of a bit-parallel warp took 0.18 ms for the same handful -- paid once per
         * chunk of the host-buffer path.  A long queue (repeat-rich genomes) keeps the bit-parallel kernel. */
        b.retry_mode = 2;
        AS_TRY(launch_bp_w<5>(h, b, n, bps, 32));
        GactArgs c = b;
        c.retry_mode = 1;
        c.work_counter = cnt + 11;
        return as_launch_gact_warp_list(h, c, max_qlen);
    }
    if (words <= 5) return launch_bp_w<5>(h, a, n, bps, kb);
    if (words <= 10) return launch_bp_w<10>(h, a, n, bps, kb);
    return launch_bp_w<16>(h, a, n, bps, kb);
}
