/*
 * as_align.cu -- PART 2 and PART 3 of the per-read loop and the aligner behind cigar_align:
 *   locus/strand   seq_lookup alnmain.c:151-176, _rev_comp_in_place alnmain.c:27-60, alnmain.c:425-438
 *   GACT           cigar_align mutils.c:94-105 -> simple_gact (un-vendored lisanhu/GACT; spec: DESIGN.md)
 *   result fill    alnmain.c:458-477
 *
 * GACT kernel: one warp per read, tiles of the read processed in sequence (each tile's anchor is where the
 * previous traceback stopped).  Inside a tile lane L owns R = T/32 consecutive query rows and sweeps the
 * target columns on a skewed anti-diagonal schedule (lane L is at column s-L in step s): the R cell values
 * of the previous column live in registers, the strip boundary moves between lanes with one __shfl_up_sync
 * per step, and 2-bit traceback codes (the column's op codes, packed R per word) are staged in shared
 * memory, laid out [lane][column] so that a step's 32 stores hit 32 different banks.
 */
#include "as_common.cuh"

/* ---------------------------------------------------------------------------------------------- */
/* PART 2: locus lookup + strand fix                                                               */
/* ---------------------------------------------------------------------------------------------- */
__device__ __forceinline__ uint32_t comp_ascii(uint32_t c) {             /* alnmain.c:31-51 */
    /* branch-free (a switch made the 32 lanes of a warp take five different paths per byte): the 2-bit code of the byte
     * picks the letter it would have to be and its complement out of two constants; anything else is 'N' */
    const uint32_t code = ((c >> 1) ^ (c >> 2)) & 3u;                    /* A 0, C 1, G 2, T 3 (as_pack4) */
    const uint32_t sel = code | 0x4440u;                                 /* byte `code` of the constant, zeros above it */
    const uint32_t is = __byte_perm(0x54474341u, 0u, sel);               /* "ACGT"[code] */
    const uint32_t comp = __byte_perm(0x41434754u, 0u, sel);             /* "TGCA"[code] */
    return (c & 0xDFu) == is ? comp : (uint32_t)'N';
}

/* reverse the sixteen 2-bit groups of a word */
__device__ __forceinline__ uint32_t rev2(uint32_t x) {
    const uint32_t r = __brev(x);
    return ((r >> 1) & 0x55555555u) | ((r & 0x55555555u) << 1);
}

/*
 * Thread per read for the lookup (vote -> mta binary search -> result), then the warp turns to the strand-1 reads among
 * its 32, one at a time: ASCII reverse complement by mirrored pairs (both halves coalesced), and the packed copy straight
 * from the packed words (reverse the 2-bit groups, complement) unless the read holds a byte outside ACGT, whose code the
 * packed form cannot complement -- those re-pack from the ASCII, as do reads of more than 512 bases.
 */
__global__ void __launch_bounds__(256) k_locus(DevIndex ix, char *__restrict__ ascii, size_t stride,
                                               const uint32_t *__restrict__ lens, size_t n, uint32_t *__restrict__ pack,
                                               size_t wstride, const VoteOut *__restrict__ vote,
                                               as_result *__restrict__ res, uint64_t *__restrict__ toff) {
    const uint32_t lane = threadIdx.x & 31u;
    const size_t warp0 = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const size_t nwarps = ((size_t)gridDim.x * blockDim.x) >> 5;
    for (size_t base = warp0 * 32; base < n; base += nwarps * 32) {
        const size_t r = base + lane;
        uint32_t len = 0;
        int strand = 0;
        if (r < n) {
            const VoteOut v = vote[r];
            len = lens[r];
            const uint64_t loc = v.best_key;
            /* seq_lookup: the entries tile .cat, so only the last one starting at or before loc can match */
            int lo = 0, hi = ix.mta_len;     /* first entry with offset > loc */
            while (lo < hi) {
                int mid = (lo + hi) >> 1;
                if (ix.mta[mid].offset <= loc) lo = mid + 1; else hi = mid;
            }
            int seq_id = -1;
            uint64_t off = 0, fwd = 0;
            if (lo > 0 && v.phases >= 0 && loc + len >= loc) {           /* u64 wrap of loc+len: reference UB, unmapped here */
                const int i = lo - 1;
                const uint64_t start = ix.mta[i].offset, sl = ix.mta[i].seq_len, end = start + 2 * sl;
                if (loc >= start && loc + len <= start + sl) {           /* alnmain.c:158-164 */
                    seq_id = i; strand = 0; fwd = loc; off = loc - start;
                } else if (loc >= start + sl && loc + len <= end) {      /* alnmain.c:165-172 */
                    seq_id = i; strand = 1; off = end - loc - len; fwd = off + start;
                }
            }
            as_result o;
            o.loc = loc;
            o.off = off;
            o.seq_id = seq_id;
            o.strand = strand;
            o.flag = seq_id < 0 ? 4 : (strand ? 16 : 0);                 /* alnmain.c:466-474; GACT may still fail it */
            o.mapq = seq_id < 0 ? 0 : 255;
            o.ed = -1;
            o.valid = seq_id >= 0;
            o.n_cigar_op = 0;
            o.myers = -2;
            o.phases = v.phases;
            o.from_outer = v.from_outer;
            res[r] = o;
            toff[r] = seq_id < 0 ? AS_TOFF_NONE : fwd;
        }
        /* ---- strand-1 reads (alnmain.c:433-438) ---- */
        if (!ascii && wstride <= 16) {
            /* packed-only batch of short reads (as_map_batch_packed): every lane turns its own read -- all its words are
             * requested before any is written, so the 32 reads' loads are in flight together instead of one read per round
             * trip.  Out word w = complement of input bases [len-16w-16, len-16w), last first. */
            uint32_t *pw = pack + r * wstride;
            const uint32_t nw = strand == 1 ? (len + 15) >> 4 : 0u;
            uint32_t y[16];
#pragma unroll
            for (int w = 0; w < 16; w++) {
                y[w] = 0;
                if ((uint32_t)w < nw) {
                    const int st = (int)len - 16 * w - 16;
                    const uint32_t cnt = st >= 0 ? 16u : (uint32_t)(16 + st);
                    const uint32_t x = as_bases16(pw, (uint64_t)(st > 0 ? st : 0));
                    y[w] = ~rev2(x) >> (2u * (16u - cnt));
                    if (cnt < 16u) y[w] &= (1u << (2u * cnt)) - 1u;
                }
            }
#pragma unroll
            for (int w = 0; w < 16; w++)
                if ((uint32_t)w < nw) pw[w] = y[w];
            continue;
        }
        /* ... of this warp's 32, four at a time: a group of 8 lanes turns one read -- ASCII reverse complement by mirrored
         * pairs, then the packed copy straight from the packed words (all of a read's words are loaded before any is stored:
         * an output word reads input words from the other end of the read).  A round is a chain of dependent round trips to
         * memory (load, complement, store, re-pack), so four reads per round is four times fewer of them. */
        uint32_t todo = __ballot_sync(0xFFFFFFFFu, strand == 1);
        const uint32_t grp = lane >> 3, sub = lane & 7u, gmask = 0xFFu << (8u * grp);
        while (todo) {
            /* this group's read of the round: the (grp+1)-th set bit of todo, if there is one */
            uint32_t t = todo;
            for (uint32_t k = 0; k < grp; k++) t &= t - 1;
            const bool mine = t != 0;
            const int src = mine ? __ffs((int)t) - 1 : 0;
            const uint32_t rl = __shfl_sync(0xFFFFFFFFu, len, src);
            const size_t rr = base + (size_t)src;
            uint8_t *s = (uint8_t *)ascii + rr * stride;
            uint32_t *pw = pack + rr * wstride;
            const uint32_t half = (ascii && mine) ? rl >> 1 : 0u;          /* packed-only batch (as_map_batch_packed): no ASCII copy */
            bool other = false;                                          /* a byte outside ACGT/acgt */
            for (uint32_t i = sub; i < half; i += 8) {
                const uint32_t ca = comp_ascii(s[i]), cb = comp_ascii(s[rl - 1 - i]);
                other |= (ca == 'N') | (cb == 'N');
                s[i] = (uint8_t)cb;
                s[rl - 1 - i] = (uint8_t)ca;
            }
            if (ascii && mine && (rl & 1u) && sub == 0) {
                const uint32_t cm = comp_ascii(s[rl >> 1]);
                other |= cm == 'N';
                s[rl >> 1] = (uint8_t)cm;
            }
            const uint32_t nw = mine ? (rl + 15) >> 4 : 0u;
            const bool repack = (__ballot_sync(0xFFFFFFFFu, other) & gmask) != 0 || nw > 32;
            /* out word w = complement of input bases [rl-16w-16, rl-16w), last first */
            uint32_t y[4];
#pragma unroll
            for (int q = 0; q < 4; q++) {
                const uint32_t w = sub + 8u * q;
                y[q] = 0;
                if (!repack && w < nw) {
                    const int st = (int)rl - 16 * (int)w - 16;
                    const uint32_t cnt = st >= 0 ? 16u : (uint32_t)(16 + st);
                    const uint32_t x = as_bases16(pw, (uint64_t)(st > 0 ? st : 0));
                    y[q] = ~rev2(x) >> (2u * (16u - cnt));
                    if (cnt < 16u) y[q] &= (1u << (2u * cnt)) - 1u;
                }
            }
            __syncwarp();
            if (!repack) {
#pragma unroll
                for (int q = 0; q < 4; q++) {
                    const uint32_t w = sub + 8u * q;
                    if (w < nw) pw[w] = y[q];
                }
            } else {
                for (uint32_t w = sub; w < nw; w += 8) {
                    const uint32_t b0 = w << 4, lim = rl - b0 < 16 ? rl - b0 : 16;
                    pw[w] = as_pack16(s + b0, lim);
                }
            }
            __syncwarp();
            /* retire the (up to) four reads of this round */
            for (int k = 0; k < 4 && todo; k++) todo &= todo - 1;
        }
    }
}

/* long reads: one warp per read, so that the byte work of a 10-kbp reverse complement spreads over all warps */
__global__ void __launch_bounds__(256) k_locus_long(DevIndex ix, char *__restrict__ ascii, size_t stride,
                                               const uint32_t *__restrict__ lens, size_t n, uint32_t *__restrict__ pack,
                                               size_t wstride, const VoteOut *__restrict__ vote,
                                               as_result *__restrict__ res, uint64_t *__restrict__ toff,
                                               uint32_t *__restrict__ tmp_rows) {
    const uint32_t lane = threadIdx.x & 31u;
    const uint64_t warp0 = (blockIdx.x * (uint64_t)blockDim.x + threadIdx.x) >> 5;
    const uint64_t nwarps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
    for (uint64_t r = warp0; r < n; r += nwarps) {
        const VoteOut v = vote[r];
        const uint32_t len = lens[r];
        const uint64_t loc = v.best_key;
        /* seq_lookup: the entries tile .cat, so only the last one starting at or before loc can match */
        int lo = 0, hi = ix.mta_len;     /* first entry with offset > loc */
        while (lo < hi) {
            int mid = (lo + hi) >> 1;
            if (ix.mta[mid].offset <= loc) lo = mid + 1; else hi = mid;
        }
        int seq_id = -1, strand = 0;
        uint64_t off = 0, fwd = 0;
        if (lo > 0 && v.phases >= 0 && loc + len >= loc) {               /* u64 wrap of loc+len: reference UB, unmapped here */
            const int i = lo - 1;
            const uint64_t start = ix.mta[i].offset, sl = ix.mta[i].seq_len, end = start + 2 * sl;
            if (loc >= start && loc + len <= start + sl) {               /* alnmain.c:158-164 */
                seq_id = i; strand = 0; fwd = loc; off = loc - start;
            } else if (loc >= start + sl && loc + len <= end) {          /* alnmain.c:165-172 */
                seq_id = i; strand = 1; off = end - loc - len; fwd = off + start;
            }
        }
        if (lane == 0) {
            as_result o;
            o.loc = loc;
            o.off = off;
            o.seq_id = seq_id;
            o.strand = strand;
            o.flag = seq_id < 0 ? 4 : (strand ? 16 : 0);                 /* alnmain.c:466-474; GACT may still fail it */
            o.mapq = seq_id < 0 ? 0 : 255;
            o.ed = -1;
            o.valid = seq_id >= 0;
            o.n_cigar_op = 0;
            o.myers = -2;
            o.phases = v.phases;
            o.from_outer = v.from_outer;
            res[r] = o;
            toff[r] = seq_id < 0 ? AS_TOFF_NONE : fwd;
        }
        if (strand == 1 && !ascii) {
            /* packed-only batch: out word w = complement of input bases [len-16w-16, len-16w), last first -- built in this
             * warp's scratch row (an output word reads input words from the other end of the read), then copied back */
            uint32_t *pw = pack + r * wstride;
            uint32_t *tw = tmp_rows + warp0 * wstride;
            const uint32_t nw = (len + 15) >> 4;
            for (uint32_t w = lane; w < nw; w += 32) {
                const long st = (long)len - 16 * (long)w - 16;
                const uint32_t cnt = st >= 0 ? 16u : (uint32_t)(16 + st);
                const uint32_t x = as_bases16(pw, (uint64_t)(st > 0 ? st : 0));
                uint32_t y = ~rev2(x) >> (2u * (16u - cnt));
                if (cnt < 16u) y &= (1u << (2u * cnt)) - 1u;
                tw[w] = y;
            }
            __syncwarp();
            for (uint32_t w = lane; w < nw; w += 32) pw[w] = tw[w];
            __syncwarp();
        } else if (strand == 1) {                                        /* alnmain.c:433-438 */
            uint8_t *s = (uint8_t *)ascii + r * stride;
            const uint32_t half = len >> 1;
            for (uint32_t i = lane; i < half; i += 32) {
                uint32_t a = s[i], b = s[len - 1 - i];
                s[i] = (uint8_t)comp_ascii(b);
                s[len - 1 - i] = (uint8_t)comp_ascii(a);
            }
            if ((len & 1u) && lane == 0) s[half] = (uint8_t)comp_ascii(s[half]);
            __syncwarp();
            uint32_t *pw = pack + r * wstride;
            const uint32_t nw = (len + 15) >> 4;
            for (uint32_t w = lane; w < nw; w += 32) {
                const uint32_t b0 = w << 4, lim = len - b0 < 16 ? len - b0 : 16;
                pw[w] = as_pack16(s + b0, lim);
            }
        }
    }
}

int as_launch_locus(as_gpu *h, char *d_ascii, size_t stride, const uint32_t *d_lens, size_t n,
                    uint32_t *d_pack, size_t wstride, const VoteOut *d_vote, as_result *d_res, uint64_t *d_toff) {
    if (n == 0) return AS_OK;
    if (wstride > 32 + 3) {                              /* reads of more than 512 bases in the batch */
        uint64_t blocks = (n + 7) / 8, cap = (uint64_t)h->sm_count * 8;
        const unsigned grid = (unsigned)(blocks < cap ? blocks : cap);
        uint32_t *tmp_rows = nullptr;
        if (!d_ascii) {                                  /* one scratch row per warp for the packed reverse complement */
            AS_TRY(as_scratch(h, &h->locus_tmp, (size_t)grid * 8 * wstride * sizeof(uint32_t)));
            tmp_rows = (uint32_t *)h->locus_tmp.p;
        }
        k_locus_long<<<grid, 256, 0, h->stream>>>(h->ix, d_ascii, stride, d_lens, n, d_pack, wstride, d_vote, d_res, d_toff,
                                                  tmp_rows);
        h->launches++;
        AS_CUDA(cudaGetLastError());
        return AS_OK;
    }
    uint64_t blocks = (n + 255) / 256, cap = (uint64_t)h->sm_count * 8;
    k_locus<<<(unsigned)(blocks < cap ? blocks : cap), 256, 0, h->stream>>>(h->ix, d_ascii, stride, d_lens, n, d_pack, wstride,
                                                                          d_vote, d_res, d_toff);
    h->launches++;
    AS_CUDA(cudaGetLastError());
    return AS_OK;
}

/* ---------------------------------------------------------------------------------------------- */
/* GACT                                                                                            */
/* ---------------------------------------------------------------------------------------------- */
template <int R, typename PT>
__global__ void __launch_bounds__(32) k_gact(GactArgs a) {
    extern __shared__ __align__(16) unsigned char smem[];
    const int T = a.T, inner = a.T - a.O, TS = a.TS;
    PT *ptr = reinterpret_cast<PT *>(smem);                    /* [32][TS] */
    uint8_t *ts = smem + (size_t)32 * TS * sizeof(PT);         /* [TS] target codes of the tile */
    uint8_t *tmp = ts + TS;                                    /* [TS+TQ] committed ops, reversed */
    const int L = threadIdx.x;
    PT *myptr = ptr + (size_t)L * TS;
    unsigned long long st_cells = 0;

    for (;;) {
        uint32_t read = 0;
        if (L == 0) read = atomicAdd(a.work_counter, 1u);
        read = __shfl_sync(0xFFFFFFFFu, read, 0);
        if (a.n_dev) {                                           /* list mode: the reads queued by k_gact_bp */
            const uint32_t cnt = *a.n_dev;
            if (read >= cnt || (a.retry_mode == 1 && cnt > AS_GACT_RETRY_WARP_MAX)) break;
            read = a.order[read];
        }
        if (read >= a.n) break;
        const int qlen = (int)a.qlens[read];
        const int tlen = a.tlens ? (int)a.tlens[read] : qlen;
        const uint64_t toff = a.toff ? a.toff[read] : 0;
        uint8_t *ops = a.ops ? a.ops + (size_t)read * a.ops_stride : nullptr;
        int n = 0, score = 0;
        if (toff == AS_TOFF_NONE) {
            /* seq_lookup failed: flag 4 was set by PART 2; reference behaviour is undefined (alnmain.c:409,430-446) */
            continue;
        }
        const uint32_t *qp = a.qpack + (size_t)read * a.q_wstride;
        int i0 = 0, j0 = 0;
        while (i0 < qlen || j0 < tlen) {
            const int Tq = qlen - i0 < T ? qlen - i0 : T;
            const int Tt = tlen - j0 < T ? tlen - j0 : T;
            const bool q_done = (i0 + Tq == qlen), t_done = (j0 + Tt == tlen);
            const bool last = q_done && t_done;
            /* stage the target codes of the tile */
            for (int x = L; x < Tt; x += 32) ts[x] = (uint8_t)as_base_at(a.tbase, toff + (uint64_t)(j0 + x));
            /* this lane's query rows */
            int qv[R], h[R];
            {
                const int row0 = L * R;
                uint32_t w = row0 < Tq ? as_bases16(qp, (uint64_t)(i0 + row0)) : 0u;
#pragma unroll
                for (int r = 0; r < R; r++) {
                    qv[r] = (row0 + r < Tq) ? (int)((w >> (2 * r)) & 3u) : 4;
                    h[r] = row0 + r + 1;                         /* D[i][0] = i */
                }
            }
            __syncwarp();
            int prev_recv = L * R;                               /* D[L*R][0] */
            int rowmin_key = (Tq << 12) | (Tt << 1);             /* bottom-row candidates, column 0 first */
            const int nsteps = Tt + 31;
            for (int s = 1; s <= nsteps; s++) {
                int recv = __shfl_up_sync(0xFFFFFFFFu, h[R - 1], 1);
                const int j = s - L;
                if (j >= 1 && j <= Tt) {
                    const int tc = ts[j - 1];
                    int up = L == 0 ? j : recv;
                    int diag = L == 0 ? j - 1 : prev_recv;
                    prev_recv = recv;
                    uint32_t word = 0;
#pragma unroll
                    for (int r = 0; r < R; r++) {
                        const int left = h[r];
                        const int neq = qv[r] != tc;
                        const int d = diag + neq;
                        const int u1 = up + 1, l1 = left + 1;
                        const int m = min(u1, l1);
                        const int best = min(d, m);
                        /* op code of the step taken into this cell: 0 '=', 3 'X', 1 up ('I'), 2 left ('D') */
                        const uint32_t code = d <= m ? (neq ? 3u : 0u) : (u1 <= l1 ? 1u : 2u);
                        word |= code << (2 * r);
                        diag = left;
                        up = best;
                        h[r] = best;
                    }
                    myptr[j - 1] = (PT)word;
                    const int key = (h[R - 1] << 12) | ((Tt - j) << 1);
                    rowmin_key = min(rowmin_key, key);
                }
            }
            __syncwarp();
            st_cells += (unsigned long long)Tq * (unsigned long long)Tt;
            /* ---- end cell ---- */
            int ei = Tq, ej = Tt;
            if (!last) {
                int best = 0x7FFFFFFF;
                if (!q_done) best = __shfl_sync(0xFFFFFFFFu, rowmin_key, 31);   /* Tq == T == 32R: row T is lane 31's last */
                if (!t_done) {
                    int ck = 0x7FFFFFFF;
#pragma unroll
                    for (int r = 0; r < R; r++) {
                        const int i = L * R + r + 1;
                        if (i <= Tq) ck = min(ck, (h[r] << 12) | ((Tq - i) << 1) | 1);
                    }
                    if (L == 0) ck = min(ck, (Tt << 12) | (Tq << 1) | 1);       /* i = 0: D[0][Tt] = Tt */
#pragma unroll
                    for (int d = 16; d >= 1; d >>= 1) ck = min(ck, __shfl_xor_sync(0xFFFFFFFFu, ck, d));
                    best = min(best, ck);
                }
                const int dist = (best >> 1) & 0x7FF;
                if (best & 1) { ei = Tq - dist; ej = Tt; } else { ei = Tq; ej = Tt - dist; }
            }
            /* ---- traceback (uniform across the warp; lane 0 records) ---- */
            int ci = ei, cj = ej, m = 0, ai = ei, aj = ej;
            bool anchored = false;
            while (ci > 0 || cj > 0) {
                uint32_t op;
                if (ci == 0) op = 2;
                else if (cj == 0) op = 1;
                else {
                    const int rl = (ci - 1) / R, rr = (ci - 1) - rl * R;
                    op = ((uint32_t)ptr[(size_t)rl * TS + (cj - 1)] >> (2 * rr)) & 3u;
                }
                const int si = ci - (op != 2u), sj = cj - (op != 1u);
                if (last || (si < inner && sj < inner)) {
                    if (!anchored) { ai = ci; aj = cj; anchored = true; }
                    if (L == 0) tmp[m] = (uint8_t)op;
                    m++;
                    score += (op != 0u);
                }
                ci = si;
                cj = sj;
            }
            __syncwarp();
            if (m == 0 && !last) { score = -1; break; }
            if (ops) {
                for (int x = L; x < m; x += 32)
                    if ((size_t)(n + x) < a.ops_stride) ops[n + x] = tmp[m - 1 - x];
            }
            n += m;
            i0 += ai;
            j0 += aj;
            __syncwarp();
            if (last) break;
        }
        if (L == 0) {
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
    if (L == 0 && st_cells) atomicAdd(a.stats + 5, st_cells);
}

template <int R>
static int launch_gact_r(as_gpu *h, const GactArgs &a, size_t n) {
    int rc;
    uint64_t cap = a.n_dev ? AS_GACT_RETRY_WARP_MAX : (uint64_t)h->sm_count * 32;
    unsigned grid = (unsigned)(n < cap ? n : cap);
    if constexpr (R <= 8) {
        size_t smem = (size_t)32 * a.TS * 2 + 2 * (size_t)a.TS + a.TQ + 16;
        rc = cudaFuncSetAttribute(k_gact<R, uint16_t>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (rc != cudaSuccess) { as_set_error("cudaFuncSetAttribute(gact): %s", cudaGetErrorString((cudaError_t)rc)); return AS_ECUDA; }
        k_gact<R, uint16_t><<<grid, 32, smem, h->stream>>>(a);
    } else {
        size_t smem = (size_t)32 * a.TS * 4 + 2 * (size_t)a.TS + a.TQ + 16;
        rc = cudaFuncSetAttribute(k_gact<R, uint32_t>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (rc != cudaSuccess) { as_set_error("cudaFuncSetAttribute(gact): %s", cudaGetErrorString((cudaError_t)rc)); return AS_ECUDA; }
        k_gact<R, uint32_t><<<grid, 32, smem, h->stream>>>(a);
    }
    h->launches++;
    AS_CUDA(cudaGetLastError());
    return AS_OK;
}

static int gact_warp_dispatch(as_gpu *h, const GactArgs &a, size_t n, uint32_t max_qlen) {
    /* rows per lane: the whole tile when some query reaches T rows, else just enough for the longest query */
    const int T = a.T;
    int rows = (int)(max_qlen < (uint32_t)T ? max_qlen : (uint32_t)T);
    int R = (rows + 31) / 32;
    if (max_qlen >= (uint32_t)T) R = T / 32;
    if (R < 1) R = 1;
    /* a shorter tile buffer for short queries would not change results: T only bounds Tq/Tt */
    switch (R) {
    case 1: return launch_gact_r<1>(h, a, n);
    case 2: return launch_gact_r<2>(h, a, n);
    case 3: return launch_gact_r<3>(h, a, n);
    case 4: return launch_gact_r<4>(h, a, n);
    case 5: return launch_gact_r<5>(h, a, n);
    case 6: return launch_gact_r<6>(h, a, n);
    case 7: return launch_gact_r<7>(h, a, n);
    case 8: return launch_gact_r<8>(h, a, n);
    case 9: return launch_gact_r<9>(h, a, n);
    case 10: return launch_gact_r<10>(h, a, n);
    case 11: return launch_gact_r<11>(h, a, n);
    case 12: return launch_gact_r<12>(h, a, n);
    case 13: return launch_gact_r<13>(h, a, n);
    case 14: return launch_gact_r<14>(h, a, n);
    case 15: return launch_gact_r<15>(h, a, n);
    default: return launch_gact_r<16>(h, a, n);
    }
}

int as_launch_gact_warp_list(as_gpu *h, const GactArgs &a, uint32_t max_qlen) {
    return gact_warp_dispatch(h, a, a.n, max_qlen);
}

int as_launch_gact(as_gpu *h, const as_params *p, const uint32_t *d_qpack, size_t q_wstride, const uint32_t *d_qlens,
                   const uint32_t *d_tbase, const uint64_t *d_toff, const uint32_t *d_tlens, size_t n,
                   uint32_t max_qlen, uint32_t max_tlen, uint8_t *d_ops, size_t ops_stride, int32_t *d_nops,
                   int32_t *d_score, as_result *d_res) {
    if (n == 0) return AS_OK;
    const int T = p->gact_T, O = p->gact_O;
    if (T < 32 || T > 512 || (T & 31) || O < 0 || O >= T) {
        as_set_error("GACT: tile size T=%d must be a multiple of 32 in [32,512] and 0 <= O=%d < T", T, O);
        return AS_EINVAL;
    }
    GactArgs a;
    a.qpack = d_qpack; a.q_wstride = q_wstride; a.qlens = d_qlens;
    a.tbase = d_tbase; a.toff = d_toff; a.tlens = d_tlens;
    a.n = n; a.T = T; a.O = O;
    a.TS = (int)(max_tlen < (uint32_t)T ? max_tlen : (uint32_t)T);
    a.TS = (a.TS + 2) & ~1;      /* even: lane stride TS-1 is odd, so a step's stores hit distinct banks */
    a.TQ = (int)(max_qlen < (uint32_t)T ? max_qlen : (uint32_t)T);
    a.ops = d_ops; a.ops_stride = ops_stride; a.n_ops = d_nops; a.score = d_score; a.res = d_res;
    a.ops_zeroed = 1;            /* both callers clear the rows first (run_stages, as_cigar_align_batch) */
    AS_TRY(as_scratch(h, &h->misc[5], 64));
    AS_CUDA(cudaMemsetAsync(h->misc[5].p, 0, 64, h->stream));
    a.work_counter = (unsigned int *)h->misc[5].p;
    a.stats = h->d_stats;
    a.band = p->gact_band;
    a.order_mode = p->gact_order;
    a.blocks_per_sm = p->gact_blocks_per_sm;
    a.fail_count = nullptr;
    a.order = nullptr;
    a.retry_list = nullptr;
    a.retry_count = nullptr;
    a.n_dev = nullptr;
    a.retry_mode = 0;
    a.emit_runs = h->gact_emit_runs ? 1 : 0;
    h->gact_emit_runs = false;
    /* two implementations of the same tile DP: bit-parallel, one read per lane (as_gact_bp.cu) for batches that fill
     * the GPU with lanes; anti-diagonal, one read per warp (below) for small batches */
    const bool force_warp = (p->flags & AS_F_GACT_WARP) != 0, force_bp = (p->flags & AS_F_GACT_BITPAR) != 0;
    if (!force_warp && (force_bp || n >= 2048)) return as_launch_gact_bp(h, a, n, max_qlen, max_tlen);
    return gact_warp_dispatch(h, a, n, max_qlen);
}
