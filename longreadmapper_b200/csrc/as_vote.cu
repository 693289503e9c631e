/*
 * as_vote.cu -- PART 1 of the per-read loop (alnmain.c:333-405): phase-by-phase seeding, SA-hit voting
 * into 16-bp buckets (histo.c:26-56), top-2 selection (histo.c:84-96) and the phase control with its
 * last-phase override (alnmain.c:371-404).
 *
 * Mapping: a group of G lanes (8, 16 or 32, chosen from the longest read of the batch) owns one read and
 * walks its phases; within a phase every lane owns one seed at a time: lc-table gather, then up to
 * seed_len-hlen backward-search steps (two 32-byte sector gathers each), then the group spreads the SA
 * rows of all surviving seeds over its lanes and votes into a shared-memory hash table keyed by bucket
 * (atomicCAS claim, atomicAdd count, atomicMin key / first-seen order).  The reference's vote list is
 * order-dependent only through tie-breaks in TOP2; `ord` = (seed index in phase, SA row offset) restores
 * that order exactly.  Reads whose distinct buckets overflow the shared table are
 * queued and redone by the same code over a large table in global memory (BIG instantiation).
 */
#include "as_common.cuh"
#include <type_traits>

#define VOTE_THREADS 128
#ifndef VOTE_CLAIM
#define VOTE_CLAIM 1
#endif
/* intervals of at most this many rows are finished against the text.  With the sax table the rows' SA values and
 * contexts are neighbours in one or two sectors, so two rows cost what one does and save the FM steps that would
 * have split them (c4: 369M -> 153M steps, PART 1 35.9 -> 33.5 ms; 3 and 4 rows bring nothing more).  Without the
 * table a second row is two more gathers: worth it when the index sits in L2 for good part (one super block),
 * not at human scale (measured 39 -> 44 ms), see vw below */
#ifndef VOTE_VW
#define VOTE_VW 2
#endif
#ifndef VOTE_MIN_BLOCKS
#define VOTE_MIN_BLOCKS 6
#endif
#ifndef VOTE_CARVEOUT
#define VOTE_CARVEOUT (-1)               /* >= 0: preferred shared-memory carve-out of the kernel, per cent */
#endif
#define MAXP 34                      /* seed_len <= 32 -> at most 33 phases */
#define EMPTY_BUCKET 0xFFFFFFFFFFFFFFFFULL

template <typename USED_T>
struct VoteTableT {
    unsigned long long *bucket;
    uint32_t *val, *ord, *low;
    USED_T *used;                       /* slots claimed in this phase, in claim order */
    uint32_t *n_used;
    uint32_t cap, shift;                /* cap = 1 << (64 - shift) */
};

struct VoteArgs {
    const uint32_t *pack;
    size_t wstride;
    const uint32_t *lens;
    size_t n;
    int sl;
    uint32_t thres;
    VoteOut *out;
    unsigned int *work_counter;
    /* overflow queue */
    unsigned int *ovf_count;
    uint32_t *ovf_list;
    /* BIG */
    unsigned char *big_mem;
    size_t big_stride;
    uint32_t big_cap;
    unsigned long long *stats;   /* as_stats layout */
    /* group kernel in list mode: its reads are list[0 .. *list_count) (the reads the thread-per-read kernel deferred) */
    const uint32_t *list;
    const unsigned int *list_count;
    /* thread-per-read kernel: reads it hands to the group kernel */
    uint32_t *dlist;
    unsigned int *dlist_count;
    int use2;                    /* seeds start from the deep table (DevIndex.lc2) instead of the 12-mer table */
    int use_ectx;                /* decide one- and two-row entries from the context bases they carry (as_lc2.cu) */
};

__device__ __forceinline__ uint32_t hash_bucket(uint64_t b, uint32_t shift) {
    return (uint32_t)((b * 0x9E3779B97F4A7C15ULL) >> shift);
}

template <typename USED_T, typename VT>
__device__ __forceinline__ bool table_insert(const VT &t, uint64_t bucket, uint32_t low, uint32_t ord) {
    uint32_t slot = hash_bucket(bucket, t.shift);
    for (uint32_t probes = 0; probes < t.cap; probes++) {
        unsigned long long prev = atomicCAS(t.bucket + slot, EMPTY_BUCKET, (unsigned long long)bucket);
        if (prev == EMPTY_BUCKET || prev == bucket) {
            if (prev == EMPTY_BUCKET) t.used[atomicAdd(t.n_used, 1u)] = (USED_T)slot;
            atomicAdd(t.val + slot, 1u);
            atomicMin(t.ord + slot, ord);
            atomicMin(t.low + slot, low);
            return true;
        }
        slot = (slot + 1) & (t.cap - 1);
    }
    return false;
}

/* top-2 under (val desc, ord asc), as histo_find_2_max leaves them (histo.c:84-96) */
__device__ __forceinline__ void top2_push(uint64_t &a0, uint64_t &a1, uint64_t c) {
    if (c > a0) {
        a1 = a0;
        a0 = c;
    } else if (c > a1) {
        a1 = c;
    }
}

/* warp-wide max of a small unsigned value (all 32 lanes take part) */
__device__ __forceinline__ uint32_t warp_max_u32(uint32_t v) {
    return __reduce_max_sync(0xFFFFFFFFu, v);                       /* REDUX.MAX: one instruction */
}

/*
 * One seed of a phase: the SA interval lc_aln / fmi_aln reach for it (lchash.c:89-104, fmidx.c:295-313), or -- when the text
 * decided the tail of a one-row walk -- the position of its only hit.  Shared by the two PART 1 kernels: the group kernel
 * loads the table entry here, the thread-per-read kernel hands in the entry it loaded ahead (have_e).
 * Out: rr rows [kk, ll] survive (rr == 0: none); have_pos: rr == 1 and hit_pos = SA of that row (no SA access needed).
 */
template <bool SMALL>
__device__ __forceinline__ void vote_resolve_seed(const DevIndex &ix, const VoteArgs &a, const BsTab *s_tab, const uint32_t *s_tabChi,
                                                  const int sl, const int left0, const uint64_t hmask, const uint64_t seed,
                                                  const bool have_e, const unsigned long long e_in, uint64_t &kk, uint64_t &ll,
                                                  uint64_t &rr, bool &have_pos, uint64_t &hit_pos, unsigned long long &st_seeds,
                                                  unsigned long long &st_steps, unsigned int &st_fm, unsigned int &st_txt) {
    bool absent = false;
    uint32_t etag = 0, ectx = 0;
    int tt = left0 - 1;                                             /* next seed base to walk */
    if (a.use2) {
        /* deep table: the interval lc_access + (hlen2 - hlen) steps of fmi_aln would have reached */
        const unsigned long long e = have_e ? e_in : __ldg(ix.lc2 + ((seed >> (2 * left0)) & hmask));
        const uint32_t info = (uint32_t)(e >> 40);
        etag = info >> 22;                                          /* as_lc2.cu: 0 count, 1 / 2 rows + context */
        ectx = info & 0x3FFFFFu;
        const uint32_t cnt = etag ? etag : ectx;
        kk = e & 0xFFFFFFFFFFULL;
        ll = kk + cnt - 1;
        absent = e == 0;
        if (info == 0x3FFFFFu) {                                     /* too many rows for the entry: 12-mer table */
            tt = sl - ix.hlen - 1;
            const ulonglong2 kl = __ldg(ix.lc + ((seed >> (2 * (tt + 1))) & ((1ULL << (2 * ix.hlen)) - 1ULL)));
            kk = kl.x;
            ll = kl.y;
            absent = (kk == 0 && ll == 0);
        } else if (!absent) {
            st_steps += (unsigned)(ix.hlen2 - ix.hlen);            /* steps of the reference's loop it stands for */
        }
    } else if (left0 >= 0) {                                        /* lchash.c:93-101 */
        ulonglong2 kl = __ldg(ix.lc + ((seed >> (2 * left0)) & hmask));
        kk = kl.x;
        ll = kl.y;
        absent = (kk == 0 && ll == 0);
    }
    st_seeds++;
    if (!absent) {
        /* Backward search (fmidx.c:300-310) until the interval is empty, the seed is used up, or at most
         * VOTE_VW rows are left.  A row r of the interval survives the rest of the walk iff the remaining bases
         * [0,tt] of the seed equal the tt+1 bases in front of SA[r] -- the text decides -- and the row it
         * maps to has SA value SA[r]-(tt+1).  So the tail of the walk is one SA
         * gather and one text gather instead of 2(tt+1) dependent sector gathers, and the vote needs no
         * further SA access.  st_steps still counts the steps the reference's loop executes (up to and
         * including the failing one); st_fm counts the FM steps actually done here. */
        bool fail = false;
        /* entry with context that covers the rest of the seed: no FM step is taken (the interval has <= 2 rows);
         * a one-row entry carries the row's text POSITION in place of the row (as_lc2.cu): its seed needs no
         * further gather unless more than 11 bases are left (then one text gather) */
        const bool e_pos = etag == 1;
        const bool e_ok = tt >= 0 && ((e_pos && tt < 11) || (a.use_ectx && etag == 2 && tt < 5));
        const uint32_t vw = (SMALL || ix.sax != nullptr || e_ok) ? (uint32_t)VOTE_VW : 1u;
        if (e_pos && tt < 0) {                                      /* the table answered the whole seed */
            have_pos = true;
            hit_pos = kk;
        }
        if (SMALL) {
            uint32_t k32 = (uint32_t)kk, l32 = (uint32_t)ll;
            while (tt >= 0 && l32 - k32 >= vw) {
                const uint32_t c = (uint32_t)(seed >> (2 * tt)) & 3u;
                as_bs_step32t(ix, *reinterpret_cast<const uint4 *>(s_tab + c), k32, l32);
                st_steps++;
                st_fm++;
                tt--;
                if (k32 > l32) { fail = true; break; }
            }
            kk = k32;
            ll = l32;
        } else {
            while (tt >= 0 && ll - kk >= (uint64_t)vw) {
                const uint32_t c = (uint32_t)(seed >> (2 * tt)) & 3u;
                as_bs_step64t(ix, *reinterpret_cast<const uint4 *>(s_tab + c), s_tabChi[c], c, kk, ll);
                st_steps++;
                st_fm++;
                tt--;
                if (kk > ll) { fail = true; break; }
            }
        }
        if (!fail && tt >= 0) {
            /* 1..VOTE_VW rows left, tt+1 bases to go: every row's tail is read off the text -- the SA gathers
             * of the rows are independent, so are the text gathers, where the FM walk is a chain */
            const int nrem = tt + 1;
            const uint32_t wdt = (uint32_t)(ll - kk) + 1u;
            uint64_t pr[VOTE_VW];
            uint32_t pv[VOTE_VW];                  /* with sax: the nrem bases in front of SA[r], no text gather */
            /* the table entry itself carries the context of its one or two rows (as_lc2.cu, tags 1 and 2):
             * nothing is fetched to decide, and only a surviving row's position is */
            const bool from_e = e_ok;
            const bool ctx = from_e || (!e_pos && ix.sax != nullptr && nrem <= 12);
            bool in_text = true;
#pragma unroll
            for (int r = 0; r < VOTE_VW; r++) {
                pv[r] = 0;
                if ((uint32_t)r >= wdt) {
                    pr[r] = ~0ULL;
                } else if (e_pos) {
                    pr[r] = kk;                                      /* the position; >= 32 by construction */
                    pv[r] = ectx >> (2 * (11 - (nrem < 11 ? nrem : 11)));
                } else if (from_e) {
                    pr[r] = 1ULL << 62;                             /* not known yet; the builder checked SA >= context */
                    pv[r] = ((ectx >> (10 * r)) & 0x3FFu) >> (2 * (5 - nrem));            /* tag 2 */
                } else if (ctx) {
                    const unsigned long long e = __ldg(ix.sax + kk + (uint64_t)r);
                    pr[r] = e & 0xFFFFFFFFFFULL;
                    pv[r] = (uint32_t)(e >> 40) >> (2 * (12 - nrem));
                } else {
                    pr[r] = as_locate(ix, kk + (uint64_t)r);
                }
                in_text &= pr[r] >= (uint64_t)nrem;
            }
            bool resolved = false;
            if (in_text) {
                int nsurv = 0, steps = 0;
                uint64_t pos1 = 0;
                uint32_t rsurv = 0;
#pragma unroll
                for (int r = 0; r < VOTE_VW; r++) {
                    if ((uint32_t)r < wdt) {
                        int sr;
                        if (nrem <= 16) {
                            const uint32_t tb = ctx ? pv[r] : as_bases16(ix.text2, pr[r] - (uint64_t)nrem);
                            const uint32_t x = ((uint32_t)seed ^ tb) & (nrem == 16 ? 0xFFFFFFFFu : ((1u << (2 * nrem)) - 1u));
                            sr = x == 0 ? nrem : nrem - ((31 - __clz(x)) >> 1);
                            if (x == 0) { nsurv++; pos1 = pr[r] - (uint64_t)nrem; rsurv = (uint32_t)r; }
                        } else {
                            const uint64_t x = (seed ^ as_bases32(ix.text2, pr[r] - (uint64_t)nrem)) & ((1ULL << (2 * nrem)) - 1ULL);
                            sr = x == 0 ? nrem : nrem - ((63 - __clzll((long long)x)) >> 1);
                            if (x == 0) { nsurv++; pos1 = pr[r] - (uint64_t)nrem; rsurv = (uint32_t)r; }
                        }
                        steps = max(steps, sr);      /* the reference's loop stops when the last row drops out */
                    }
                }
                if (nsurv <= 1) {
                    st_steps += steps;
                    st_txt++;
                    resolved = true;
                    if (nsurv == 1) {
                        if (from_e && !e_pos) {
                            const uint64_t row = kk + rsurv;
                            pos1 = (ix.sax ? (__ldg(ix.sax + row) & 0xFFFFFFFFFFULL) : as_locate(ix, row)) - (uint64_t)nrem;
                        }
                        have_pos = true;
                        hit_pos = pos1;
                        ll = kk;                                    /* one row survives: rr = 1 */
                    } else {
                        fail = true;
                    }
                }
            }
            if (!resolved) {       /* two or more rows survive, or the text starts inside the seed: walk it */
                while (tt >= 0) {
                    const uint32_t c = (uint32_t)(seed >> (2 * tt)) & 3u;
                    as_bs_step64t(ix, *reinterpret_cast<const uint4 *>(s_tab + c), s_tabChi[c], c, kk, ll);
                    st_steps++;
                    st_fm++;
                    tt--;
                    if (kk > ll) { fail = true; break; }
                }
            }
        }
        if (fail) { kk = 1; ll = 0; }
        rr = kk > ll ? 0 : ll - kk + 1;
    }
}

/*
 * The 32/G groups of a warp run as one state machine: every trip of the outer loop is ONE PHASE of the read each
 * group currently owns (groups that finished a read pick the next one at the top of the trip).  All loop bounds
 * inside a trip are warp-wide maxima and the per-group work is predicated, so the warp never splits into serial
 * sub-warps: reads that need 1 phase and reads that need 21 share full-width instructions.
 */
template <int G, bool BIG, bool SMALL>
__global__ void __launch_bounds__(BIG ? 32 : VOTE_THREADS, BIG ? 1 : (G == 32 ? VOTE_MIN_BLOCKS + 1 : VOTE_MIN_BLOCKS)) k_vote(DevIndex ix, VoteArgs a) {
    constexpr int CAP = 8 * G;
    constexpr int GROUPS = VOTE_THREADS / G;
    constexpr uint32_t FULL = 0xFFFFFFFFu;
    typedef typename std::conditional<BIG, uint32_t, uint8_t>::type USED_T;
    __shared__ uint8_t s_used[BIG ? 1 : GROUPS * CAP];
    __shared__ uint32_t s_nused[BIG ? 1 : GROUPS];
    __shared__ unsigned long long s_bucket[BIG ? 1 : GROUPS * CAP];
    __shared__ uint32_t s_val[BIG ? 1 : GROUPS * CAP], s_ord[BIG ? 1 : GROUPS * CAP], s_low[BIG ? 1 : GROUPS * CAP];
    __shared__ unsigned long long s_okey[(BIG ? 1 : GROUPS) * MAXP];
    __shared__ uint32_t s_oval[(BIG ? 1 : GROUPS) * MAXP];
    __shared__ BsTab s_tab[4];                 /* per-character constants of a backward-search step */
    __shared__ uint32_t s_tabChi[4];

    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t g = lane & (G - 1);
    const uint32_t gmask = (G == 32) ? 0xFFFFFFFFu : (((1u << G) - 1u) << (lane & ~(uint32_t)(G - 1)));
    const int grp = threadIdx.x / G;

    VoteTableT<USED_T> t;
    if (BIG && *(volatile unsigned int *)a.ovf_count == 0) return;      /* nothing overflowed: do not even clear the table */
    if (BIG) {
        unsigned char *m = a.big_mem + (size_t)blockIdx.x * a.big_stride;
        t.cap = a.big_cap;
        t.bucket = (unsigned long long *)m;
        t.val = (uint32_t *)(m + (size_t)t.cap * 8);
        t.ord = t.val + t.cap;
        t.low = t.ord + t.cap;
        t.used = (USED_T *)(t.low + t.cap);
        t.n_used = (uint32_t *)(t.low + 2 * (size_t)t.cap);
    } else {
        t.cap = CAP;
        t.bucket = s_bucket + grp * CAP;
        t.val = s_val + grp * CAP;
        t.ord = s_ord + grp * CAP;
        t.low = s_low + grp * CAP;
        t.used = (USED_T *)(s_used + grp * CAP);
        t.n_used = s_nused + grp;
    }
    t.shift = 64 - (31 - __clz(t.cap));
    unsigned long long *okey = s_okey + (BIG ? 0 : grp) * MAXP;
    uint32_t *oval = s_oval + (BIG ? 0 : grp) * MAXP;

    /* the table starts empty and every phase resets exactly the slots it claimed */
    for (uint32_t s = g; s < t.cap; s += G) {
        t.bucket[s] = EMPTY_BUCKET;
        t.val[s] = 0;
        t.ord[s] = 0xFFFFFFFFu;
        t.low[s] = 0xFFFFFFFFu;
    }
    if (g == 0) *t.n_used = 0;
    as_bs_tab_init(ix, s_tab, s_tabChi, threadIdx.x);
    __syncthreads();

    const int sl = a.sl, P = sl + 1;
    const int hl = a.use2 ? ix.hlen2 : ix.hlen;          /* bases answered by the table */
    const int left0 = sl - hl;
    const uint64_t hmask = (1ULL << (2 * hl)) - 1ULL;
    unsigned long long st_seeds = 0, st_steps = 0, st_hits = 0, st_ovf = 0;
    unsigned int st_queued = 0;                  /* reads this warp handed to the overflow (BIG) pass */
    unsigned int st_fm = 0, st_txt = 0;          /* FM steps done here / tails decided by the text */

    /* state of the read this group owns */
    bool have = false, drained = false;
    uint32_t claim_next = 0, claim_end = 0;
    uint32_t read = 0, len = 0, num_seeds = 0, span = 0, n_outer = 0, span_q = 0, span_r = 0;
    const uint32_t *pr = a.pack;
    int iter = 0;

    for (;;) {
        /* ---- groups without a read take the next one ---- */
        if (!have && !drained) {
            /* reads are claimed VOTE_CLAIM at a time: one same-address atomic per group and claim, not per read */
            if (claim_next == claim_end) {
                uint32_t w0 = 0;
                if (g == 0) w0 = atomicAdd(a.work_counter, (uint32_t)VOTE_CLAIM);
                claim_next = __shfl_sync(gmask, w0, 0, G);
                claim_end = claim_next + VOTE_CLAIM;
            }
            const uint32_t widx = claim_next++;
            const uint32_t limit = BIG ? *a.ovf_count : (a.list_count ? *a.list_count : (uint32_t)a.n);
            if (widx < limit) {
                read = BIG ? a.ovf_list[widx] : (a.list ? a.list[widx] : widx);
                len = a.lens[read];
                pr = a.pack + (size_t)read * a.wstride;
                num_seeds = len / (uint32_t)P;                                   /* alnmain.c:371 */
                span = len > (uint32_t)sl ? len - (uint32_t)sl : 0;               /* seeds start at j < len - sl (:353) */
                span_q = span / (uint32_t)P;                                      /* one division per read, not per phase */
                span_r = span - span_q * (uint32_t)P;
                n_outer = 0;
                iter = 0;
                have = true;
            } else {
                drained = true;
            }
        }
        if (!__any_sync(FULL, have)) break;

        /* ---- one phase: seeds iter, iter+P, ... of the read (alnmain.c:353) ---- */
        /* seeds of the phase = ceil((span - iter) / P) = span/P + (iter < span%P), because 0 <= iter < P */
        const uint32_t m_count = have ? span_q + ((uint32_t)iter < span_r ? 1u : 0u) : 0u;
        const uint32_t m_max = warp_max_u32(m_count);
        bool ok = true, reg_vote = false;
        uint32_t v0val = 0, v1val = 0;
        uint64_t c0key = 0;
        for (uint32_t c0 = 0; c0 < m_max; c0 += G) {
            const uint32_t m = c0 + g;
            const uint32_t j = (uint32_t)iter + m * (uint32_t)P;
            uint64_t kk = 1, ll = ix.N - 1, rr = 0;                             /* alnmain.c:354 */
            bool have_pos = false;                                               /* hit_pos = SA of the seed's one row */
            uint64_t hit_pos = 0;
            /* (a group whose table overflowed in an earlier round of the phase resolves no more seeds: see below) */
            const bool gfull = !BIG && c0 > 0 && (__ballot_sync(FULL, !ok) & gmask) != 0;
            if (m < m_count && !gfull) {
                const uint64_t seed = as_bases32(pr, j);
                vote_resolve_seed<SMALL>(ix, a, s_tab, s_tabChi, sl, left0, hmask, seed, false, 0ULL, kk, ll, rr, have_pos, hit_pos,
                                         st_seeds, st_steps, st_fm, st_txt);
            }
            __syncwarp();
            const uint32_t rr_eff = (rr > 0 && rr < a.thres) ? (uint32_t)rr : 0u;   /* alnmain.c:360 */
            if (m_max <= (uint32_t)G && !__any_sync(FULL, rr_eff > 1)) {
                /* the whole phase is this one round and every surviving seed has exactly one SA row: at most one
                 * vote per lane, so the vote list is built in registers.  Lanes voting for the same bucket of the
                 * same read find each other with MATCH.ANY; the entry's val is their number, its list position the
                 * lowest such lane (votes are cast in seed order = lane order, histo.c:42-56), its key the smallest
                 * key among them.  TOP2 (histo.c:84-96) is then a 3..5-round butterfly on 32-bit words. */
                const bool valid = rr_eff != 0;
                uint64_t key = 0;
                if (valid) {
                    key = (have_pos ? hit_pos : as_locate(ix, kk)) - (uint64_t)j;   /* alnmain.c:363-365, u64 wrap kept */
                    st_hits++;
                }
                const uint64_t tag = valid ? ((key >> 4) | ((uint64_t)(lane / G) << 60)) : (0xF000000000000000ULL | lane);
                const uint32_t peers = __match_any_sync(FULL, tag);
                const uint32_t low = (uint32_t)key & 15u;
                const uint32_t b3 = __ballot_sync(FULL, low & 8u), b2 = __ballot_sync(FULL, low & 4u),
                               b1 = __ballot_sync(FULL, low & 2u), b0 = __ballot_sync(FULL, low & 1u);
                uint32_t cand = peers, minlow = 0, z;                           /* min of `low` over the peers */
                z = cand & ~b3; if (z) cand = z; else minlow |= 8u;
                z = cand & ~b2; if (z) cand = z; else minlow |= 4u;
                z = cand & ~b1; if (z) cand = z; else minlow |= 2u;
                z = cand & ~b0; if (z) cand = z; else minlow |= 1u;
                const bool leader = valid && ((uint32_t)__ffs((int)peers) - 1u) == lane;
                uint32_t a0 = leader ? (((uint32_t)__popc(peers) << 5) | (31u - lane)) : 0u, a1 = 0u;
#pragma unroll
                for (int d = G >> 1; d >= 1; d >>= 1) {
                    const uint32_t o0 = __shfl_xor_sync(FULL, a0, d), o1 = __shfl_xor_sync(FULL, a1, d);
                    const uint32_t mn = min(a0, o0), mx1 = max(a1, o1);
                    a0 = max(a0, o0);
                    a1 = max(mn, mx1);
                }
                v0val = a0 >> 5;
                v1val = a1 >> 5;
                const uint32_t wl = 31u - (a0 & 31u);
                const uint64_t wkey = __shfl_sync(FULL, key, wl);
                const uint32_t wlow = __shfl_sync(FULL, minlow, wl);
                c0key = v0val ? ((wkey & ~15ULL) | wlow) : 0ULL;
                reg_vote = true;
            } else if (!__any_sync(FULL, rr_eff > 1)) {
                /* every surviving seed of the warp has exactly one SA row */
                if (rr_eff) {
                    uint64_t key = (have_pos ? hit_pos : as_locate(ix, kk)) - (uint64_t)j;   /* alnmain.c:363-365, u64 wrap kept */
                    st_hits++;
                    if (!table_insert<USED_T>(t, key >> 4, (uint32_t)key & 15u, m * a.thres)) ok = false;
                }
            } else {
                /* spread all SA rows of the group's seeds over its lanes */
                uint32_t incl = rr_eff;
#pragma unroll
                for (int d = 1; d < G; d <<= 1) {
                    uint32_t o = __shfl_up_sync(gmask, incl, d, G);
                    if ((int)g >= d) incl += o;
                }
                const uint32_t base = incl - rr_eff;
                const uint32_t V = __shfl_sync(gmask, incl, G - 1, G);
                const uint32_t V_max = warp_max_u32(V);
                /* the overflow pass has one warp per read and four per SM: VU rows per lane are located before any is
                 * inserted, so their SA gathers (and the inserts' first atomics) overlap instead of queueing up */
                constexpr int VU = BIG ? 8 : 1;
                for (uint32_t v0 = 0; v0 < V_max; v0 += G * VU) {
                    uint64_t vkey[VU];
                    uint32_t vord[VU];
                    bool vok[VU];
                    /* a group whose table is full leaves the rest of the phase alone: the read goes to the overflow pass
                     * whatever else it votes, and every further vote would probe the whole table before it failed (c2r:
                     * 1750 votes of a 250-copy family into 64 slots -- the list pass took 10 ms for that alone) */
                    const bool full = !BIG && (__ballot_sync(FULL, !ok) & gmask) != 0;
                    if (full) ok = false;
#pragma unroll
                    for (int u = 0; u < VU; u++) {
                        const uint32_t v = v0 + (uint32_t)u * G + g;
                        /* owner = largest lane whose exclusive base <= v */
                        uint32_t lo = 0;
#pragma unroll
                        for (int d = G >> 1; d >= 1; d >>= 1) {
                            uint32_t b = __shfl_sync(gmask, base, lo + d, G);
                            if (b <= v) lo += d;
                        }
                        const uint32_t ob = __shfl_sync(gmask, base, lo, G);
                        const uint32_t okl = __shfl_sync(gmask, (uint32_t)kk, lo, G);
                        const uint32_t okh = __shfl_sync(gmask, (uint32_t)(kk >> 32), lo, G);
                        const uint32_t opl = __shfl_sync(gmask, (uint32_t)hit_pos, lo, G);
                        const uint32_t oph = __shfl_sync(gmask, (uint32_t)(hit_pos >> 32), lo, G);
                        const bool opk = __shfl_sync(gmask, (int)have_pos, lo, G) != 0;
                        vok[u] = v < V && !full;
                        vkey[u] = 0;
                        vord[u] = 0;
                        if (vok[u]) {
                            const uint32_t om = c0 + lo;
                            const uint32_t oj = (uint32_t)iter + om * (uint32_t)P;
                            const uint32_t off = v - ob;
                            uint64_t row = (((uint64_t)okh << 32) | okl) + off;
                            vkey[u] = (opk ? (((uint64_t)oph << 32) | opl) : as_locate(ix, row)) - (uint64_t)oj;
                            vord[u] = om * a.thres + off;
                            st_hits++;
                        }
                    }
                    if (BIG) {
                        /* The warp owns the table.  The first compare-and-swap of all VU votes of a lane leaves before any
                         * answer is back (a vote was four dependent round trips to L2: claim, list slot, and again for the
                         * next vote; c2r: 23 k cycles per 128 votes); a vote whose slot belongs to another bucket probes on
                         * by itself; the slots claimed by the batch are appended to the list by ONE addition per batch
                         * (prefix sum over the lanes' counts) instead of one atomic per claim. */
                        uint32_t slot[VU];
                        unsigned long long prev[VU];
#pragma unroll
                        for (int u = 0; u < VU; u++) {
                            slot[u] = hash_bucket(vkey[u] >> 4, t.shift);
                            prev[u] = EMPTY_BUCKET;
                            if (vok[u]) prev[u] = atomicCAS(t.bucket + slot[u], EMPTY_BUCKET, (unsigned long long)(vkey[u] >> 4));
                        }
                        uint32_t newmask = 0;
#pragma unroll
                        for (int u = 0; u < VU; u++) {
                            if (vok[u]) {
                                const unsigned long long b = (unsigned long long)(vkey[u] >> 4);
                                unsigned long long pv = prev[u];
                                uint32_t sl2 = slot[u], probes = 1;
                                while (pv != EMPTY_BUCKET && pv != b && probes < t.cap) {        /* someone else's slot */
                                    sl2 = (sl2 + 1) & (t.cap - 1);
                                    pv = atomicCAS(t.bucket + sl2, EMPTY_BUCKET, b);
                                    probes++;
                                }
                                if (pv == EMPTY_BUCKET || pv == b) {
                                    slot[u] = sl2;
                                    if (pv == EMPTY_BUCKET) newmask |= 1u << u;
                                    atomicAdd(t.val + sl2, 1u);
                                    atomicMin(t.ord + sl2, vord[u]);
                                    atomicMin(t.low + sl2, (uint32_t)vkey[u] & 15u);
                                } else {
                                    ok = false;
                                }
                            }
                        }
                        __syncwarp();
                        uint32_t mine = (uint32_t)__popc(newmask), incl2 = mine;
#pragma unroll
                        for (int d = 1; d < 32; d <<= 1) {
                            const uint32_t o = __shfl_up_sync(FULL, incl2, d);
                            if ((int)lane >= d) incl2 += o;
                        }
                        const uint32_t tot = __shfl_sync(FULL, incl2, 31);
                        if (tot) {
                            uint32_t at = *(volatile uint32_t *)t.n_used + incl2 - mine;
#pragma unroll
                            for (int u = 0; u < VU; u++)
                                if (newmask & (1u << u)) t.used[at++] = (USED_T)slot[u];
                            __syncwarp();
                            if (lane == 0) *(volatile uint32_t *)t.n_used = *(volatile uint32_t *)t.n_used + tot;
                            __syncwarp();
                        }
                    } else {
#pragma unroll
                        for (int u = 0; u < VU; u++)
                            if (vok[u] && !table_insert<USED_T>(t, vkey[u] >> 4, (uint32_t)vkey[u] & 15u, vord[u])) ok = false;
                    }
                }
            }
        }
        __syncwarp();
        const bool overflow = have && (__ballot_sync(gmask, !ok) & gmask) != 0;

        if (!reg_vote) {                     /* warp-uniform: reg_vote follows from m_max and a warp vote */
        /* ---- TOP2 of the phase (histo.c:84-96) over the slots the phase claimed ---- */
        uint64_t a0 = 0, a1 = 0;
        uint32_t slot0 = 0;
        const uint32_t nu = *t.n_used;
        const uint32_t nu_max = warp_max_u32(nu);
        for (uint32_t u0 = 0; u0 < nu_max; u0 += G) {
            const uint32_t u = u0 + g;
            if (u < nu) {
                uint32_t s = t.used[u];
                uint64_t c = ((uint64_t)t.val[s] << 32) | (0xFFFFFFFFu - t.ord[s]);
                if (c > a0) slot0 = s;
                top2_push(a0, a1, c);
            }
        }
#pragma unroll
        for (int d = G >> 1; d >= 1; d >>= 1) {
            uint64_t b0 = __shfl_xor_sync(gmask, a0, d, G);
            uint64_t b1 = __shfl_xor_sync(gmask, a1, d, G);
            uint32_t bs = __shfl_xor_sync(gmask, slot0, d, G);
            uint64_t mn = a0 < b0 ? a0 : b0;
            uint64_t mx1 = a1 > b1 ? a1 : b1;
            if (b0 > a0) {
                a0 = b0;
                slot0 = bs;
            }
            a1 = mn > mx1 ? mn : mx1;
        }
        v0val = (uint32_t)(a0 >> 32);
        v1val = (uint32_t)(a1 >> 32);
        if (v0val) c0key = ((uint64_t)t.bucket[slot0] << 4) | t.low[slot0];
        __syncwarp();
        /* ---- reset: the claimed slots, or the whole table when the phase overflowed it ---- */
        if (__any_sync(FULL, overflow)) {
            if (overflow) {
                for (uint32_t s = g; s < t.cap; s += G) {
                    t.bucket[s] = EMPTY_BUCKET;
                    t.val[s] = 0;
                    t.ord[s] = 0xFFFFFFFFu;
                    t.low[s] = 0xFFFFFFFFu;
                }
            }
        }
        for (uint32_t u0 = 0; u0 < nu_max; u0 += G) {
            const uint32_t u = u0 + g;
            if (u < nu && !overflow) {
                uint32_t s = t.used[u];
                t.bucket[s] = EMPTY_BUCKET;
                t.val[s] = 0;
                t.ord[s] = 0xFFFFFFFFu;
                t.low[s] = 0xFFFFFFFFu;
            }
        }
        __syncwarp();
        if (g == 0) *t.n_used = 0;
        __syncwarp();
        }

        /* ---- phase control (alnmain.c:371-404) ---- */
        bool finished = false, broke = false;
        uint64_t best_key = 0, best_val = 0;
        uint32_t vote_outer = 0;             /* this group adds c0key to its outer list */
        if (have && !overflow && num_seeds > 0) {                               /* alnmain.c:373 */
            /* score = (double)v / num_seeds > 0.6 (alnmain.c:371-378) decided in integers: 5v > 3n makes the quotient
             * exceed 0.6 by >= 1/(5n), far more than an ulp; 5v == 3n makes the correctly rounded quotient the double
             * nearest 0.6, i.e. the literal itself (not greater); 5v < 3n puts it below.  Same verdict, no fp64 divide. */
            const uint64_t v = (uint64_t)v0val + v1val;
            if (5ull * v > 3ull * (uint64_t)num_seeds) {                        /* :378-384 */
                best_key = c0key;
                best_val = v0val;
                broke = true;
            } else if (v0val != 0) {                                            /* :386-388 histo_add(outer, cand[0].key) */
                vote_outer = 1;
            }
        }
        if (__any_sync(FULL, vote_outer != 0)) {
            const uint32_t no_max = warp_max_u32(vote_outer ? n_outer : 0u);
            const uint64_t b = c0key >> 4;
            uint32_t found = n_outer;
            for (uint32_t e = 0; e < no_max; e++)
                if (vote_outer && e < n_outer && (okey[e] >> 4) == b) found = e;
            __syncwarp();
            if (vote_outer && g == 0) {
                if (found < n_outer) {
                    oval[found] += 1;
                    if (c0key < okey[found]) okey[found] = c0key;
                } else {
                    okey[n_outer] = c0key;
                    oval[n_outer] = 1;
                }
            }
            if (vote_outer && found == n_outer) n_outer++;
            __syncwarp();
        }
        if (have) {
            if (overflow) {
                finished = true;
            } else {
                if (!broke) iter++;                                            /* a break leaves iter where it was */
                finished = broke || iter >= P;
            }
        }
        if (__any_sync(FULL, finished)) {
            int from_outer = 0;
            if (finished && !overflow && iter >= P - 1) {                       /* alnmain.c:400-403 */
                uint64_t bk = 0;
                uint32_t bv = 0;
                for (uint32_t e = 0; e < n_outer; e++) {
                    if (oval[e] > bv) {
                        bv = oval[e];
                        bk = okey[e];
                    }
                }
                best_key = bk;
                best_val = bv;
                from_outer = 1;
            }
            if (finished && g == 0) {
                if (!overflow) {
                    VoteOut o;
                    o.best_key = best_key;
                    o.best_val = best_val;
                    o.phases = iter;
                    o.from_outer = from_outer;
                    a.out[read] = o;
                } else if (BIG) {
                    /* cannot happen: the big table holds every possible vote of a phase; flag it loudly */
                    VoteOut o;
                    o.best_key = 0;
                    o.best_val = 0;
                    o.phases = -1;
                    o.from_outer = -1;
                    a.out[read] = o;
                } else {
                    a.ovf_list[atomicAdd(a.ovf_count, 1u)] = read;           /* redone by the BIG pass */
                    st_queued++;
                }
            }
            if (finished && overflow && BIG) st_ovf += (g == 0);
            if (finished) have = false;
            __syncwarp();
        }
    }

    /* stats: one atomic per warp and counter */
#pragma unroll
    for (int d = 16; d >= 1; d >>= 1) {
        st_seeds += __shfl_xor_sync(0xFFFFFFFFu, st_seeds, d);
        st_steps += __shfl_xor_sync(0xFFFFFFFFu, st_steps, d);
        st_hits += __shfl_xor_sync(0xFFFFFFFFu, st_hits, d);
        st_ovf += __shfl_xor_sync(0xFFFFFFFFu, st_ovf, d);
        st_fm += __shfl_xor_sync(0xFFFFFFFFu, st_fm, d);
        st_txt += __shfl_xor_sync(0xFFFFFFFFu, st_txt, d);
        st_queued += __shfl_xor_sync(0xFFFFFFFFu, st_queued, d);
    }
    if (lane == 0) {
        if (st_seeds) atomicAdd(a.stats + 6, st_seeds);
        if (st_hits) atomicAdd(a.stats + 7, st_hits);
        if (st_steps) atomicAdd(a.stats + 8, st_steps);
        if (st_ovf) atomicAdd(a.stats + 9, st_ovf);
        if (st_fm) atomicAdd(a.stats + 12, (unsigned long long)st_fm);
        if (st_txt) atomicAdd(a.stats + 13, (unsigned long long)st_txt);
        if (st_queued) atomicAdd(a.stats + 14, (unsigned long long)st_queued);
    }
}


/* ------------------------------------------------------------------------------------------------------------------------
 * Thread-per-read PART 1 for short reads (at most VT_M seeds per phase: 150-bp reads at seed_len 20 have seven).
 *
 * The group kernel above spends ~185 warp instructions per read and phase on cooperation -- ballots, MATCH.ANY, butterflies,
 * shared-memory tables, warp-uniform loop bounds over four reads in four different states -- for seven table lookups and
 * seven votes.  Here a LANE owns a read: it requests the table entries of all the phase's seeds at once (seven independent
 * gathers in flight per lane, 224 per warp), resolves them one after the other with the same vote_resolve_seed the group
 * kernel uses, and keeps the phase's vote list (histo.c:26-56: buckets in first-seen order, a count and the smallest key
 * each) and the read's outer list in registers -- votes are cast in seed order by one thread, so first-seen order is just
 * the order of insertion and TOP2 (histo.c:84-96) is one scan.  Lanes that finish a read take the next one at the top of
 * the trip, so a warp's 32 reads are in 32 different phases and nobody waits for the slowest.
 *
 * What does not fit a thread goes to the group kernel through a queue (list mode): a seed with more than two rows left
 * after the walk (its votes would have to be spread), a phase with more than VT_NV distinct buckets, a read whose outer
 * list outgrows VT_NO.  On the uniform 4.6 Mbp genome that is a few reads in a thousand; repeat families send theirs there.
 * ---------------------------------------------------------------------------------------------------------------------- */
#define VT_THREADS 128
#define VT_M 8                        /* seeds per phase a thread takes */
#define VT_W 16                       /* packed words of a read kept in shared memory */
#define VT_NV 4                       /* distinct buckets of a phase kept in registers */
#define VT_NO 4                       /* distinct buckets of the outer list kept in registers */
#ifndef VT_MIN_BLOCKS
#define VT_MIN_BLOCKS 4
#endif

/* 32 bases from base i of a packed read in shared memory (as_bases32 for __shared__ words) */
__device__ __forceinline__ uint64_t vt_bases32(const uint32_t *w, uint32_t i) {
    const uint32_t *p = w + (i >> 4);
    const uint32_t x = p[0], y = p[1], z = p[2], sh = 2u * (i & 15u);
    return (uint64_t)__funnelshift_r(x, y, sh) | ((uint64_t)__funnelshift_r(y, z, sh) << 32);
}

template <bool SMALL>
__global__ void __launch_bounds__(VT_THREADS, VT_MIN_BLOCKS) k_vote_t(DevIndex ix, VoteArgs a) {
    constexpr uint32_t FULL = 0xFFFFFFFFu;
    __shared__ uint32_t s_rd[VT_THREADS * (VT_W + 1)];      /* row stride 17 words: a warp's accesses fall in 32 banks */
    __shared__ BsTab s_tab[4];
    __shared__ uint32_t s_tabChi[4];
    const uint32_t lane = threadIdx.x & 31u;
    uint32_t *rd = s_rd + threadIdx.x * (VT_W + 1);
    as_bs_tab_init(ix, s_tab, s_tabChi, threadIdx.x);
    __syncthreads();

    const int sl = a.sl, P = sl + 1;
    const int hl = a.use2 ? ix.hlen2 : ix.hlen;            /* bases answered by the table; <= sl (launcher) */
    const int left0 = sl - hl;
    const uint64_t hmask = (1ULL << (2 * hl)) - 1ULL;
    const uint32_t nw = a.wstride < (size_t)VT_W ? (uint32_t)a.wstride : (uint32_t)VT_W;
    unsigned long long st_seeds = 0, st_steps = 0, st_hits = 0;
    unsigned int st_fm = 0, st_txt = 0, st_deferred = 0;

    bool have = false, drained = false;
    uint32_t read = 0, num_seeds = 0, span_q = 0, span_r = 0, n_outer = 0;
    int iter = 0;
    uint64_t okey[VT_NO];
    uint32_t oval[VT_NO];
    unsigned long long r_seeds = 0, r_steps = 0;           /* counters of the read in hand: dropped if it is deferred */
    unsigned int r_hits = 0, r_fm = 0, r_txt = 0;
#pragma unroll
    for (int e = 0; e < VT_NO; e++) { okey[e] = 0; oval[e] = 0; }

    for (;;) {
        /* ---- lanes without a read take the next ones: one atomic per warp and trip ---- */
        const uint32_t need = __ballot_sync(FULL, !have && !drained);
        if (need) {
            const int leader = __ffs((int)need) - 1;
            uint32_t base = 0;
            if ((int)lane == leader) base = atomicAdd(a.work_counter, (uint32_t)__popc(need));
            base = __shfl_sync(FULL, base, leader);
            if (!have && !drained) {
                const uint32_t widx = base + (uint32_t)__popc(need & ((1u << lane) - 1u));
                if (widx < (uint32_t)a.n) {
                    read = widx;
                    const uint32_t len = a.lens[read];
                    const uint32_t *src = a.pack + (size_t)read * a.wstride;
                    for (uint32_t w = 0; w < nw; w++) rd[w] = __ldg(src + w);
                    num_seeds = len / (uint32_t)P;                               /* alnmain.c:371 */
                    const uint32_t span = len > (uint32_t)sl ? len - (uint32_t)sl : 0;   /* seeds start at j < len - sl (:353) */
                    span_q = span / (uint32_t)P;
                    span_r = span - span_q * (uint32_t)P;
                    n_outer = 0;
                    iter = 0;
                    r_seeds = r_steps = 0;
                    r_hits = r_fm = r_txt = 0;
                    have = true;
                } else {
                    drained = true;
                }
            }
        }
        if (!__any_sync(FULL, have)) break;

        /* ---- one phase of this lane's read: seeds iter, iter+P, ... (alnmain.c:353) ---- */
        const uint32_t m_count = have ? span_q + ((uint32_t)iter < span_r ? 1u : 0u) : 0u;     /* <= VT_M (launcher) */
        unsigned long long ent[VT_M];
        uint32_t sd[VT_M];                                   /* low 16 bases of the seeds: what a table entry's context is compared with */
#pragma unroll
        for (int m = 0; m < VT_M; m++) {
            ent[m] = 0;
            sd[m] = 0;
            if (a.use2 && (uint32_t)m < m_count) {          /* without the deep table vote_resolve_seed reads the 12-mer table */
                const uint64_t seed = vt_bases32(rd, (uint32_t)iter + (uint32_t)m * (uint32_t)P);
                ent[m] = __ldg(ix.lc2 + ((seed >> (2 * left0)) & hmask));
                sd[m] = (uint32_t)seed;
            }
        }
        uint64_t vb[VT_NV];
        uint32_t vc[VT_NV];                  /* count << 8 | first vote (2 * seed + row) << 4 | smallest low nibble of the keys */
        uint32_t nv = 0;
        bool defer = false;
        /* histo_add (histo.c:42-56) into the phase's list; `ord` = 2 * seed index + row: the list's first-seen order */
        auto cast = [&](uint64_t key, uint32_t ord) {
            const uint64_t b = key >> 4;
            const uint32_t lw = (uint32_t)key & 15u;
            bool found = false;
#pragma unroll
            for (int i = 0; i < VT_NV; i++) {
                if ((uint32_t)i < nv && vb[i] == b) {
                    const uint32_t c = vc[i];
                    vc[i] = ((c & ~0xFFu) + 0x100u) | min(c & 0xF0u, ord << 4) | min(c & 15u, lw);
                    found = true;
                }
            }
            if (!found) {
                if (nv == VT_NV) {
                    defer = true;
                } else {
#pragma unroll
                    for (int i = 0; i < VT_NV; i++) {
                        if ((uint32_t)i == nv) {
                            vb[i] = b;
                            vc[i] = 0x100u | (ord << 4) | lw;
                        }
                    }
                    nv++;
                }
            }
        };
        /* pass A, every seed: an absent entry, or a one-row entry whose context covers the rest of the seed, is finished on
         * the spot (as_lc2.cu tag 1: the entry holds the hit's position and the 11 bases in front of it); the others are
         * marked for pass B.  Same counters as vote_resolve_seed keeps for these two cases. */
        uint32_t todo = 0;
        const bool fastpath = a.use2 && left0 <= 11 && a.thres > 1u;
#pragma unroll
        for (int m = 0; m < VT_M; m++) {
            if ((uint32_t)m < m_count) {
                const unsigned long long e = ent[m];
                const uint32_t info = (uint32_t)(e >> 40);
                if (fastpath && e == 0) {
                    r_seeds++;
                } else if (fastpath && (info >> 22) == 1u) {
                    r_seeds++;
                    r_steps += (unsigned)(ix.hlen2 - ix.hlen);
                    uint32_t x = 0;
                    if (left0 > 0) {
                        x = (sd[m] ^ ((info & 0x3FFFFFu) >> (2 * (11 - left0)))) & ((1u << (2 * left0)) - 1u);
                        r_steps += (unsigned)(x == 0 ? left0 : left0 - ((31 - __clz(x)) >> 1));
                        r_txt++;
                    }
                    if (x == 0) {
                        const uint32_t j = (uint32_t)iter + (uint32_t)m * (uint32_t)P;
                        r_hits++;
                        cast((e & 0xFFFFFFFFFFULL) - (uint64_t)left0 - (uint64_t)j, 2u * (uint32_t)m);    /* alnmain.c:363-365 */
                    }
                } else {
                    todo |= 1u << m;
                }
            }
        }
        /* pass B, the marked seeds of each lane in turn (backward-search steps, two-row intervals, text gathers): the
         * lines they start from are requested for all of them first, so the resolver finds them in L1 */
        if (a.use2) {
#pragma unroll
            for (int m = 0; m < VT_M; m++) {
                const unsigned long long e = ent[m];
                const uint32_t info = (uint32_t)(e >> 40), etag = info >> 22;
                if (((todo >> m) & 1u) && e != 0 && etag != 1u && info != 0x3FFFFFu) {
                    const uint64_t k0 = e & 0xFFFFFFFFFFULL;
                    const uint32_t cnt = etag ? etag : (info & 0x3FFFFFu);
                    if (cnt <= (uint32_t)VOTE_VW) {
                        /* a two-row entry that decides from its own context (use_ectx) fetches a position for survivors only */
                        if (etag == 2u && a.use_ectx && left0 >= 1 && left0 <= 5) continue;
                        if (ix.sax) asm volatile("prefetch.global.L1 [%0];" ::"l"(ix.sax + k0));
                        else if (ix.sa_lo) asm volatile("prefetch.global.L1 [%0];" ::"l"(ix.sa_lo + k0));
                    } else {
                        asm volatile("prefetch.global.L1 [%0];" ::"l"(ix.fm + ((k0 - 1) >> 6)));
                        asm volatile("prefetch.global.L1 [%0];" ::"l"(ix.fm + ((k0 + cnt - 1) >> 6)));
                    }
                }
            }
        }
        while (__any_sync(FULL, todo != 0)) {
            if (todo) {
                const uint32_t m = (uint32_t)__ffs((int)todo) - 1u;
                todo &= todo - 1u;
                unsigned long long em = ent[0];
#pragma unroll
                for (int q = 1; q < VT_M; q++)
                    if (m == (uint32_t)q) em = ent[q];
                const uint32_t j = (uint32_t)iter + m * (uint32_t)P;
                const uint64_t seed = vt_bases32(rd, j);
                uint64_t kk = 1, ll = ix.N - 1, rr = 0, hit_pos = 0;             /* alnmain.c:354 */
                bool have_pos = false;
                vote_resolve_seed<SMALL>(ix, a, s_tab, s_tabChi, sl, left0, hmask, seed, a.use2 != 0, em, kk, ll, rr, have_pos, hit_pos,
                                         r_seeds, r_steps, r_fm, r_txt);
                const uint32_t rr_eff = (rr > 0 && rr < a.thres) ? (uint32_t)rr : 0u;   /* alnmain.c:360 */
                if (rr_eff > 2u) defer = true;
                for (uint32_t r = 0; r < rr_eff && r < 2u; r++) {
                    r_hits++;
                    cast((have_pos ? hit_pos : as_locate(ix, kk + r)) - (uint64_t)j, 2u * m + r);          /* alnmain.c:363-365 */
                }
            }
        }
        /* ---- TOP2 of the phase (histo.c:84-96): largest count, first seen wins a tie ---- */
        uint32_t t0 = 0, t1 = 0;
        uint64_t c0key = 0;
#pragma unroll
        for (int i = 0; i < VT_NV; i++) {
            if ((uint32_t)i < nv) {
                const uint32_t c = ((vc[i] >> 8) << 4) | (15u - ((vc[i] >> 4) & 15u));
                if (c > t0) {
                    t1 = t0;
                    t0 = c;
                    c0key = (vb[i] << 4) | (vc[i] & 15u);
                } else if (c > t1) {
                    t1 = c;
                }
            }
        }
        const uint32_t v0val = t0 >> 4, v1val = t1 >> 4;
        /* ---- phase control (alnmain.c:371-404), as in the group kernel ---- */
        bool broke = false;
        uint64_t best_key = 0, best_val = 0;
        if (have && !defer && num_seeds > 0) {                                   /* alnmain.c:373 */
            const uint64_t v = (uint64_t)v0val + v1val;
            if (5ull * v > 3ull * (uint64_t)num_seeds) {                        /* score > 0.6 in integers, see k_vote */
                best_key = c0key;
                best_val = v0val;
                broke = true;
            } else if (v0val != 0) {                                            /* :386-388 histo_add(outer, cand[0].key) */
                const uint64_t b = c0key >> 4;
                bool found = false;
#pragma unroll
                for (int e = 0; e < VT_NO; e++) {
                    if ((uint32_t)e < n_outer && (okey[e] >> 4) == b) {
                        oval[e] += 1;
                        if (c0key < okey[e]) okey[e] = c0key;
                        found = true;
                    }
                }
                if (!found) {
                    if (n_outer == VT_NO) {
                        defer = true;
                    } else {
#pragma unroll
                        for (int e = 0; e < VT_NO; e++) {
                            if ((uint32_t)e == n_outer) {
                                okey[e] = c0key;
                                oval[e] = 1;
                            }
                        }
                        n_outer++;
                    }
                }
            }
        }
        if (have) {
            bool finished;
            if (defer) {
                a.dlist[atomicAdd(a.dlist_count, 1u)] = read;                    /* redone by the group kernel */
                st_deferred++;
                finished = true;
            } else {
                if (!broke) iter++;                                            /* a break leaves iter where it was */
                finished = broke || iter >= P;
                if (finished) {
                    int from_outer = 0;
                    if (iter >= P - 1) {                                        /* alnmain.c:400-403 */
                        uint64_t bk = 0;
                        uint32_t bv = 0;
#pragma unroll
                        for (int e = 0; e < VT_NO; e++) {
                            if ((uint32_t)e < n_outer && oval[e] > bv) {
                                bv = oval[e];
                                bk = okey[e];
                            }
                        }
                        best_key = bk;
                        best_val = bv;
                        from_outer = 1;
                    }
                    VoteOut o;
                    o.best_key = best_key;
                    o.best_val = best_val;
                    o.phases = iter;
                    o.from_outer = from_outer;
                    a.out[read] = o;
                    st_seeds += r_seeds;
                    st_steps += r_steps;
                    st_hits += r_hits;
                    st_fm += r_fm;
                    st_txt += r_txt;
                }
            }
            if (finished) have = false;
        }
    }
#pragma unroll
    for (int d = 16; d >= 1; d >>= 1) {
        st_seeds += __shfl_xor_sync(FULL, st_seeds, d);
        st_steps += __shfl_xor_sync(FULL, st_steps, d);
        st_hits += __shfl_xor_sync(FULL, st_hits, d);
        st_fm += __shfl_xor_sync(FULL, st_fm, d);
        st_txt += __shfl_xor_sync(FULL, st_txt, d);
        st_deferred += __shfl_xor_sync(FULL, st_deferred, d);
    }
    if (lane == 0) {
        if (st_seeds) atomicAdd(a.stats + 6, st_seeds);
        if (st_hits) atomicAdd(a.stats + 7, st_hits);
        if (st_steps) atomicAdd(a.stats + 8, st_steps);
        if (st_fm) atomicAdd(a.stats + 12, (unsigned long long)st_fm);
        if (st_txt) atomicAdd(a.stats + 13, (unsigned long long)st_txt);
        if (st_deferred) atomicAdd(a.stats + 16, (unsigned long long)st_deferred);
    }
}

static uint32_t next_pow2(uint64_t v) {
    uint32_t p = 64;
    while (p < v && p < (1u << 30)) p <<= 1;
    return p;
}

int as_launch_vote(as_gpu *h, const as_params *p, const uint32_t *d_pack, size_t wstride, const uint32_t *d_lens,
                   size_t n, uint32_t max_len, VoteOut *d_vote) {
    if (n == 0) return AS_OK;
    if (p->seed_len < 1 || p->seed_len > 32 || p->thres < 1 || p->thres > 65535) {
        as_set_error("as_map_batch: seed_len must be in [1,32] and thres in [1,65535]");
        return AS_EINVAL;
    }
    if (n >= 0xFFFFFFFFull) return AS_ERANGE;
    const int sl = (int)p->seed_len, P = sl + 1;
    /* counters: [0] work counter, [1] overflow count, [2] work counter of the BIG pass */
    AS_TRY(as_scratch(h, &h->counters, 64));
    AS_TRY(as_scratch(h, &h->ovf_list, n * sizeof(uint32_t)));
    AS_CUDA(cudaMemsetAsync(h->counters.p, 0, 64, h->stream));

    /* BIG table: every vote of one phase fits at load factor <= 1/2 */
    uint64_t seeds_per_phase = max_len > (uint32_t)sl ? ((uint64_t)(max_len - sl) + P - 1) / P : 1;
    uint64_t max_votes = seeds_per_phase * (uint64_t)(p->thres > 1 ? p->thres - 1 : 1);
    /* the vote order key (seed index in the phase) * thres + (row offset) is a 32-bit word, and the overflow table holds
     * at most 2^30 slots: batches beyond either are refused instead of silently breaking the TOP2 tie order */
    if (seeds_per_phase * (uint64_t)p->thres >= (1ULL << 32) || 2 * max_votes > (1ULL << 30)) {
        as_set_error("as_map_batch: a read of %u bases with thres %u needs more vote slots than the overflow table has", max_len, p->thres);
        return AS_ERANGE;
    }
    uint32_t big_cap = next_pow2(max_votes + max_votes / 4);            /* load factor <= 0.8 */
    size_t big_stride = ((size_t)big_cap * 24 + 64 + 255) & ~(size_t)255;
    size_t budget = (size_t)2 << 30;
    size_t big_warps = budget / big_stride;
    /* one warp per queued read.  The pass is latency-bound (a located row and a compare-and-swap per vote), so its time
     * falls with the warps in flight -- as long as their tables stay in L2: 16 warps per SM on 8192-slot tables (465 MB) ran
     * SLOWER than 4 (c2r: 38 vs 29 ms).  Short reads: up to 8 warps per SM within 96 MB of tables; long reads (tables of
     * megabytes each): 4 per SM within the 2 GB budget, as before */
    {
        const size_t l2_warps = ((size_t)96 << 20) / big_stride;
        size_t cap = (size_t)h->sm_count * 4;
        if (l2_warps >= (size_t)h->sm_count) cap = l2_warps < (size_t)h->sm_count * 8 ? l2_warps : (size_t)h->sm_count * 8;
        if (big_warps > cap) big_warps = cap;
    }
    if (big_warps > n) big_warps = n;
    if (big_warps < 1) big_warps = 1;
    AS_TRY(as_scratch(h, &h->ovf_table, big_warps * big_stride));

    VoteArgs a;
    a.pack = d_pack;
    a.wstride = wstride;
    a.lens = d_lens;
    a.n = n;
    a.sl = sl;
    a.thres = p->thres;
    a.out = d_vote;
    a.work_counter = (unsigned int *)h->counters.p;
    a.ovf_count = (unsigned int *)h->counters.p + 1;
    a.ovf_list = (uint32_t *)h->ovf_list.p;
    a.big_mem = (unsigned char *)h->ovf_table.p;
    a.big_stride = big_stride;
    a.big_cap = big_cap;
    a.stats = h->d_stats;
    a.list = nullptr;
    a.list_count = nullptr;
    a.dlist = nullptr;
    a.dlist_count = nullptr;
    a.use2 = (h->ix.lc2 && sl >= h->ix.hlen2 && !(p->flags & AS_F_NO_LC2)) ? 1 : 0;
    /* Deciding a seed from its table entry saves the SA/context gather of every seed that dies there, at the price of
     * fetching a survivor's position after the compare instead of along with the context.  It pays when that gather is
     * a DRAM sector (c3: PART 1 7.4 -> 6.5 ms, c4: 33.2 -> 31.7), not when the SA table sits in L2 (c2: 2.35 -> 2.55) */
    a.use_ectx = h->ix.N * 8ULL > 2ULL * (unsigned long long)h->l2_bytes ? 1 : 0;
    if (p->vote_ectx >= 0) a.use_ectx = p->vote_ectx != 0;

    /* indexes of one super block (<= 2^30 rows) take the 32-bit backward-search path */
    const bool small = h->ix.N <= (1ULL << AS_SB_SHIFT) && !(p->flags & AS_F_WIDE_INDEX);
    /* lanes per read from the seeds of one phase of the longest read */
    int G = seeds_per_phase <= 8 ? 8 : seeds_per_phase <= 16 ? 16 : 32;
    uint64_t groups_needed = n;
    uint64_t groups_per_block = VOTE_THREADS / G;
    uint64_t blocks = (groups_needed + groups_per_block - 1) / groups_per_block;
    unsigned grid = 0;
    /* Short reads: the thread-per-read kernel takes the batch; what it defers comes back as a list for the group kernel.
     * Applies when every phase has at most VT_M seeds, the words a seed can touch fit the shared-memory row, and the table
     * (deep or 12-mer) is not longer than the seed (counters: [3] work counter of the list pass, [4] length of the list) */
    const bool thread_ok = sl >= (a.use2 ? h->ix.hlen2 : h->ix.hlen) && seeds_per_phase <= VT_M && sl < 32 && max_len > (uint32_t)sl &&
                           ((max_len - (uint32_t)sl - 1u) >> 4) + 3u <= (uint32_t)VT_W &&
                           /* default: indexes of one super block.  At human scale PART 1 runs at the DRAM's random-gather rate
                            * in either kernel (c4: 33.7 ms here, 32.6 ms in the group kernel) */
                           (p->vote_thread > 0 || (p->vote_thread < 0 && small));
    if (thread_ok) {
        AS_TRY(as_scratch(h, &h->vote_dlist, n * sizeof(uint32_t)));
        a.dlist = (uint32_t *)h->vote_dlist.p;
        a.dlist_count = (unsigned int *)h->counters.p + 4;
        int occ = 0;
        uint64_t tb = (n + VT_THREADS - 1) / VT_THREADS;
        if (small) {
            if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_vote_t<true>, VT_THREADS, 0) != cudaSuccess || occ < 1) occ = VT_MIN_BLOCKS;
            if (p->vote_blocks_per_sm > 0 && p->vote_blocks_per_sm < occ) occ = p->vote_blocks_per_sm;
            const uint64_t cap = (uint64_t)h->sm_count * (uint64_t)occ;
            k_vote_t<true><<<(unsigned)(tb < cap ? tb : cap), VT_THREADS, 0, h->stream>>>(h->ix, a);
        } else {
            if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_vote_t<false>, VT_THREADS, 0) != cudaSuccess || occ < 1) occ = VT_MIN_BLOCKS;
            if (p->vote_blocks_per_sm > 0 && p->vote_blocks_per_sm < occ) occ = p->vote_blocks_per_sm;
            const uint64_t cap = (uint64_t)h->sm_count * (uint64_t)occ;
            k_vote_t<false><<<(unsigned)(tb < cap ? tb : cap), VT_THREADS, 0, h->stream>>>(h->ix, a);
        }
        h->launches++;
        AS_CUDA(cudaGetLastError());
        a.list = a.dlist;
        a.list_count = a.dlist_count;
        a.work_counter = (unsigned int *)h->counters.p + 3;
    }
    /* reads are claimed from a work counter, so the grid is exactly the resident blocks (registers / shared memory) */
#define VOTE_LAUNCH1(GG, SM)                                                                                     \
    do {                                                                                                         \
        int occ = 0;                                                                                             \
        if (VOTE_CARVEOUT >= 0)                                                                                  \
            cudaFuncSetAttribute(k_vote<GG, false, SM>, cudaFuncAttributePreferredSharedMemoryCarveout, VOTE_CARVEOUT); \
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_vote<GG, false, SM>, VOTE_THREADS, 0) != cudaSuccess || \
            occ < 1)                                                                                             \
            occ = VOTE_MIN_BLOCKS;                                                                               \
        const uint64_t cap = (uint64_t)h->sm_count * (uint64_t)occ;                                              \
        grid = (unsigned)(blocks < cap ? blocks : cap);                                                          \
        k_vote<GG, false, SM><<<grid, VOTE_THREADS, 0, h->stream>>>(h->ix, a);                                   \
    } while (0)
#define VOTE_LAUNCH(GG)                 \
    do {                                \
        if (small) VOTE_LAUNCH1(GG, true); \
        else VOTE_LAUNCH1(GG, false);   \
    } while (0)
    if (G == 8) VOTE_LAUNCH(8);
    else if (G == 16) VOTE_LAUNCH(16);
    else VOTE_LAUNCH(32);
#undef VOTE_LAUNCH
#undef VOTE_LAUNCH1
    h->launches++;
    AS_CUDA(cudaGetLastError());
    /* overflow pass: exits at once when the queue is empty */
    a.work_counter = (unsigned int *)h->counters.p + 2;
    if (small) k_vote<32, true, true><<<(unsigned)big_warps, 32, 0, h->stream>>>(h->ix, a);
    else k_vote<32, true, false><<<(unsigned)big_warps, 32, 0, h->stream>>>(h->ix, a);
    h->launches++;
    AS_CUDA(cudaGetLastError());
    return AS_OK;
}
