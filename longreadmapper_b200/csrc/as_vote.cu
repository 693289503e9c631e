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
/* A/B switches (tools/ab_vote.sh): VOTE_PF = while a one-round phase is resolved, the deep-table entries of the read's
 * NEXT phase are requested into L2 (the next phase's seed is this one moved by a base, so its table index comes out of
 * the register that holds this seed); VOTE_L2HINT = deep-table loads carry an evict-first L2 policy and the SA + context
 * loads an evict-last one, so the table stream does not wash the small tables out of L2 */
#ifndef VOTE_PF
#define VOTE_PF 1
#endif
#ifndef VOTE_L2HINT
#define VOTE_L2HINT 0
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
    int use2;                    /* seeds start from the deep table (DevIndex.lc2) instead of the 12-mer table */
    int use_ectx;                /* decide one- and two-row entries from the context bases they carry (as_lc2.cu) */
};

__device__ __forceinline__ unsigned long long vote_ld_hint(const unsigned long long *p, unsigned long long pol) {
    unsigned long long v;
    asm volatile("ld.global.nc.L2::cache_hint.b64 %0, [%1], %2;" : "=l"(v) : "l"(p), "l"(pol));
    return v;
}

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
#if VOTE_L2HINT
    unsigned long long pol_first, pol_last;
    asm("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol_first));
    asm("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol_last));
#define VOTE_LD_LC2(p) vote_ld_hint(p, pol_first)
#define VOTE_LD_SAX(p) vote_ld_hint(p, pol_last)
#else
#define VOTE_LD_LC2(p) __ldg(p)
#define VOTE_LD_SAX(p) __ldg(p)
#endif
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
            const uint32_t limit = BIG ? *a.ovf_count : (uint32_t)a.n;
            if (widx < limit) {
                read = BIG ? a.ovf_list[widx] : widx;
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
#if VOTE_PF
        /* next phase = same seeds one base on (j + 1): sl + 1 <= 32 bases are in the seed register.  Wasted only for the
         * phase that ends the read (one in ~6 on 150-bp reads at 2 %) */
        const bool pf_next = m_max <= (uint32_t)G && sl < 32 && iter + 1 < P;
#endif
        uint32_t v0val = 0, v1val = 0;
        uint64_t c0key = 0;
        for (uint32_t c0 = 0; c0 < m_max; c0 += G) {
            const uint32_t m = c0 + g;
            const uint32_t j = (uint32_t)iter + m * (uint32_t)P;
            uint64_t kk = 1, ll = ix.N - 1, rr = 0;                             /* alnmain.c:354 */
            bool have_pos = false;                                               /* hit_pos = SA of the seed's one row */
            uint64_t hit_pos = 0;
            if (m < m_count) {
                const uint64_t seed = as_bases32(pr, j);
                bool absent = false;
                uint32_t etag = 0, ectx = 0;
                int tt = left0 - 1;                                             /* next seed base to walk */
                if (a.use2) {
                    /* deep table: the interval lc_access + (hlen2 - hlen) steps of fmi_aln would have reached */
                    const unsigned long long e = VOTE_LD_LC2(ix.lc2 + ((seed >> (2 * left0)) & hmask));
#if VOTE_PF
                    if (pf_next) asm volatile("prefetch.global.L2 [%0];" ::"l"(ix.lc2 + ((seed >> (2 * left0 + 2)) & hmask)));
#endif
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
                                const unsigned long long e = VOTE_LD_SAX(ix.sax + kk + (uint64_t)r);
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
                for (uint32_t v0 = 0; v0 < V_max; v0 += G) {
                    const uint32_t v = v0 + g;
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
                    if (v < V) {
                        const uint32_t om = c0 + lo;
                        const uint32_t oj = (uint32_t)iter + om * (uint32_t)P;
                        const uint32_t off = v - ob;
                        uint64_t row = (((uint64_t)okh << 32) | okl) + off;
                        uint64_t key = (opk ? (((uint64_t)oph << 32) | opl) : as_locate(ix, row)) - (uint64_t)oj;
                        st_hits++;
                        if (!table_insert<USED_T>(t, key >> 4, (uint32_t)key & 15u, om * a.thres + off)) ok = false;
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
    uint32_t big_cap = next_pow2(2 * max_votes);
    size_t big_stride = ((size_t)big_cap * 24 + 64 + 255) & ~(size_t)255;
    size_t budget = (size_t)2 << 30;
    size_t big_warps = budget / big_stride;
    if (big_warps > (size_t)h->sm_count * 4) big_warps = (size_t)h->sm_count * 4;
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
    a.use2 = (h->ix.lc2 && sl >= h->ix.hlen2 && !(p->flags & AS_F_NO_LC2)) ? 1 : 0;
    /* Deciding a seed from its table entry saves the SA/context gather of every seed that dies there, at the price of
     * fetching a survivor's position after the compare instead of along with the context.  It pays when that gather is
     * a DRAM sector (c3: PART 1 7.4 -> 6.5 ms, c4: 33.2 -> 31.7), not when the SA table sits in L2 (c2: 2.35 -> 2.55) */
    a.use_ectx = h->ix.N * 8ULL > 2ULL * (unsigned long long)h->l2_bytes ? 1 : 0;
    if (p->vote_ectx >= 0) a.use_ectx = p->vote_ectx != 0;

    /* lanes per read from the seeds of one phase of the longest read */
    int G = seeds_per_phase <= 8 ? 8 : seeds_per_phase <= 16 ? 16 : 32;
    uint64_t groups_needed = n;
    uint64_t groups_per_block = VOTE_THREADS / G;
    uint64_t blocks = (groups_needed + groups_per_block - 1) / groups_per_block;
    unsigned grid = 0;
    /* indexes of one super block (<= 2^30 rows) take the 32-bit backward-search path */
    const bool small = h->ix.N <= (1ULL << AS_SB_SHIFT) && !(p->flags & AS_F_WIDE_INDEX);
    /* reads are claimed from a work counter, so the grid is exactly the resident blocks (registers / shared memory) */
#define VOTE_LAUNCH1(GG, SM)                                                                                     \
    do {                                                                                                         \
        int occ = 0;                                                                                             \
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
