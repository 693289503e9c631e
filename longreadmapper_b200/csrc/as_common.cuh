/*
 * as_common.cuh -- internal declarations shared by the translation units of libaccseq_b200.so.
 * sm_100a only.  Nothing here is part of the public C ABI (include/accseq_b200.h).
 */
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>
#include <stdio.h>
#include <string.h>

#include "accseq_b200.h"

/* ------------------------------------------------------------------------------------------------
 * HBM layout of the index (DESIGN.md "HBM layout")
 * ---------------------------------------------------------------------------------------------- */

/* One 32-byte sector per 64 BWT rows: occurrence counts at the block start (relative to the enclosing
 * super block) and the 64 rows as two bit planes.  One occ() = one sector.  Replaces the reference's
 * O table (4 x u64 per 32 rows) + byte-per-row BWT (fmidx.h:16-21; occ at fmidx.c:277-293). */
struct __align__(32) FmBlock {
    uint32_t cnt[4];   /* # of A,C,G,T in rows [sb_start, 64*blk) */
    uint64_t lo;       /* bit r = low bit of the code of row 64*blk + r */
    uint64_t hi;       /* bit r = high bit                                 */
};
#define AS_MAX_CHUNKS 32
#define AS_SB_SHIFT 30 /* super block = 2^30 rows; absolute counts as u64[4] per super block */

struct DevIndex {
    uint64_t N;                 /* rows = .cat length (fmi->length) */
    uint64_t primary;           /* row whose BWT char is '$' (coded as A in the planes) */
    uint64_t C[4];              /* dna_fmi.c['A'],['C'],['G'],['T'] */
    const FmBlock *fm;
    const uint64_t *super;      /* [n_super][4] */
    const ulonglong2 *lc;       /* (k,l) per 12-mer, indexed by the LSB-first code of the k-mer */
    int hlen;
    const unsigned long long *lc2;   /* deep table (as_lc2.cu): row | rows << 40 per hlen2-mer, LSB-first code; may be NULL */
    int hlen2;
    const unsigned long long *sax;   /* SA[row] | 12 preceding text bases << 40 (as_lc2.cu); may be NULL */
    const uint32_t *sa_lo;      /* low 32 bits of SA[row]; NULL on a handle that keeps only the sampled SA */
    const uint8_t *sa_hi;       /* bits 32..39, NULL when N <= 2^32 */
    const uint32_t *csa_lo;     /* sampled SA: SA[csa_ratio * i] (fmidx.c:152-163), low 32 bits; NULL when not sampled */
    const uint8_t *csa_hi;
    uint32_t csa_ratio;
    const uint32_t *text2;      /* .cat text, 2 bits per base, 16 bases per word, LSB first */
    const as_mta *mta;
    int mta_len;
};

struct ScratchBuf {
    void *p = nullptr;
    size_t cap = 0;
};

struct as_gpu {
    int device = 0;
    cudaStream_t own_stream = nullptr;
    cudaStream_t stream = nullptr;
    DevIndex ix{};
    as_open_opts opts{};        /* as_gpu_open_opts / as_gpu_build_opts */
    /* owned device allocations of the index */
    void *d_lc2 = nullptr, *d_sax = nullptr;
    void *d_csa_lo = nullptr, *d_csa_hi = nullptr;   /* sampled SA of as_csa_access_batch: SA[csa_ratio * i], 32 + 8 bits */
    int csa_ratio = 0;
    void *d_fm = nullptr, *d_super = nullptr, *d_lc = nullptr, *d_sa_lo = nullptr, *d_sa_hi = nullptr,
         *d_text2 = nullptr, *d_mta = nullptr;
    /* device scratch, grown on demand */
    ScratchBuf reads, lens, results, store, qpack, tpack, toff, vote, counters, ovf_list, ovf_table, misc[6];
    ScratchBuf gact_planes, gact_tb;     /* bit-parallel GACT: traceback planes / packed tile ops, per resident warp */
    ScratchBuf gact_order;               /* bit-parallel GACT: reads by decreasing length (multi-tile batches) */
    bool gact_emit_runs = false;         /* run_stages -> as_launch_gact: the caller only reads run bytes (AS_F_RLE_CIGAR, dense gather) */
    ScratchBuf vote_dlist;               /* PART 1: reads the thread-per-read kernel deferred to the group kernel */
    ScratchBuf gact_retry;               /* bit-parallel GACT: one-tile reads the one-word diagonal band could not prove exact */
    ScratchBuf path_planes;              /* as_edit_alignment_batch: Pv / Ph words of every (column, block) of a slice of pairs */
    ScratchBuf locus_tmp;                /* packed-only long reads: one scratch row per warp for the reverse complement */
    ScratchBuf pack_in, cres, runs, run_tot;   /* as_map_batch_packed: tight packed rows in, compact records + dense run bytes out */
    /* pinned staging for small D2H/H2D */
    void *h_pin = nullptr;
    size_t h_pin_cap = 0;
    /* stats: device accumulators (as_stats layout, u64 x 10) */
    unsigned long long *d_stats = nullptr;
    cudaEvent_t ev[10]{};
    float last_ms[8]{};
    double dbg_us[4]{};
    double last_host_us[3]{};   /* as_map_batch_packed: host time enqueueing / waiting / reading events */
    /* chunked pipeline of as_map_batch: H2D of chunk i+1 | kernels of chunk i | D2H of chunk i-1 */
    cudaStream_t s_h2d = nullptr, s_d2h = nullptr, s_d2h2 = nullptr;
    cudaEvent_t cev[AS_MAX_CHUNKS][8]{};   /* per chunk: [0..4] stage marks, [5] H2D done, [6] kernels done, [7] fixed D2H done */
    cudaEvent_t ev_d2h0 = nullptr;         /* first D2H issued */
    bool cev_ready = false;
    int n_chunks_last = 0;
    size_t reads_done_last = 0, reads_touched_last = 0;     /* as_map_batch: leading reads whose copies back to the caller were all / partly enqueued */
    uint64_t last_h2d_bytes = 0, last_d2h_bytes = 0;   /* bytes the last as_map_batch moved over PCIe */
    bool timing_pending = false;           /* as_map_batch_device enqueued a call whose events were not read yet */
    int *d_chunk_max = nullptr;            /* per chunk: max n_cigar_op (trims the CIGAR-column D2H) */
    int *h_chunk_max = nullptr;            /* pinned mirror */
    unsigned long long *h_run_tot = nullptr;   /* pinned: run bytes per chunk (as_map_batch_packed) */
    unsigned int *h_band = nullptr, *d_band = nullptr;   /* adaptive GACT band: mapped pinned word + its device alias */
    int sm_count = 148;
    size_t l2_bytes = 126u << 20;
    uint64_t launches = 0;      /* kernels launched by this handle since open */
    /* second lane of as_map_batch_packed: a handle that shares this one's index (ix is a struct of device pointers) but owns
     * its own stream and scratch, so that consecutive chunks run on two streams and the tail of one chunk's kernels fills
     * with the next chunk's blocks */
    as_gpu *twin = nullptr;
    bool is_twin = false;
};

/* error plumbing */
void as_set_error(const char *fmt, ...);
#define AS_CUDA(call)                                                                               \
    do {                                                                                            \
        cudaError_t e_ = (call);                                                                    \
        if (e_ != cudaSuccess) {                                                                    \
            as_set_error("%s:%d: %s failed: %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); \
            return AS_ECUDA;                                                                        \
        }                                                                                           \
    } while (0)
#define AS_TRY(call)              \
    do {                          \
        int r_ = (call);          \
        if (r_ != AS_OK) return r_; \
    } while (0)

as_gpu *as_handle_new(int device);                        /* as_index.cu: device checks + stream/events/stats */
int as_build_big(as_gpu *h, const char *content, uint64_t N, const as_mta *mta, int mta_len, int hlen,
                 uint64_t part_max);                      /* as_build_big.cu: partitioned suffix sort, 40-bit rows */
int as_build_lc2(as_gpu *h);                              /* as_lc2.cu: deep k-mer table from the device index */
int as_csa_sample(as_gpu *h, int csa_ratio);              /* as_lookup.cu: SA[csa_ratio * i] into the handle */
int as_launch_split_sa(as_gpu *h, const uint8_t *d_in, int entry_bytes, uint64_t n, uint32_t *d_lo, uint8_t *d_hi);
int as_launch_pack_text(as_gpu *h, const uint8_t *d_in, uint64_t n, uint32_t *d_out);   /* ASCII text -> 2-bit words */
int as_scratch(as_gpu *h, ScratchBuf *b, size_t bytes);   /* ensure capacity (device memory) */
int as_pinned(as_gpu *h, size_t bytes);                   /* ensure h->h_pin capacity */

/* stage launchers (as_lookup.cu / as_vote.cu / as_align.cu); all asynchronous on h->stream */
struct VoteOut {
    uint64_t best_key, best_val;
    int32_t phases, from_outer;
};
int as_launch_pack(as_gpu *h, const char *d_ascii, size_t stride, const uint32_t *d_lens, size_t n,
                   uint32_t *d_pack, size_t wstride);
int as_launch_myers_path(as_gpu *h, const uint32_t *d_qpack, size_t q_wstride, const uint32_t *d_qlens, const uint32_t *d_tbase,
                         const uint64_t *d_toff, const uint32_t *d_tlens, size_t n, uint32_t max_qlen, uint32_t max_tlen,
                         const int32_t *d_ks, int32_t *d_dist, uint8_t *d_ops, size_t ops_stride, int32_t *d_nops);
int as_launch_vote(as_gpu *h, const as_params *p, const uint32_t *d_pack, size_t wstride, const uint32_t *d_lens,
                   size_t n, uint32_t max_len, VoteOut *d_vote);
int as_launch_locus(as_gpu *h, char *d_ascii, size_t stride, const uint32_t *d_lens, size_t n,
                    uint32_t *d_pack, size_t wstride, const VoteOut *d_vote, as_result *d_res, uint64_t *d_toff);
int as_launch_gact(as_gpu *h, const as_params *p, const uint32_t *d_qpack, size_t q_wstride, const uint32_t *d_qlens,
                   const uint32_t *d_tbase, const uint64_t *d_toff, const uint32_t *d_tlens, size_t n,
                   uint32_t max_qlen, uint32_t max_tlen, uint8_t *d_ops, size_t ops_stride, int32_t *d_nops,
                   int32_t *d_score, as_result *d_res);
#define AS_TOFF_NONE 0xFFFFFFFFFFFFFFFFULL
struct GactArgs {
    const uint32_t *qpack;
    size_t q_wstride;
    const uint32_t *qlens;
    const uint32_t *tbase;
    const uint64_t *toff;        /* per read base offset into tbase, AS_TOFF_NONE = skip */
    const uint32_t *tlens;       /* NULL: tlen = qlen (alnmain.c:440-446) */
    size_t n;
    int T, O;
    int TS;                      /* columns kept per lane in shared memory: even, >= min(T, longest target) */
    int TQ;                      /* min(T, longest query) */
    uint8_t *ops;
    size_t ops_stride;
    int ops_zeroed;              /* the op rows were cleared before the launch: single-tile reads only write their non-'=' ops */
    int32_t *n_ops, *score;      /* pair mode outputs (may be NULL) */
    as_result *res;              /* map mode output (may be NULL) */
    unsigned int *work_counter;
    unsigned long long *stats;
    /* bit-parallel kernel with a banded first attempt per tile (as_gact_bp.cu) */
    int band;                    /* < 0: full matrix only; 0: adaptive; > 0: fixed half-width (rounded up to a level) */
    const uint32_t *order;       /* slot -> read, longest first (NULL: identity); lanes of a warp then finish together */
    int order_mode;              /* as_params.gact_order: -1 auto, 0 off, 1 on */
    int blocks_per_sm;           /* as_params.gact_blocks_per_sm: 0 = occupancy query */
    unsigned int *fail_count;    /* policy counters of the launch: warp-tiles refilled, warp-tiles, lane-tiles served by
                                  * half the band, lane-tiles */
    /* one-tile reads, lowest band level: a read whose one-word diagonal band cannot be proven exact is queued here and
     * aligned by a second launch with wider bands, instead of making its whole warp fill the tile again */
    uint32_t *retry_list;
    unsigned int *retry_count;
    const unsigned int *n_dev;   /* non-NULL: slots [0, *n_dev) of `order` are the launch's reads (the queue above) */
    int emit_runs;               /* one-tile reads of the short-read kernels leave their CIGAR as RUN BYTES (k_rle_ops' format)
                                  * at the start of their row and n_cigar_op = runs | AS_NOPS_RUNS; the rows need no clearing */
    int retry_mode;              /* launches over the queue: 1 = only when it holds at most AS_GACT_RETRY_WARP_MAX reads (the
                                  * warp-per-read kernel: a handful of reads finish in microseconds), 2 = only when more */
};
#define AS_GACT_RETRY_WARP_MAX 1024u
#define AS_NOPS_RUNS 0x40000000      /* as_result.n_cigar_op on the device, between GACT and k_rle_ops: the row already holds run bytes */
#define AS_GACT_BAND0 16

/* bit-parallel GACT (as_gact_bp.cu): one read per lane; `a` is filled by as_launch_gact */
int as_launch_gact_bp(as_gpu *h, GactArgs &a, size_t n, uint32_t max_qlen, uint32_t max_tlen);
int as_gact_order_reserve(as_gpu *h, size_t n);
/* warp-per-read GACT (as_align.cu) over the reads a.order[0 .. *a.n_dev) */
int as_launch_gact_warp_list(as_gpu *h, const GactArgs &a, uint32_t max_qlen);
int as_launch_myers(as_gpu *h, const uint32_t *d_qpack, size_t q_wstride, const uint32_t *d_qlens,
                    const uint32_t *d_tbase, const uint64_t *d_toff, const uint32_t *d_tlens, size_t n,
                    uint32_t max_qlen, const int32_t *d_ks, int32_t *d_dist, as_result *d_res);

/* ------------------------------------------------------------------------------------------------
 * device helpers
 * ---------------------------------------------------------------------------------------------- */
#ifdef __CUDACC__

/* A0 C1 G2 T3 for upper or lower case ASCII (lchash.c:38-45, fmidx.c:278-282 mapper[]) */
__device__ __forceinline__ uint32_t as_base_code(uint32_t c) { return ((c >> 1) & 3u) ^ ((c >> 2) & 1u); }

__device__ __forceinline__ void as_ld_block(const FmBlock *b, uint4 &cnt, uint64_t &lo, uint64_t &hi) {
    /* one 32-byte sector, two 16-byte read-only loads */
    const uint4 *p = reinterpret_cast<const uint4 *>(b);
    cnt = __ldg(p);
    uint4 w = __ldg(p + 1);
    lo = ((uint64_t)w.y << 32) | w.x;
    hi = ((uint64_t)w.w << 32) | w.z;
}

/* occ(c, loc) = #{ rows r <= loc : bwt[r] == c }   (fmidx.c:277-293 _occ_access, inclusive of loc) */
__device__ __forceinline__ uint64_t as_occ(const DevIndex &ix, uint32_t c, uint64_t loc) {
    uint64_t blk = loc >> 6;
    uint32_t r = (uint32_t)loc & 63u;
    uint32_t v8[8];
    asm("ld.global.nc.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
        : "=r"(v8[0]), "=r"(v8[1]), "=r"(v8[2]), "=r"(v8[3]), "=r"(v8[4]), "=r"(v8[5]), "=r"(v8[6]), "=r"(v8[7])
        : "l"(ix.fm + blk));
    uint4 cnt = make_uint4(v8[0], v8[1], v8[2], v8[3]);
    const uint64_t lo = ((uint64_t)v8[5] << 32) | v8[4], hi = ((uint64_t)v8[7] << 32) | v8[6];
    uint64_t m = ((c & 1u) ? lo : ~lo) & ((c & 2u) ? hi : ~hi);
    m &= (r == 63u) ? ~0ULL : ((1ULL << (r + 1)) - 1ULL);
    uint32_t base = c == 0 ? cnt.x : c == 1 ? cnt.y : c == 2 ? cnt.z : cnt.w;
    uint64_t v = __ldg(ix.super + 4 * (loc >> AS_SB_SHIFT) + c) + base + (uint64_t)__popcll(m);
    /* the '$' row is stored as code 0 */
    if (c == 0 && (ix.primary >> 6) == blk && ((uint32_t)ix.primary & 63u) <= r) v -= 1;
    return v;
}

/* one backward-search step for character code c (fmidx.c:300-307) */
__device__ __forceinline__ void as_bs_step(const DevIndex &ix, uint32_t c, uint64_t &k, uint64_t &l) {
    uint64_t ok = as_occ(ix, c, k - 1);
    uint64_t ol = as_occ(ix, c, l);
    /* select chain: a dynamic index into the by-value kernel parameter would force a local-memory copy */
    const uint64_t cc = c == 0 ? ix.C[0] : c == 1 ? ix.C[1] : c == 2 ? ix.C[2] : ix.C[3];
    k = cc + ok + 1;
    l = cc + ol;
}

/* ---- 32-bit fast path: indexes of at most 2^30 rows (one super block, so the block counts are absolute and every
 * row number fits a register).  Same values as as_occ / as_bs_step, about a third of the instructions. ---- */
#ifndef AS_OCC_256
#define AS_OCC_256 1
#endif
struct BsChar32 {           /* per-character constants of one backward-search step */
    uint32_t xlo, xhi;      /* xor masks that turn the two planes into "row holds c" */
    uint32_t C;             /* dna_fmi.c[c] */
    uint32_t c;
};
__device__ __forceinline__ BsChar32 as_bs_char32(const DevIndex &ix, uint32_t c) {
    BsChar32 b;
    b.c = c;
    b.xlo = (c & 1u) ? 0u : 0xFFFFFFFFu;
    b.xhi = (c & 2u) ? 0u : 0xFFFFFFFFu;
    b.C = (uint32_t)(c == 0 ? ix.C[0] : c == 1 ? ix.C[1] : c == 2 ? ix.C[2] : ix.C[3]);
    return b;
}
__device__ __forceinline__ uint32_t as_shr_clamp(uint32_t v, uint32_t s) {   /* PTX shr: shifts >= 32 give 0 */
    uint32_t r;
    asm("shr.b32 %0, %1, %2;" : "=r"(r) : "r"(v), "r"(s));
    return r;
}
__device__ __forceinline__ uint32_t as_shl_clamp(uint32_t v, uint32_t s) {
    uint32_t r;
    asm("shl.b32 %0, %1, %2;" : "=r"(r) : "r"(v), "r"(s));
    return r;
}
/* the whole 32-byte FM block in ONE 256-bit request (sm_100: LDG.E.256): one L1 tag lookup per occ instead of two */
__device__ __forceinline__ void as_ld_block256(const FmBlock *b, uint32_t (&v)[8]) {
    asm("ld.global.nc.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
        : "l"(b));
}
__device__ __forceinline__ uint32_t as_occ32(const DevIndex &ix, const BsChar32 &b, uint32_t loc) {
    const uint32_t blk = loc >> 6, r = loc & 63u;
    uint32_t v8[8];                                                        /* cnt[0..3], lo.x lo.y hi.x hi.y */
#if AS_OCC_256
    as_ld_block256(ix.fm + blk, v8);
    const uint32_t base = b.c == 0 ? v8[0] : b.c == 1 ? v8[1] : b.c == 2 ? v8[2] : v8[3];
#else
    const uint32_t *p = reinterpret_cast<const uint32_t *>(ix.fm + blk);
    const uint32_t base = __ldg(p + b.c);                                   /* same 32-byte sector as the planes */
    const uint4 w = __ldg(reinterpret_cast<const uint4 *>(p) + 1);
    v8[4] = w.x; v8[5] = w.y; v8[6] = w.z; v8[7] = w.w;
#endif
    const uint32_t m0 = (v8[4] ^ b.xlo) & (v8[6] ^ b.xhi), m1 = (v8[5] ^ b.xlo) & (v8[7] ^ b.xhi);
    const uint32_t k0 = ~as_shl_clamp(0xFFFFFFFEu, r);                       /* rows 0..r of the low half */
    const uint32_t k1 = as_shr_clamp(0xFFFFFFFFu, 63u - r);                  /* rows 32..r of the high half */
    uint32_t v = base + __popc(m0 & k0) + __popc(m1 & k1);
    /* the '$' row is stored as code 0 */
    if (b.c == 0 && loc >= (uint32_t)ix.primary && blk == (uint32_t)(ix.primary >> 6)) v -= 1;
    return v;
}
__device__ __forceinline__ void as_bs_step32(const DevIndex &ix, uint32_t c, uint32_t &k, uint32_t &l) {
    const BsChar32 b = as_bs_char32(ix, c);
    const uint32_t ok = as_occ32(ix, b, k - 1), ol = as_occ32(ix, b, l);
    k = b.C + ok + 1;
    l = b.C + ol;
}

/* ---- table-driven backward search (k_vote): the per-character constants of a step live in a 4-entry shared-memory
 * table, so a step costs one LDS.128 instead of three select chains over kernel parameters, and the block count of
 * the character is picked from the loaded sector with three bit-selects on the same masks. ---- */
struct __align__(16) BsTab {
    uint32_t xlo, xhi;      /* all-ones when bit 0 / bit 1 of the character code is clear */
    uint32_t C;             /* dna_fmi.c[c] (low half; the 64-bit path adds Chi) */
    uint32_t plim;          /* c == A: rows [primary, primary + plim) share primary's block ('$' is coded A); else 0 */
};
__device__ __forceinline__ void as_bs_tab_init(const DevIndex &ix, BsTab *tab, uint32_t *tabChi, uint32_t tid) {
    if (tid < 4) {
        const uint64_t C = tid == 0 ? ix.C[0] : tid == 1 ? ix.C[1] : tid == 2 ? ix.C[2] : ix.C[3];
        BsTab b;
        b.xlo = (tid & 1u) ? 0u : 0xFFFFFFFFu;
        b.xhi = (tid & 2u) ? 0u : 0xFFFFFFFFu;
        b.C = (uint32_t)C;
        b.plim = tid == 0 ? 64u - ((uint32_t)ix.primary & 63u) : 0u;
        tab[tid] = b;
        tabChi[tid] = (uint32_t)(C >> 32);
    }
}
__device__ __forceinline__ uint32_t as_bitsel(uint32_t a, uint32_t b, uint32_t m) { return (a & m) | (b & ~m); }
/* both FM blocks of a backward-search step in one asm statement, so the two 256-bit requests are issued back to back
 * (left to itself ptxas sank the second load below the first one's consumers and serialised the two round trips) */
__device__ __forceinline__ void as_ld_block256x2(const FmBlock *pa, const FmBlock *pb, uint32_t (&a)[8], uint32_t (&b)[8]) {
    asm volatile("ld.global.nc.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%16];\n\t"
                 "ld.global.nc.v8.b32 {%8,%9,%10,%11,%12,%13,%14,%15}, [%17];"
                 : "=r"(a[0]), "=r"(a[1]), "=r"(a[2]), "=r"(a[3]), "=r"(a[4]), "=r"(a[5]), "=r"(a[6]), "=r"(a[7]),
                   "=r"(b[0]), "=r"(b[1]), "=r"(b[2]), "=r"(b[3]), "=r"(b[4]), "=r"(b[5]), "=r"(b[6]), "=r"(b[7])
                 : "l"(pa), "l"(pb));
}
/* count of the character in the loaded block up to row r, plus the block's own count: ~14 instructions */
__device__ __forceinline__ uint32_t as_occ_blk(const uint32_t (&v8)[8], const uint4 ct, uint32_t r) {
    const uint32_t base = as_bitsel(as_bitsel(v8[0], v8[1], ct.x), as_bitsel(v8[2], v8[3], ct.x), ct.y);
    const uint32_t k0 = ~as_shl_clamp(0xFFFFFFFEu, r);                       /* rows 0..r of the low half */
    const uint32_t k1 = as_shr_clamp(0xFFFFFFFFu, 63u - r);                  /* rows 32..r of the high half */
    const uint32_t m0 = ((v8[4] ^ ct.x) & k0) & (v8[6] ^ ct.y), m1 = ((v8[5] ^ ct.x) & k1) & (v8[7] ^ ct.y);
    return base + __popc(m0) + __popc(m1);
}
__device__ __forceinline__ void as_bs_step32t(const DevIndex &ix, const uint4 ct, uint32_t &k, uint32_t &l) {
    const uint32_t a = k - 1, b = l, p = (uint32_t)ix.primary;
    uint32_t va[8], vb[8];
    as_ld_block256x2(ix.fm + (a >> 6), ix.fm + (b >> 6), va, vb);
    uint32_t ok = as_occ_blk(va, ct, a & 63u);
    uint32_t ol = as_occ_blk(vb, ct, b & 63u);
    ok -= (a - p) < ct.w ? 1u : 0u;                                          /* the '$' row is stored as code 0 */
    ol -= (b - p) < ct.w ? 1u : 0u;
    k = ct.z + ok + 1;
    l = ct.z + ol;
}
__device__ __forceinline__ void as_bs_step64t(const DevIndex &ix, const uint4 ct, uint32_t Chi, uint32_t c,
                                              uint64_t &k, uint64_t &l) {
    const uint64_t a = k - 1, b = l;
    const uint64_t C = ((uint64_t)Chi << 32) | ct.z;
    uint32_t va[8], vb[8];
    as_ld_block256x2(ix.fm + (a >> 6), ix.fm + (b >> 6), va, vb);
    uint64_t ok = __ldg(ix.super + 4 * (a >> AS_SB_SHIFT) + c) + as_occ_blk(va, ct, (uint32_t)a & 63u);
    uint64_t ol = __ldg(ix.super + 4 * (b >> AS_SB_SHIFT) + c) + as_occ_blk(vb, ct, (uint32_t)b & 63u);
    ok -= (a - ix.primary) < (uint64_t)ct.w ? 1u : 0u;
    ol -= (b - ix.primary) < (uint64_t)ct.w ? 1u : 0u;
    k = C + ok + 1;
    l = C + ol;
}

__device__ __forceinline__ uint64_t as_sa(const DevIndex &ix, uint64_t row) {
    uint64_t v = __ldg(ix.sa_lo + row);
    if (ix.sa_hi) v |= (uint64_t)__ldg(ix.sa_hi + row) << 32;
    return v;
}

/* SA[row] on any handle: the suffix array itself, or -- on a handle that keeps only the sampled array (as_open_opts.sampled_sa,
 * fmidx.c:152-163) -- csa_access' walk (fmidx.c:315-331, in the form its own test expects: LF(row) = C[ch] + occ(ch, row)):
 * follow LF until a sampled row; every step reads ONE FM sector (the row's BWT code and its occ come from the same block). */
__device__ __forceinline__ uint64_t as_locate(const DevIndex &ix, uint64_t row) {
    if (ix.sa_lo) return as_sa(ix, row);
    const uint64_t ratio = ix.csa_ratio;
    uint64_t counter = 0;
    while (row % ratio != 0) {
        if (row == ix.primary) return counter;                          /* bwt[row] == '$': the suffix starts at 0 */
        uint4 cnt;
        uint64_t lo, hi;
        as_ld_block(ix.fm + (row >> 6), cnt, lo, hi);
        const uint32_t r = (uint32_t)row & 63u;
        const uint32_t c = (uint32_t)((lo >> r) & 1u) | ((uint32_t)((hi >> r) & 1u) << 1);
        uint64_t m = ((c & 1u) ? lo : ~lo) & ((c & 2u) ? hi : ~hi);
        m &= (r == 63u) ? ~0ULL : ((1ULL << (r + 1)) - 1ULL);
        const uint32_t base = c == 0 ? cnt.x : c == 1 ? cnt.y : c == 2 ? cnt.z : cnt.w;
        uint64_t occ = __ldg(ix.super + 4 * (row >> AS_SB_SHIFT) + c) + base + (uint64_t)__popcll(m);
        if (c == 0 && (ix.primary >> 6) == (row >> 6) && ((uint32_t)ix.primary & 63u) <= r) occ -= 1;   /* '$' is stored as A */
        const uint64_t cc = c == 0 ? ix.C[0] : c == 1 ? ix.C[1] : c == 2 ? ix.C[2] : ix.C[3];
        row = cc + occ;
        counter++;
    }
    const uint64_t j = row / ratio;
    uint64_t v = __ldg(ix.csa_lo + j);
    if (ix.csa_hi) v |= (uint64_t)__ldg(ix.csa_hi + j) << 32;
    return v + counter;
}

/* 2-bit packed sequences: 16 bases per u32, base i at bits [2(i&15), 2(i&15)+1] of word i>>4 */
__device__ __forceinline__ uint32_t as_base_at(const uint32_t *w, uint64_t i) {
    return (__ldg(w + (i >> 4)) >> (2u * ((uint32_t)i & 15u))) & 3u;
}
/* up to 16 bases starting at base offset i (needs one word of slack past the end) */
__device__ __forceinline__ uint32_t as_bases16(const uint32_t *w, uint64_t i) {
    uint32_t a = __ldg(w + (i >> 4)), b = __ldg(w + (i >> 4) + 1);
    return __funnelshift_r(a, b, 2u * ((uint32_t)i & 15u));
}
/* 32 bases starting at base offset i as a u64 (needs two words of slack) */
__device__ __forceinline__ uint64_t as_bases32(const uint32_t *w, uint64_t i) {
    const uint32_t *p = w + (i >> 4);
    uint32_t a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2);
    uint32_t s = 2u * ((uint32_t)i & 15u);
    return ((uint64_t)__funnelshift_r(b, c, s) << 32) | __funnelshift_r(a, b, s);
}

/* four ASCII bases in a word -> their four 2-bit codes in the low byte (A0 C1 G2 T3, as_base_code per byte lane) */
__device__ __forceinline__ uint32_t as_pack4(uint32_t x) {
    const uint32_t c = ((x >> 1) & 0x03030303u) ^ ((x >> 2) & 0x01010101u);     /* one code per byte */
    return (c * 0x01041040u) >> 24;                                               /* b0 | b1<<2 | b2<<4 | b3<<6, no carries */
}
/* lim <= 16 ASCII bases at src (any alignment) -> one packed word: aligned 32-bit loads shifted into place.  Reads whole
 * words, so up to 3 bytes either side of the bases are touched (inside the 256-byte aligned batch buffer). */
__device__ __forceinline__ uint32_t as_pack16(const uint8_t *src, uint32_t lim) {
    const uint32_t mis = (uint32_t)(reinterpret_cast<uintptr_t>(src) & 3u);
    const uint32_t *p = reinterpret_cast<const uint32_t *>(src - mis);
    const uint32_t nload = (mis + lim + 3) >> 2;                                  /* <= 5 words */
    uint32_t q[5], v = 0;
#pragma unroll
    for (int t = 0; t < 5; t++) q[t] = (uint32_t)t < nload ? p[t] : 0u;
    const uint32_t sh = 8u * mis;
#pragma unroll
    for (int t = 0; t < 4; t++) v |= as_pack4(__funnelshift_r(q[t], q[t + 1], sh)) << (8 * t);
    if (lim < 16) v &= (1u << (2 * lim)) - 1u;
    return v;
}

#endif /* __CUDACC__ */
