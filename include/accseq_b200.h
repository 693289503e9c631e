/*
 * include/accseq_b200.h -- C ABI of libaccseq_b200.so: the B200 (sm_100a) implementation of AccSeqV9's
 * per-read mapping hot path.
 *
 * The reference (lisanhu/LongReadMapper) has no plugin/FFI seam for this path: the loop nest is inlined in
 * `single_end()` (alnmain.c:277-551).  The boundary is therefore cut where SURVEY.md 8(b) puts it --
 * after `refactor_reads_seq` (alnmain.c:302-303) and before SAM printing (alnmain.c:485) -- and each
 * entry below names the reference code it replaces.  Plain pointers and sizes only; no CUDA or torch
 * types.  All functions return 0 on success or a negative AS_E* code (the reference signals no errors).
 * There is NO CPU fallback: every compute entry fails with AS_ENODEV when no sm_100 device is usable.
 *
 * Host pointers may be pageable or pinned (as_host_alloc); pinned buffers reach full PCIe speed.
 * One host thread per as_gpu handle; results are written in input order.
 */
#ifndef ACCSEQ_B200_H
#define ACCSEQ_B200_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define AS_OK        0
#define AS_EINVAL   -22
#define AS_ENOMEM   -12
#define AS_ENODEV   -19
#define AS_EIO       -5
#define AS_ECUDA   -200     /* a CUDA runtime call failed; as_last_error() has the text */
#define AS_ERANGE   -34

typedef struct as_gpu as_gpu;       /* opaque: one per GPU, owns the device copy of the index */

/* ---- the index as the reference holds it in host memory after init() (alnmain.c:179-256) ---- */
typedef struct as_mta {             /* mta_entry (accaln.h:67-71) without the name string */
    uint64_t offset;                /* start of the forward copy inside .cat */
    uint64_t seq_len;
} as_mta;

typedef struct as_index_view {
    uint64_t        length;         /* dna_fmi.length = N (fmidx.h:16-21) */
    const uint64_t *C;              /* dna_fmi.c, 256 entries indexed by ASCII code */
    const uint64_t *O;              /* dna_fmi.o, o_len entries, 4 per block of o_ratio rows */
    uint64_t        o_len;
    int             o_ratio;
    const char     *bwt;            /* dna_fmi.bwt, one byte per row: A C G T $ */
    const uint64_t *lc;             /* lc_hash.lc (lchash.h:12-16): (k,l) pairs, 2*4^hlen entries.  lc, sa and content may be
                                     * NULL: a lookup-only handle (fmi_aln / csa walks), which as_map_batch refuses */
    uint64_t        lc_len;
    int             hlen;
    const void     *sa;             /* suffix array, N entries */
    int             sa_entry_bytes; /* 8: sa_buf->mem (ui40_t structs, sa_use.h:17-20) or u64;  5: raw .sa5 bytes;  4: u32 */
    const char     *content;        /* context.content: the .cat text, con_len bytes */
    uint64_t        con_len;
    const as_mta   *mta;            /* context.mta */
    int             mta_len;
} as_index_view;

/* Upload the index once and repack it into the device layouts (DESIGN.md "HBM layout").
 * Replaces nothing in the reference (it keeps host arrays); lookups return identical values. */
int  as_gpu_open(as_gpu **h, int device, const as_index_view *ix);
void as_gpu_close(as_gpu *h);

/* Open-time options (what used to be environment variables: a drop-in library must not read its host's environment).
 * Zero-initialised = the defaults; as_gpu_open / as_gpu_build use the defaults. */
typedef struct as_open_opts {
    int32_t  lc2_hlen;        /* depth of the deep k-mer table: 0 = min(16, ceil(log4 N)+1), +1 while the table stays within 4 GB; <= hlen disables it; 13..16 forces */
    int32_t  no_sax;          /* 1: do not build the SA + context table (PART 1 then gathers SA and text separately) */
    uint64_t build_part_max;  /* as_gpu_build only: > 0 forces the partitioned suffix sort with at most this many suffixes
                               * per partition (the path texts of >= 2^32-1 rows take) */
    int32_t  sampled_sa;      /* 0: the full suffix array stays in HBM (5 bytes/row, + 8 for the SA + context table);
                               * r > 0: keep only SA[r*i] (csa, fmidx.c:152-163) and resolve vote positions by the LF walk of
                               * csa_access (fmidx.c:315-331): no SA, no SA + context table -- ~1/r of the SA bytes per GPU */
    int32_t  l2_fetch_gran;   /* > 0: cudaLimitMaxL2FetchGranularity of the device (32, 64 or 128 bytes); 0 leaves the device's
                               * setting alone.  PART 1 is random 8-byte gathers: anything fetched beyond the 32-byte sector is waste */
    int32_t  reserved[4];
} as_open_opts;
int  as_gpu_open_opts(as_gpu **h, int device, const as_index_view *ix, const as_open_opts *opts);
/* A handle WITHOUT an index: serves the aligners on explicit pairs (as_cigar_align_batch, as_edit_distance_batch) and,
 * after as_gpu_attach_sa, as_sa_access_batch.  What libaccseq_compat's cigar_align / sa_access run on. */
int  as_gpu_open_aligner(as_gpu **h, int device);
/* Tables added to an open handle: a suffix array (n entries of 4, 5 or 8 bytes) for a handle opened without one, and the
 * caller's sampled suffix array csa[i] = SA[csa_ratio*i] (dna_fmi.csa as fmi_read loads it, fmidx.c:266-271). */
int  as_gpu_attach_sa(as_gpu *h, const void *sa, uint64_t n, int sa_entry_bytes);
int  as_gpu_attach_csa(as_gpu *h, const uint64_t *csa, uint64_t csa_len, int csa_ratio);

/* Build the index ON the GPU from the .cat text (asindex.c:97-110: forward + reverse complement per sequence,
 * one final '$') -- replaces fmi_build (fmidx.c:166-199: pSAscan suffix sort psascan/sa_use.cc:8-18, BWT
 * :76-98, C table :101-125, O table :128-150) and lc_build (lchash.c:52-73) of the reference's `accidx`.
 * Texts of 2^32-1 rows and more (3.1 Gbp genomes) are sorted in partitions with 40-bit rows.  The handle is ready for as_map_batch; as_gpu_export returns the same index in
 * the reference's layouts, byte-identical to the files accidx writes. */
int  as_gpu_build(as_gpu **h, int device, const char *content, uint64_t con_len, const as_mta *mta, int mta_len,
                  int hlen);
int  as_gpu_build_opts(as_gpu **h, int device, const char *content, uint64_t con_len, const as_mta *mta, int mta_len,
                       int hlen, const as_open_opts *opts);
typedef struct as_index_out {       /* caller-owned buffers; a NULL member is skipped */
    uint64_t *C;                    /* 256 entries */
    uint64_t *O;                    /* 4*(N/o_ratio+1) entries (fmidx.c:128-150) */
    int       o_ratio;              /* any positive ratio; the reference uses 32 (asindex.c:138) */
    char     *bwt;                  /* N bytes */
    uint64_t *lc;                   /* 2*4^hlen entries, MSB-first code order (lchash.c:36-49) */
    void     *sa;                   /* N entries of sa_entry_bytes (5 = .sa5 bytes, 8 = u64 / ui40_t structs) */
    int       sa_entry_bytes;
    uint64_t *csa;                  /* ceil(N/csa_ratio) entries written: SA[csa_ratio*i] (fmidx.c:153-163) */
    int       csa_ratio;
} as_index_out;
int  as_gpu_export(as_gpu *h, as_index_out *out);
/* rows of the index (dna_fmi.length) and the hash length of a handle */
uint64_t as_gpu_index_length(as_gpu *h);
/* Route all work of this handle to a caller-owned CUDA stream (a cudaStream_t cast to void*); NULL = the
 * handle's own stream.  Lets a host framework time/overlap the calls on its own stream. */
int  as_gpu_set_stream(as_gpu *h, void *cuda_stream);
const char *as_last_error(void);
int  as_device_count(void);

/* ---- parameters of the path ---- */
typedef struct as_params {
    uint32_t seed_len;   /* context.seed_len, default 20 (alnmain.c:577-580) */
    uint32_t thres;      /* context.uninformative_thres == histo_cap, default 300 */
    int32_t  gact_T;     /* GACT tile size (DESIGN.md GACT spec), default 320 */
    int32_t  gact_O;     /* GACT overlap, default 120 */
    uint32_t flags;      /* AS_F_* */
    int32_t  gact_band;  /* bit-parallel GACT: half-width of the banded first attempt at a tile; 0 = adaptive (default),
                          * < 0 = full matrix only.  Results never depend on it (a tile the band cannot prove exact is
                          * filled again with the whole matrix). */
    /* scheduling knobs (A/B runs and tests; results never depend on them).  as_params_default sets the "auto" values. */
    int32_t  vote_ectx;  /* PART 1 decides one- and two-row deep-table entries from their context bases: -1 = only when the
                          * index is larger than L2 (default), 0 = never, 1 = always */
    int32_t  gact_order; /* hand reads to GACT by decreasing length: -1 = only for reads of several tiles (default), 0, 1 */
    int32_t  gact_blocks_per_sm;  /* resident GACT blocks per SM; 0 = occupancy query (default) */
    uint32_t chunk_reads;         /* as_map_batch: fixed chunk size; 0 = the ramp-up / ramp-down plan (default) */
    int32_t  chunk_tail;          /* as_map_batch: how many times the last chunk is split 2:1; -1 = default (1).
                                   * as_map_batch_packed: chunks between the first and the last one (default 2); 100 * d + m =
                                   * first and last chunk of n / d reads (default 8), m chunks in between */
    int32_t  vote_thread;         /* PART 1 of short reads (<= 8 seeds per phase) by the thread-per-read kernel, the rest by the
                                   * group kernel: -1 = when it applies and the index has at most 2^30 rows (default), 0 = group
                                   * kernel only, 1 = whenever it applies */
    int32_t  gact_runs;           /* as_map_batch_packed: one-tile reads leave their run bytes straight from GACT's traceback
                                   * (no op bytes, no run-length pass): -1 / 1 = on (default), 0 = off */
    int32_t  vote_blocks_per_sm;  /* resident blocks per SM of the thread-per-read PART 1 kernel; 0 = occupancy query (default).
                                   * Its blocks and GACT's are persistent and between them fill an SM's registers: fewer of either
                                   * lets the two lanes of as_map_batch_packed run PART 1 of one chunk beside GACT of another */
} as_params;
#define AS_F_MYERS        1u   /* also run the edlib-equivalent NW distance with k=(int)(0.05*len*2) per read */
#define AS_F_NO_ALIGN     2u   /* stop after PART 1+2 (no GACT): results carry loc/off/strand/flag only */
#define AS_F_GACT_WARP    4u   /* force the one-read-per-warp anti-diagonal GACT kernel (default for n < 2048) */
#define AS_F_GACT_BITPAR  8u   /* force the bit-parallel one-read-per-lane GACT kernel (default for n >= 2048) */
#define AS_F_WIDE_INDEX  16u   /* force the 64-bit backward-search path (default only for indexes of > 2^30 rows) */
#define AS_F_NO_LC2      32u   /* PART 1 starts every seed from the reference's 12-mer table, not the deep table */
#define AS_F_RLE_CIGAR   64u   /* store_mem rows hold run bytes ((len << 2) | class; class 0 M, 1 I, 2 D; len 1..63, longer
                                * runs continue in the next byte) and n_cigar_op counts them: what parse_cigar prints
                                * (alnmain.c:497-498), ~10 bytes instead of ~150 per short read across PCIe */
#define AS_F_KEEP_READS 128u   /* as_map_batch leaves the host reads untouched: strand-1 reads are NOT copied back reverse-
                                * complemented (alnmain.c:433-438); the SAM emitter reverse-complements them as it prints */
void as_params_default(as_params *p);

/* ---- per-read output: the fields of `result` (accaln.h:95-103) that the path computes ---- */
typedef struct as_result {
    uint64_t loc;         /* best.key: voted .cat coordinate (alnmain.c:427) */
    uint64_t off;         /* result.off: 0-based forward offset in the sequence; SAM POS = off+1 */
    int32_t  seq_id;      /* g_name = mta[seq_id].seq_name; -1 when seq_lookup fails (reference: UB) */
    int32_t  strand;      /* seq_meta.strand */
    int32_t  flag;        /* 0, 16 or 4   (alnmain.c:466-474) */
    int32_t  mapq;        /* 255 or 0 */
    int32_t  ed;          /* result.ed = GACT score */
    int32_t  valid;
    int32_t  n_cigar_op;  /* cigar.n_cigar_op: columns written to the read's slice of store_mem */
    int32_t  myers;       /* edlib NW distance (-1 above k), or -2 when AS_F_MYERS is off */
    int32_t  phases;      /* value of `iter` when the phase loop ended (parity aid) */
    int32_t  from_outer;  /* 1 when best came from the outer vote list (alnmain.c:400-403) */
} as_result;

/*
 * as_map_batch -- replaces alnmain.c:330-479 (PART 1-3) for n reads.
 *   reads      host buffer in refactor_reads_seq layout (alnmain.c:87-103): read i at reads + i*stride,
 *              upper-case ACGT, lens[i] bases.  Reads mapped to strand 1 are reverse-complemented IN
 *              PLACE, as alnmain.c:433-438 does (the OpenACC region's copy(reads_mem)).
 *   results    n entries.
 *   store_mem  caller-owned CIGAR column storage, read i at store_mem + i*store_stride, one op byte per
 *              column (alnmain.c:314-325 uses store_stride = 2*max_read_len); may be NULL.  Columns at and past
 *              results[i].n_cigar_op are unspecified, as in the reference (it never initialises `store`): only the
 *              columns some read of the chunk uses are copied back.
 * The batch is processed in chunks (as_params.chunk_reads overrides the size) pipelined over three streams; results do
 * not depend on the chunking.
 * On an error the call returns only after every copy it had enqueued on the caller's buffers has ended, and
 * as_gpu_last_reads_done says how far it got: reads [0, done) have their results, their CIGAR columns and -- strand 1 --
 * their in-place reverse complement (alnmain.c:433-438); reads [touched, n) are exactly as the caller passed them and
 * their results are untouched; the chunk in between (if any) is partly delivered.  Argument errors (a read longer than its
 * row, a too narrow store_stride) are found before anything is copied back: done = touched = 0.  After a successful call
 * done = touched = n.
 */
int as_map_batch(as_gpu *h, const as_params *p, char *reads, size_t stride, const uint32_t *lens, size_t n,
                 as_result *results, uint8_t *store_mem, size_t store_stride);
/* How many leading reads of the last as_map_batch call were delivered to the caller's buffers (see above); the reference has
 * no counterpart: its loop either finishes or the process dies (alnmain.c:330-479). */
size_t as_gpu_last_reads_done(as_gpu *h, size_t *touched);

/*
 * as_map_batch_packed -- the same path for hosts that hold reads 2-bit packed (what the FASTQ ingest of libaccseq_host.so
 * produces while it parses) and want records, not buffers, back: 44 B per 150-bp read cross PCIe on the way in and about
 * 22 B on the way out, instead of 155 + 104 (refactor_reads_seq buffer in, as_result + run-byte rows out).
 *   packed     read i at packed + i*wstride: 16 bases per 32-bit word, base j at bits [2(j&15), 2(j&15)+1] of word j>>4,
 *              A0 C1 G2 T3 (bytes outside ACGT take the code of ((c>>1)&3)^((c>>2)&1), as the ASCII path does; the reference's
 *              behaviour for them is undefined, lchash.c:38-45); bits past lens[i] must be zero; wstride >= ceil(max len / 16)
 *   results    n compact records (below)
 *   runs       run bytes ((len << 2) | class, as AS_F_RLE_CIGAR) of read 0, read 1, ... back to back: read i's start at the
 *              sum of results[j].n_runs over j < i.  runs_cap = capacity; *runs_len = bytes written.  AS_ERANGE when the
 *              capacity does not suffice (runs_len then holds the size needed; results are complete).
 * Strand-1 reads are NOT handed back reverse-complemented: the emitter does that as it prints (as AS_F_KEEP_READS).
 * Every as_result field can be recomputed from a record (as_result_from_compact in accseq_host.h); phases / from_outer /
 * myers are not carried.
 */
typedef struct as_result_compact {
    uint64_t pos;        /* bits 0..39 off (0-based forward offset, SAM POS - 1); bits 40..55 seq_id + 1 (0 = seq_lookup failed);
                          * bits 56..63 the low 8 bits of flag (0, 16 or 4) */
    int32_t  ed;         /* GACT score, -1 = failure */
    uint32_t n_runs;     /* bits 0..26 run bytes of this read; bits 27..30 flag bits 8..11 (none set by this path); bit 31 the
                          * strand (a read whose alignment failed keeps FLAG 4, yet prints reverse-complemented, alnmain.c:433-438) */
} as_result_compact;
#define AS_RC_NRUNS(c)  ((c).n_runs & 0x07FFFFFFu)
#define AS_RC_STRAND(c) ((c).n_runs >> 31)
#define AS_RC_FLAG(c)   ((unsigned)((c).pos >> 56) | ((((c).n_runs >> 27) & 0xFu) << 8))
#define AS_RC_SEQ_ID(c) ((int)(((c).pos >> 40) & 0xFFFFu) - 1)
#define AS_RC_OFF(c)    ((c).pos & 0xFFFFFFFFFFULL)
int as_map_batch_packed(as_gpu *h, const as_params *p, const uint32_t *packed, size_t wstride, const uint32_t *lens, size_t n,
                        as_result_compact *results, uint8_t *runs, size_t runs_cap, uint64_t *runs_len);

/* Same path with every buffer already resident in device memory (pointers are device addresses); used to
 * time the kernels without PCIe and by frameworks that keep reads on the GPU.
 * d_reads: the packing kernel reads the batch as aligned 32-bit words, so the allocation that holds it must start on a
 * 4-byte boundary and end on one (cudaMalloc and every sub-allocator built on it do; a row pointer need not be aligned).
 * Bytes outside ACGT/acgt are outside the reference's defined behaviour (DESIGN.md parity exclusions): they seed and
 * align as the 2-bit code of their bits 1-2 ('N' -> G), print unchanged in SEQ. */
int as_map_batch_device(as_gpu *h, const as_params *p, char *d_reads, size_t stride, const uint32_t *d_lens,
                        size_t n, uint32_t max_len, as_result *d_results, uint8_t *d_store, size_t store_stride);

/* as_map_batch_device only enqueues work on the handle's stream; this (or a synchronize of the caller's
 * stream) completes it. */
int as_gpu_sync(as_gpu *h);
/* Number of CUDA kernels this handle has launched since as_gpu_open (bench.py's gpu_launches). */
uint64_t as_gpu_launch_count(as_gpu *h);

/* Per-shard counters for the end-of-run statistics (alnmain.c:541 "Sensitivity: valid/total"). */
typedef struct as_stats {
    uint64_t total, valid, n_flag4, n_flag16, sum_ed, dp_cells, seeds, sa_hits, bs_steps, overflow_reads;
    uint64_t dp_cells_eval;   /* DP cells the GACT kernels actually filled (banded pass + redone reads); dp_cells is
                               * the tile area Tq x Tt of the spec */
    uint64_t gact_redone;     /* warp-tiles (one tile of 32 reads) filled a second time with the whole matrix */
    uint64_t fm_steps_exec;   /* backward-search steps the GPU really walked through the FM index; bs_steps counts the
                               * steps of the reference's loop (fmidx.c:300-310), the rest were decided by the text */
    uint64_t text_verified;   /* seeds whose one-row interval was finished by one SA + one text gather */
    uint64_t overflow_queued; /* reads whose votes of one phase did not fit the shared-memory table (8 slots per lane of the
                               * read's group) and were redone by the overflow pass over a table in global memory: repeat
                               * families.  overflow_reads counts reads the overflow pass itself could not hold (never) */
    uint64_t vote_deferred;   /* short reads the thread-per-read PART 1 kernel handed to the group kernel (a seed with more
                               * than two hits, more distinct vote buckets in a phase or in the outer list than a thread keeps) */
    uint64_t gact_queued;     /* one-tile reads the one-word band of GACT's first attempt could not prove exact (walk cost > 15 or
                               * more than 7 net indels) and that the second launch aligned again with wider bands */
} as_stats;
int as_gpu_stats(as_gpu *h, as_stats *out);          /* accumulated since open / last reset */
int as_gpu_stats_reset(as_gpu *h);
/* Device time of the stages of the last as_map_batch* call, milliseconds (CUDA events on the call's streams):
 * [0] pack+seed+vote (PART 1)  [1] locus/strand (PART 2)  [2] GACT (+ result fill, PART 3)  [3] Myers  [4] unused
 *     -- kernel time, summed over the chunks of the call
 * [5] span of the H2D copies  [6] span of the D2H copies (as_map_batch pipelines chunks: H2D of chunk i+1, kernels of
 *     chunk i and D2H of chunk i-1 overlap, so [5] and [6] overlap the kernel time)  [7] whole call */
int as_gpu_last_timing(as_gpu *h, float ms[8]);
/* When the events of chunk `chunk` of the last as_map_batch / as_map_batch_packed call fired, ms after the call's start:
 * [0] kernels begin [1] PART 1 done [2] PART 2 done [3] GACT done [4] Myers done [5] H2D of the chunk done [6] kernels done
 * [7] fixed-size D2H of the chunk done.  Returns the number of chunks (a negative code on error). */
int as_gpu_last_chunk_timeline(as_gpu *h, int chunk, float ms[8]);
/* Host-side time of the last as_map_batch_packed call, microseconds: [0] until every kernel and copy was enqueued,
 * [1] waiting for the device, [2] reading the event times.  [0] + [1] + [2] = the caller's wall time of the call. */
int as_gpu_last_host_timing(as_gpu *h, double us[3]);
/* Bytes the last as_map_batch moved over PCIe in each direction. */
int as_gpu_last_transfer(as_gpu *h, uint64_t *h2d_bytes, uint64_t *d2h_bytes);

/* ---- batched forms of the lookup primitives (parity tests, other callers) ---- */
/* lc_aln (lchash.c:89-104) on n seeds of qlen bases each, seeds + i*qlen; k,l,r as lc_aln leaves them
 * (k,l start at 1, N-1 as alnmain.c:354 does). */
int as_lc_aln_batch(as_gpu *h, const char *seeds, int qlen, size_t n, uint64_t *k, uint64_t *l, uint64_t *r);
/* fmi_aln (fmidx.c:295-313) from the full range. */
int as_fmi_aln_batch(as_gpu *h, const char *qry, int qlen, size_t n, uint64_t *k, uint64_t *l, uint64_t *r);
/* fmi_aln continuing from the caller's intervals: k[i], l[i] are read and overwritten (fmidx.c:303; lc_aln calls it
 * that way after the table lookup, lchash.c:103) */
int as_fmi_aln_from_batch(as_gpu *h, const char *qry, int qlen, size_t n, uint64_t *k, uint64_t *l, uint64_t *r);
/* lc_build (lchash.c:52-73): the 4^hlen backward searches of the reference, one thread each; lc receives 2*4^hlen
 * values in the reference's (MSB-first) order.  Needs only the FM index of the handle. */
int as_lc_build(as_gpu *h, int hlen, uint64_t *lc);
/* sa_access (fmidx.c:18-33) */
int as_sa_access_batch(as_gpu *h, const uint64_t *rows, size_t n, uint64_t *pos);
/* csa_access (fmidx.c:315-331): SA[row] from the sampled suffix array csa[i] = SA[csa_ratio*i] (_build_csa,
 * fmidx.c:152-163; the reference builds ratio 4, fmidx.c:194) by walking LF to a sampled row.  The sample is taken
 * on the device at the first call with a given ratio and kept in the handle.  literal = 0: LF(row) =
 * c[ch] + occ(ch,row), the walk the reference's test expects (csa_access(i) == sa[i], test/test-fmidx.cc:104-106);
 * literal = 1: the statement as the reference wrote it (`- 1`, and 0 after more than 5*ratio steps). */
int as_csa_access_batch(as_gpu *h, int csa_ratio, int literal, const uint64_t *rows, size_t n, uint64_t *pos);
/* PART 1 only (alnmain.c:333-405): best.key / best.val / phases / from_outer per read. */
int as_vote_batch(as_gpu *h, const as_params *p, const char *reads, size_t stride, const uint32_t *lens,
                  size_t n, uint64_t *best_key, uint64_t *best_val, int32_t *phases, int32_t *from_outer);
/* cigar_align (mutils.c:94-105) on n (query,target) pairs: pair i is q + i*q_stride (qlens[i] bases) vs
 * t + i*t_stride (tlens[i] bases), ASCII ACGT.  ops + i*ops_stride receives the column ops. */
int as_cigar_align_batch(as_gpu *h, const as_params *p, const char *q, size_t q_stride, const uint32_t *qlens,
                         const char *t, size_t t_stride, const uint32_t *tlens, size_t n,
                         uint8_t *ops, size_t ops_stride, int32_t *n_ops, int32_t *score);
/* edlib-equivalent NW edit distance (edlib.cpp:694-895): dist[i] = distance, or -1 if ks[i] >= 0 and the
 * distance exceeds ks[i].  ks may be NULL (no bound). */
int as_edit_distance_batch(as_gpu *h, const char *q, size_t q_stride, const uint32_t *qlens,
                           const char *t, size_t t_stride, const uint32_t *tlens, size_t n,
                           const int32_t *ks, int32_t *dist);
/* edlib's ALIGNMENT PATH for the same pairs: edlibAlign(EDLIB_MODE_NW, EDLIB_TASK_PATH) (edlib.cpp:141-260 -> obtainAlignment
 * :1110-1150 -> obtainAlignmentTraceback :909-1108).  ops + i*ops_stride receives the alignment of pair i in forward order, one
 * EDLIB_EDOP_* code per column (0 match, 1 insert = a query base against a gap, 2 delete, 3 mismatch: what
 * edlibAlignmentToCigar, edlib.cpp:262-311, turns into a CIGAR); n_ops[i] = its length; dist[i] as as_edit_distance_batch
 * (-1 and no alignment when ks[i] >= 0 is exceeded, edlib.cpp:705-708).  Tie-breaks are edlib's: from the last cell backwards,
 * up (insert) before left (delete) before the diagonal.  Range: queries of at most AS_PATH_MAX_QLEN bases and pairs within the
 * 1 MiB of alignment data below which edlib itself uses this traceback (20 * ceil(qlen/64) * tlen + 8 * tlen < 2^20,
 * edlib.cpp:1131-1133; beyond it edlib switches to Hirschberg's divide and conquer and the call returns AS_ERANGE);
 * ops_stride >= qlen + tlen of every pair.  Empty sequences (outside what edlibAlign handles) give all deletes / all inserts. */
#define AS_PATH_MAX_QLEN 256
int as_edit_alignment_batch(as_gpu *h, const char *q, size_t q_stride, const uint32_t *qlens,
                            const char *t, size_t t_stride, const uint32_t *tlens, size_t n,
                            const int32_t *ks, int32_t *dist, uint8_t *ops, size_t ops_stride, int32_t *n_ops);
/* The same distances (the same kernel) for callers that cannot allocate -- code under CUDA graph capture, tight serving
 * loops, harnesses that forbid cudaMalloc.  NO COUNTERPART IN THE REFERENCE.
 * Needs no handle and calls no cudaMalloc / cudaFree / allocating routine: the scratch the batched entry takes from the
 * handle (2-bit packed sequences, offsets, checked lengths) is one file-scope device array of fixed size (~5.8 MB that every
 * GPU loading the library holds until exit).  Limits: n <= AS_NOALLOC_MAX_PAIRS (more: AS_ERANGE, nothing launched), every
 * length <= AS_NOALLOC_MAX_LEN (a pair beyond it gets dist[i] = AS_ERANGE).  All pointers are DEVICE addresses on the current
 * device (q / t: ASCII rows as in as_edit_distance_batch, inside allocations that start and end on 4-byte boundaries; d_ks
 * may be NULL); the work is only enqueued on `cuda_stream` (a cudaStream_t, NULL = the default stream) -- synchronise it
 * before reading d_dist.  Every call shares the one array, which keeps the last call's contents: calls that overlap on the
 * GPU (two streams, two threads) race; calls on one stream, or otherwise ordered, do not.  The first call on a device
 * resolves the array's device address (cudaGetSymbolAddress), so make one call before starting a graph capture. */
#define AS_NOALLOC_MAX_PAIRS 32768
#define AS_NOALLOC_MAX_LEN   256
int as_edit_distance_device_noalloc(void *cuda_stream, const char *d_q, size_t q_stride, const uint32_t *d_qlens,
                                    const char *d_t, size_t t_stride, const uint32_t *d_tlens, size_t n,
                                    const int32_t *d_ks, int32_t *d_dist);

/* pinned host memory helpers */
void *as_host_alloc(size_t bytes);
void  as_host_free(void *p);

#ifdef __cplusplus
}
#endif
#endif
