/*
 * oracle/oracle.h -- CPU restatement of the AccSeqV9 per-read mapping hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product:
 * only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this library, and only as the checker or
 * the timed CPU baseline.  The product (longreadmapper_b200/csrc) never links
 * or calls it and fails loudly when its CUDA library is missing.
 *
 * Parity status:
 *   - index lookups, voting, phase control, locus/strand, FLAG/MAPQ/POS:
 *     PINNED -- tests/test_oracle_vs_ref.py diffs every function below against
 *     the reference's own object code compiled into oracle/_ref/ (recipe:
 *     oracle/Makefile) and against SAM fixtures under tests/golden/ produced
 *     by the compiled reference `accaln`.
 *   - Myers/edlib NW edit distance: PINNED against the reference's in-tree
 *     edlib (oracle/_ref/libedlib_ref.so) and edlib/ed_test.c's known answer.
 *   - GACT CIGAR / ED: PARITY UNPINNED.  The reference's aligner lives in the
 *     un-vendored submodule lisanhu/GACT @ 5ee0690bf95a8342afdfcd721108edab9f386034
 *     (reference .gitmodules:4-6, .SUBMODULES.json:9-15; directory empty).
 *     orc_gact() implements the documented stand-in spec (DESIGN.md "GACT
 *     spec"), anchored on the call contract visible at mutils.c:94-105 and
 *     alnmain.c:440-446,498.
 *
 * Every function cites the reference file:line it restates.
 */
#ifndef ACCSEQ_ORACLE_H
#define ACCSEQ_ORACLE_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct {
    char     *name;
    uint64_t  name_len;
    uint64_t  offset;   /* offset of the forward copy inside .cat  (asindex.c:89-93) */
    uint64_t  seq_len;
} orc_mta;

typedef struct {
    /* .cat  (asindex.c:97-110) */
    uint64_t  n;        /* N = 2G+1 */
    char     *cat;
    /* .mfi  (fmidx.c:221-244) */
    uint64_t  C[256];
    int       o_ratio;
    uint64_t  o_len;
    uint64_t *O;
    char     *bwt;
    int       csa_ratio;
    uint64_t  csa_len;
    uint64_t *csa;
    /* .lch  (lchash.c:106-112) */
    int       hlen;
    uint64_t  lc_len;
    uint64_t *lc;
    /* .sa5 expanded  (sa_use.h:17-46) */
    uint64_t *sa;
    /* .mta */
    int       n_mta;
    orc_mta  *mta;
    /* raw .sa5 bytes, read when `sa` is NULL: lets a test borrow a human-scale index without 8 bytes per row */
    const uint8_t *sa5;
} orc_index;

/* ---- index construction / IO (asindex.c:78-153, fmidx.c:76-199, lchash.c:52-73) ---- */
int   orc_suffix_array(const uint8_t *text, uint64_t n, uint64_t *sa);   /* text[n-1] unique smallest */
orc_index *orc_index_build(int nseq, const char *const *names, const char *const *seqs,
                           const uint64_t *lens, int o_ratio, int hlen);
int   orc_index_write(const orc_index *ix, const char *fasta_path);      /* <fa>.mta .cat .cat.sa5 .cat.mfi .cat.lch */
orc_index *orc_index_load(const char *fasta_path);
void  orc_index_free(orc_index *ix);

/* ---- lookups (fmidx.c:277-313, lchash.c:12-16,36-49,89-104, fmidx.c:18-33) ---- */
uint64_t orc_occ(const orc_index *ix, char c, uint64_t loc);
uint64_t orc_fmi_aln(const orc_index *ix, const char *qry, int len, uint64_t *k, uint64_t *l);
uint64_t orc_lc_aln(const orc_index *ix, const char *qry, int qlen, uint64_t *k, uint64_t *l);
uint64_t orc_sa_access(const orc_index *ix, uint64_t row);
uint64_t orc_csa_access(const orc_index *ix, int ratio, uint64_t loc, int literal);   /* fmidx.c:315-331 */
void orc_csa_access_batch(const orc_index *ix, int ratio, const uint64_t *rows, size_t n, int literal, uint64_t *pos);
void  orc_lc_aln_batch(const orc_index *ix, const char *seeds, int qlen, size_t n,
                       uint64_t *k, uint64_t *l, uint64_t *r);

/* ---- voting (histo.c:9-96) ---- */
typedef struct { uint64_t key, val, bucket; } orc_entry;
typedef struct { orc_entry *e; uint32_t cap, size; } orc_histo;
void     orc_histo_init(orc_histo *h, uint32_t cap);
void     orc_histo_add(orc_histo *h, uint64_t key);
uint64_t orc_histo_top2(const orc_histo *h, orc_entry out[2]);
void     orc_histo_free(orc_histo *h);

/* ---- PART 1 (alnmain.c:333-405) ---- */
typedef struct {
    uint64_t best_key;    /* best.key: voted .cat coordinate of the read start */
    uint64_t best_val;
    int32_t  phases;      /* value of `iter` when the phase loop ended (0..P) */
    int32_t  from_outer;  /* 1 if best came from the outer list (alnmain.c:400-403) */
    uint64_t n_seeds;     /* seeds looked up */
    uint64_t n_hits;      /* SA rows voted */
    uint64_t n_steps;     /* backward-search steps actually executed (occ pairs) */
} orc_vote;
void orc_vote_read(const orc_index *ix, const char *seq, uint32_t len,
                   uint32_t seed_len, uint32_t thres, orc_vote *out);

/* ---- PART 2 (alnmain.c:151-176) ---- */
int orc_seq_lookup(const orc_index *ix, uint64_t loc, uint32_t qlen,
                   int *seq_id, int *strand, uint64_t *off, uint64_t *fwd_loc);
void orc_revcomp(char *seq, uint32_t len);                  /* alnmain.c:27-60 */

/* ---- alignment: GACT stand-in spec and edlib-equivalent NW distance ---- */
#define ORC_GACT_T 320
#define ORC_GACT_O 120
/* ops: 0 '=', 1 'I' (consumes query), 2 'D' (consumes target), 3 'X'  (edlib.h EDLIB_EDOP_*) */
int  orc_gact(const char *q, int qlen, const char *t, int tlen, int T, int O,
              uint8_t *ops, int *n_ops, uint64_t *cells);
int  orc_edit_distance_nw(const char *q, int qlen, const char *t, int tlen, int k); /* edlib.cpp:694-895 semantics */
/* edlib's NW alignment path (EDLIB_TASK_PATH, edlib.cpp:909-1108): ops[0..*n_ops) forward order, codes 0 = 1 I 2 D 3 X */
int  orc_edit_alignment_nw(const char *q, int qlen, const char *t, int tlen, int k, uint8_t *ops, int *n_ops);
int  orc_parse_cigar(const uint8_t *ops, int n_ops, char *buf, size_t buflen);      /* edlib.cpp:262-311, STANDARD */

/* ---- whole path for one batch (alnmain.c:330-479) ---- */
typedef struct {
    uint64_t loc;        /* best.key (alnmain.c:427) */
    uint64_t off;        /* 0-based forward offset inside the sequence (POS-1) */
    int32_t  seq_id;     /* index into mta, -1 if seq_lookup failed */
    int32_t  strand;
    int32_t  flag, mapq, ed, valid;
    int32_t  n_ops;
    int32_t  myers;      /* edlib NW(k=(int)(0.05*len*2)) on the same pair: verification statistic */
    int32_t  phases, from_outer;
} orc_result;
/* reads: n sequences at stride `stride` bytes (refactor_reads_seq layout, alnmain.c:87-103).
 * ops_store: n * ops_stride bytes (alnmain.c:316-320 uses 2*max_read_len). reads are reverse-complemented
 * in place when strand==1, exactly as alnmain.c:433-438 does. */
void orc_map_batch(const orc_index *ix, char *reads, size_t stride, const uint32_t *lens, size_t n,
                   uint32_t seed_len, uint32_t thres, int gact_T, int gact_O, int do_myers,
                   orc_result *res, uint8_t *ops_store, size_t ops_stride,
                   uint64_t *total_cells);

#ifdef __cplusplus
}
#endif
#endif
