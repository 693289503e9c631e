/*
 * include/accseq_host.h -- host-side C helpers around the hot path (libaccseq_host.so, no CUDA needed):
 * seeded synthetic inputs, the reference's on-disk index formats, FASTQ ingest and SAM emit.
 */
#ifndef ACCSEQ_HOST_H
#define ACCSEQ_HOST_H

#include <stdint.h>
#include <stddef.h>
#include <stdio.h>
#include "accseq_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

/* ---- synthetic inputs (distribution of scripts/dna_txt_gen.py:12, seeded) ---- */
void   as_sim_genome(uint64_t seed, uint64_t n, char *out);                 /* upper-case ACGT, no terminator */
void   as_sim_plant_repeats(uint64_t seed, char *g, uint64_t n, int n_families, uint64_t unit_len,
                            int copies, double divergence);
/* n_reads reads sampled uniformly over positions and strands; length uniform in [len_min,len_max]; per-read
 * error rate uniform in [err_min,err_max], split sub_frac substitutions, the rest half insertions half
 * deletions.  out: stride bytes per read, zero padded (refactor_reads_seq layout). Returns reads written. */
size_t as_sim_reads(uint64_t seed, const char *genome, uint64_t glen, size_t n_reads,
                    uint32_t len_min, uint32_t len_max, double err_min, double err_max, double sub_frac,
                    char *out, size_t stride, uint32_t *lens, uint64_t *truth_pos, uint8_t *truth_strand);

/* ---- the reference's index files (Notes.txt:6-25; readers: fmidx.c:246-275, lchash.c:114-127,
 *      alnmain.c:125-140,237-252) ---- */
typedef struct as_host_index {
    as_index_view view;        /* pointers into the owned buffers below */
    uint64_t *C, *O, *lc, *csa;
    char     *bwt, *content;
    uint8_t  *sa5;             /* raw 5-byte entries as on disk */
    as_mta   *mta;
    char    **names;           /* mta_len NUL-terminated sequence names */
    uint64_t  csa_len;
    int       csa_ratio;
} as_host_index;
int  as_index_load(const char *fasta_path, as_host_index **out);   /* <fa>.mta, <fa>.cat{,.mfi,.lch,.sa5} */
void as_index_free(as_host_index *ix);
/* construction side (asindex.c:78-153): the .cat image of a set of sequences, buffers for as_gpu_export,
 * and the writer of the five files */
char *as_cat_build(int nseq, const char *const *seqs, const uint64_t *lens, as_mta *mta, uint64_t *con_len);
as_host_index *as_index_alloc(uint64_t n, int hlen, int o_ratio, int csa_ratio, int mta_len);
void as_index_out_of(as_host_index *ix, as_index_out *out);
int  as_index_write(const char *fasta_path, const as_host_index *ix);

/* ---- SAM emit (alnmain.c:62-75, 485-526) ---- */
int  as_sam_header(FILE *out, const as_host_index *ix, long rg_id);
/* CIGAR text of the column ops, STANDARD M/I/D (parse_cigar contract); returns length or -1 if it does not fit */
int  as_cigar_text(const uint8_t *ops, int n_ops, char *buf, size_t buflen);
int  as_sam_record(FILE *out, const as_host_index *ix, const char *qname, size_t qname_len,
                   const char *seq, const char *qual, uint32_t len, const as_result *r,
                   const uint8_t *ops);
/* the same for batches mapped with AS_F_RLE_CIGAR / AS_F_KEEP_READS (map_flags = the as_params.flags of the batch) */
int  as_cigar_text_rle(const uint8_t *runs, int n_runs, char *buf, size_t buflen);
int  as_sam_record2(FILE *out, const as_host_index *ix, const char *qname, size_t qname_len, const char *seq,
                    const char *qual, uint32_t len, const as_result *r, const uint8_t *ops, unsigned map_flags);

/* ---- 2-bit packing for as_map_batch_packed (layout: include/accseq_b200.h) and the as_result of a compact record ---- */
void as_pack_read(const char *seq, uint32_t len, uint32_t *out, size_t nwords);      /* nwords >= ceil(len/16); the rest is zeroed */
int  as_pack_reads(const char *ascii, size_t stride, const uint32_t *lens, size_t n, uint32_t *packed, size_t wstride);
void as_result_from_compact(const as_result_compact *c, uint32_t len, const as_mta *mta, int mta_len, as_result *r);

/* ---- FASTQ ingest (accaln.c:45-58 reads_load + alnmain.c:87-103 refactor_reads_seq) ---- */
typedef struct as_read_batch {
    size_t    n;
    size_t    stride;          /* max_len + 1 */
    uint32_t  max_len;
    char     *seqs;            /* n * stride, zero padded */
    uint32_t *lens;
    char    **names;
    char    **quals;
} as_read_batch;
typedef struct as_fastq as_fastq;
as_fastq *as_fastq_open(const char *path);       /* plain or gzip */
/* loads up to max_reads records into a fresh batch (seqs from alloc(bytes) if given, else malloc) */
int  as_fastq_next_batch(as_fastq *fq, size_t max_reads, as_read_batch *b, void *(*alloc)(size_t));
void as_read_batch_free(as_read_batch *b, void (*dealloc)(void *));
void as_fastq_close(as_fastq *fq);
/* all records of a FASTA file (plain or gzip) as heap strings the caller frees: the input of accidx (asindex.c:78-117) */
int  as_fasta_load(const char *path, int *nseq, char ***names, char ***seqs, uint64_t **lens);

/* ---- a persistent thread pool whose idle workers sleep (pool.c) ---- */
typedef struct as_pool as_pool;
as_pool *as_pool_new(int threads);
int  as_pool_threads(const as_pool *p);
void as_pool_run(as_pool *p, void (*fn)(int t, int n, void *arg), void *arg);     /* fn(t, n, arg) for every t in [0, n) */
void as_pool_free(as_pool *p);

/* ---- multi-threaded ingest / emit around as_map_batch_packed (ingest.c, emit.c) ---- */
struct as_ingest;
typedef struct as_packed_batch {
    size_t      n;
    uint32_t    max_len;
    size_t      wstride;       /* ceil(max_len / 16) words per read */
    uint32_t   *packed;        /* n * wstride words, from alloc() (pinned when it is as_host_alloc) */
    uint32_t   *lens;          /* n, from alloc() */
    const char **name, **seq, **qual;   /* record text, NOT NUL-terminated: pointers into `text` */
    uint32_t   *name_len;
    const char *text;          /* the block the records live in: a slice of the mapped file, or an inflate buffer */
    size_t      text_len;
    int         text_owned;
    struct as_ingest *owner;   /* the ingest whose recycled buffers these are: it must outlive the batch */
} as_packed_batch;
typedef struct as_ingest as_ingest;
/* plain or gzip 4-line FASTQ; `threads` parse and pack; a batch is the records of about block_bytes of text (0 = 256 MB) */
as_ingest *as_ingest_open(const char *path, int threads, size_t block_bytes);
/* next batch (b->n == 0 at the end).  AS_EIO: not 4-line FASTQ -- use the serial reader (as_fastq_*), which takes kseq's
 * whole grammar */
int  as_ingest_next(as_ingest *in, as_packed_batch *b, void *(*alloc)(size_t), void (*dealloc)(void *));
void as_packed_batch_free(as_packed_batch *b, void (*dealloc)(void *));
void as_ingest_close(as_ingest *in);

typedef struct as_emitter as_emitter;
as_emitter *as_emitter_new(int threads);
void as_emitter_free(as_emitter *e);
/* SAM text (alnmain.c:488-526) of a batch mapped by one or several as_map_batch_packed calls; see emit.c */
int  as_emit_batch(as_emitter *e, const as_host_index *ix, const as_packed_batch *b, const as_result_compact *res, int n_shards,
                   const size_t *shard_begin, const uint8_t *const *shard_runs, const uint64_t *shard_runs_len,
                   const char **seg, size_t *seg_len);

/* Paired-end records (DESIGN.md "Paired-end mode"; the reference's pair_end() is a stub, alnmain.c:552-555): mate 1 then
 * mate 2 of every pair, FLAG |= 0x1, 0x40 / 0x80, 0x8 / 0x20 from the mate, 0x2 when both mates map to one sequence on
 * opposite strands within max_insert; RNEXT / PNEXT / TLEN in the columns the reference prints as `*`, 0, 0. */
int  as_emit_pairs(as_emitter *e, const as_host_index *ix, size_t n, const char *const *name1, const uint32_t *name1_len,
                   const char *const *seq1, const char *const *qual1, const uint32_t *len1, const as_result_compact *res1,
                   const uint8_t *runs1, const char *const *name2, const uint32_t *name2_len, const char *const *seq2,
                   const char *const *qual2, const uint32_t *len2, const as_result_compact *res2, const uint8_t *runs2,
                   long max_insert, const char **out, size_t *out_len);

#ifdef __cplusplus
}
#endif
#endif
