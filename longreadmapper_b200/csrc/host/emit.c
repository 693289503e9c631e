/*
 * emit.c -- multi-threaded SAM formatting of a packed batch's compact results (host tooling, plain C + the pool of pool.c).
 * Replaces the print loop alnmain.c:488-526 (one fprintf per record, parse_cigar per record) for batches mapped with
 * as_map_batch_packed: every thread formats a contiguous range of reads into its own buffer -- hand-rolled integer and
 * CIGAR printing, memcpy for SEQ/QUAL, the reverse complement of strand-1 reads done here (alnmain.c:27-60,433-438; QUAL is
 * not reversed, alnmain.c:462) -- and the buffers are written in range order, so the bytes equal what as_sam_record2 prints
 * record by record.
 */
#include "accseq_host.h"

#include <stdlib.h>
#include <string.h>

static inline char *put_u64(char *p, uint64_t v) {
    char tmp[24];
    int d = 0;
    do {
        tmp[d++] = (char)('0' + v % 10);
        v /= 10;
    } while (v);
    while (d) *p++ = tmp[--d];
    return p;
}

static inline char *put_i32(char *p, int32_t v) {
    if (v < 0) {
        *p++ = '-';
        return put_u64(p, (uint64_t)(-(int64_t)v));
    }
    return put_u64(p, (uint64_t)v);
}

/* run bytes -> CIGAR text (as_cigar_text_rle); returns the end, or NULL when a byte is malformed */
static inline char *put_cigar(char *p, const uint8_t *runs, uint32_t n) {
    static const char code[4] = {'M', 'I', 'D', '?'};
    uint32_t i = 0;
    while (i < n) {
        const uint32_t cls = runs[i] & 3u;
        uint64_t run = 0;
        if (cls == 3 || (runs[i] >> 2) == 0) return NULL;
        while (i < n && (runs[i] & 3u) == cls) run += runs[i++] >> 2;
        p = put_u64(p, run);
        *p++ = code[cls];
    }
    return p;
}

static const unsigned char COMP[256] = {
    ['A'] = 'T', ['C'] = 'G', ['G'] = 'C', ['T'] = 'A', ['a'] = 'T', ['c'] = 'G', ['g'] = 'C', ['t'] = 'A'};

/* dst[0, len) = reverse complement of src[0, len) (alnmain.c:27-60: either case complements to upper case, anything else
 * prints 'N').  16 bases per step where the CPU has SSSE3: the low nibble of a byte picks the letter it would have to be and
 * its complement out of two 16-entry tables (PSHUFB), a byte whose upper-cased value is not that letter becomes 'N', and one
 * more PSHUFB reverses the 16 bytes.  Same bytes as the table loop on every input. */
static void revcomp_scalar(char *dst, const char *src, uint32_t len) {
    for (uint32_t i = 0; i < len; i++) {
        const unsigned char x = COMP[(unsigned char)src[len - 1 - i]];
        dst[i] = x ? (char)x : 'N';
    }
}

#if defined(__x86_64__) || defined(__i386__)
#include <immintrin.h>
__attribute__((target("ssse3"))) static void revcomp_ssse3(char *dst, const char *src, uint32_t len) {
    /* low nibbles: A/a 1, C/c 3, T/t 4, G/g 7 */
    const __m128i is = _mm_setr_epi8(0, 'A', 0, 'C', 'T', 0, 0, 'G', 0, 0, 0, 0, 0, 0, 0, 0);
    const __m128i co = _mm_setr_epi8('N', 'T', 'N', 'G', 'A', 'N', 'N', 'C', 'N', 'N', 'N', 'N', 'N', 'N', 'N', 'N');
    const __m128i rev = _mm_setr_epi8(15, 14, 13, 12, 11, 10, 9, 8, 7, 6, 5, 4, 3, 2, 1, 0);
    const __m128i lo4 = _mm_set1_epi8(0x0F), up = _mm_set1_epi8((char)0xDF), en = _mm_set1_epi8('N');
    uint32_t i = 0;
    for (; i + 16 <= len; i += 16) {
        const __m128i x = _mm_loadu_si128((const __m128i *)(src + len - 16 - i));
        const __m128i nib = _mm_and_si128(x, lo4);
        const __m128i ok = _mm_cmpeq_epi8(_mm_and_si128(x, up), _mm_shuffle_epi8(is, nib));
        const __m128i c = _mm_or_si128(_mm_and_si128(ok, _mm_shuffle_epi8(co, nib)), _mm_andnot_si128(ok, en));
        _mm_storeu_si128((__m128i *)(dst + i), _mm_shuffle_epi8(c, rev));
    }
    if (i < len) revcomp_scalar(dst + i, src, len - i);      /* the first len - i bases of src, reversed, finish dst */
}
#endif

static void revcomp(char *dst, const char *src, uint32_t len) {
#if defined(__x86_64__) || defined(__i386__)
    static int have = -1;
    if (have < 0) have = __builtin_cpu_supports("ssse3") ? 1 : 0;
    if (have) {
        revcomp_ssse3(dst, src, len);
        return;
    }
#endif
    revcomp_scalar(dst, src, len);
}

typedef struct {
    char *buf;
    size_t len, cap;
} outbuf;

struct as_emitter {
    int threads;
    as_pool *pool;
    outbuf *out;
};

as_emitter *as_emitter_new(int threads) {
    as_emitter *e = calloc(1, sizeof *e);
    if (!e) return NULL;
    e->pool = as_pool_new(threads > 0 ? threads : 1);
    e->threads = as_pool_threads(e->pool);
    e->out = calloc((size_t)e->threads, sizeof *e->out);
    if (!e->out) {
        free(e);
        return NULL;
    }
    return e;
}

void as_emitter_free(as_emitter *e) {
    if (!e) return;
    for (int t = 0; t < e->threads; t++) free(e->out[t].buf);
    as_pool_free(e->pool);
    free(e->out);
    free(e);
}

/* one SAM record of alnmain.c:500-524 at p; the caller guarantees room for 3*len + name + 160 bytes */
/* paired-end fields of a record (DESIGN.md "Paired-end mode"): what FLAG gains, and RNEXT / PNEXT / TLEN */
typedef struct {
    unsigned flag_or;
    const char *rnext;           /* "=", a sequence name, or "*" */
    uint64_t pnext;
    int64_t tlen;
} mate_info;

static inline char *put_record(char *p, const as_host_index *ix, const char *name, uint32_t name_len, const char *seq,
                               const char *qual, uint32_t len, const as_result_compact *c, const uint8_t *runs,
                               const mate_info *mate) {
    const int seq_id = AS_RC_SEQ_ID(*c);
    const unsigned flag = AS_RC_FLAG(*c) | (mate ? mate->flag_or : 0u);
    const uint32_t n_runs = AS_RC_NRUNS(*c);
    memcpy(p, name, name_len);
    p += name_len;
    *p++ = '\t';
    p = put_u64(p, flag);
    *p++ = '\t';
    if (seq_id >= 0 && seq_id < ix->view.mta_len) {
        const size_t l = strlen(ix->names[seq_id]);
        memcpy(p, ix->names[seq_id], l);
        p += l;
        *p++ = '\t';
        p = put_u64(p, (c->pos & 0xFFFFFFFFFFULL) + 1);                   /* off + 1 */
    } else {
        *p++ = '*';                                                       /* seq_lookup failed: the reference prints garbage */
        *p++ = '\t';
        *p++ = '0';
    }
    *p++ = '\t';
    p = put_u64(p, (flag & 4u) ? 0u : 255u);                              /* MAPQ, alnmain.c:466-470 */
    *p++ = '\t';
    char *cig0 = p, *q = NULL;
    if (seq_id >= 0 && n_runs > 0 && runs) q = put_cigar(p, runs, n_runs);
    if (!q || (size_t)(q - cig0) + 1 > 2 * (size_t)len + 16) {           /* the reference's buffer is 2*len (alnmain.c:497) */
        p = cig0;
        *p++ = '*';
    } else {
        p = q;
    }
    if (!mate) {
        memcpy(p, "\t*\t0\t0\t", 7);                                  /* r_name "*", 0L, 0 (alnmain.c:517-519) */
        p += 7;
    } else {
        *p++ = '\t';
        const size_t l = strlen(mate->rnext);
        memcpy(p, mate->rnext, l);
        p += l;
        *p++ = '\t';
        p = put_u64(p, mate->pnext);
        *p++ = '\t';
        if (mate->tlen < 0) {
            *p++ = '-';
            p = put_u64(p, (uint64_t)(-mate->tlen));
        } else {
            p = put_u64(p, (uint64_t)mate->tlen);
        }
        *p++ = '\t';
    }
    if (AS_RC_STRAND(*c) && seq_id >= 0) {                                    /* strand 1: SEQ is printed reverse-complemented */
        revcomp(p, seq, len);
    } else {
        memcpy(p, seq, len);
    }
    p += len;
    *p++ = '\t';
    memcpy(p, qual, len);
    p += len;
    memcpy(p, "\tED:I:", 6);
    p += 6;
    p = put_i32(p, c->ed);
    *p++ = '\n';
    return p;
}

typedef struct {
    as_emitter *e;
    const as_host_index *ix;
    const as_packed_batch *b;
    const as_result_compact *res;
    int n_shards;
    const size_t *shard_begin;
    const uint8_t *const *shard_runs;
    const uint64_t *shard_runs_len;
    size_t *lo;
    uint64_t *sum, *need;
    int rc;
} emit_job;

static void job_sum(int t, int T, void *v) {
    (void)T;
    emit_job *j = v;
    uint64_t s = 0, bytes = 0;
    for (size_t i = j->lo[t]; i < j->lo[t + 1]; i++) {
        s += AS_RC_NRUNS(j->res[i]);
        bytes += 3ull * j->b->lens[i] + j->b->name_len[i] + 160;
    }
    j->sum[t] = s;
    j->need[t] = bytes;
}

static void job_format(int t, int T, void *v) {
    (void)T;
    emit_job *j = v;
    outbuf *o = &j->e->out[t];
    if (o->cap < j->need[t] + 64) {
        free(o->buf);
        o->cap = j->need[t] + j->need[t] / 8 + 4096;
        o->buf = malloc(o->cap);
    }
    o->len = 0;
    if (!o->buf) {
        o->cap = 0;
        __atomic_store_n(&j->rc, AS_ENOMEM, __ATOMIC_RELAXED);
        return;
    }
    /* the run bytes of this range's first read start, inside that read's shard, at (run bytes of all reads before it) -
     * (run bytes of all shards before the read's shard) */
    uint64_t before = 0, shards_before = 0;
    for (int u = 0; u < t; u++) before += j->sum[u];
    int s = 0;
    while (s + 1 < j->n_shards && j->shard_begin[s + 1] <= j->lo[t]) shards_before += j->shard_runs_len[s++];
    uint64_t off = before - shards_before;
    char *p = o->buf;
    const as_packed_batch *b = j->b;
    for (size_t i = j->lo[t]; i < j->lo[t + 1]; i++) {
        while (s + 1 < j->n_shards && j->shard_begin[s + 1] <= i) {       /* crossed into the next shard: its bytes start at 0 */
            s++;
            off = 0;
        }
        p = put_record(p, j->ix, b->name[i], b->name_len[i], b->seq[i], b->qual[i], b->lens[i], &j->res[i],
                       j->shard_runs[s] ? j->shard_runs[s] + off : NULL, NULL);
        off += AS_RC_NRUNS(j->res[i]);
    }
    o->len = (size_t)(p - o->buf);
}

/*
 * Formats reads [0, n) of a batch.  The results of the batch may come from several as_map_batch_packed calls (one per GPU
 * shard): shard s covers reads [shard_begin[s], shard_begin[s+1]), its run bytes lie back to back in shard_runs[s] and
 * are shard_runs_len[s] long (the call's *runs_len).  On return seg[t] / seg_len[t] (t < threads) are the text of
 * consecutive read ranges, to be written in order; they stay valid until the next call.  Returns the number of segments,
 * or a negative code.
 */
int as_emit_batch(as_emitter *e, const as_host_index *ix, const as_packed_batch *b, const as_result_compact *res, int n_shards,
                  const size_t *shard_begin, const uint8_t *const *shard_runs, const uint64_t *shard_runs_len,
                  const char **seg, size_t *seg_len) {
    if (!e || !ix || !b || !res || n_shards < 1 || !shard_begin || !shard_runs || !shard_runs_len || !seg || !seg_len)
        return AS_EINVAL;
    const int T = e->threads;
    emit_job j = {e, ix, b, res, n_shards, shard_begin, shard_runs, shard_runs_len, NULL, NULL, NULL, AS_OK};
    j.lo = malloc(((size_t)T + 1) * sizeof *j.lo);                      /* thread t formats reads [lo[t], lo[t+1]) */
    j.sum = calloc((size_t)T + 1, sizeof *j.sum);
    j.need = calloc((size_t)T + 1, sizeof *j.need);
    if (!j.lo || !j.sum || !j.need) {
        free(j.lo);
        free(j.sum);
        free(j.need);
        return AS_ENOMEM;
    }
    for (int t = 0; t <= T; t++) j.lo[t] = (size_t)(((unsigned __int128)b->n * (unsigned)t) / (unsigned)T);
    as_pool_run(e->pool, job_sum, &j);
    as_pool_run(e->pool, job_format, &j);
    for (int t = 0; t < T; t++) {
        seg[t] = e->out[t].buf;
        seg_len[t] = e->out[t].len;
    }
    free(j.lo);
    free(j.sum);
    free(j.need);
    return j.rc == AS_OK ? T : j.rc;
}

/* ------------------------------------------------------------------------------------------------
 * paired-end records (DESIGN.md "Paired-end mode": the reference's pair_end() is a stub, alnmain.c:552-555; the fields it
 * reserves are result.r_name / r_off, accaln.h:97-99)
 * ---------------------------------------------------------------------------------------------- */
static void mate_fields(const as_host_index *ix, const as_result_compact *me, uint32_t my_len, const as_result_compact *ot,
                        uint32_t ot_len, int first, long max_insert, mate_info *m) {
    const int my_id = AS_RC_SEQ_ID(*me), ot_id = AS_RC_SEQ_ID(*ot);
    const unsigned my_f = AS_RC_FLAG(*me), ot_f = AS_RC_FLAG(*ot);
    const int my_un = (my_f & 4u) != 0 || my_id < 0, ot_un = (ot_f & 4u) != 0 || ot_id < 0;
    unsigned f = 0x1u | (first ? 0x40u : 0x80u);
    if (ot_un) f |= 0x8u;
    else if (AS_RC_STRAND(*ot)) f |= 0x20u;
    m->rnext = "*";
    m->pnext = 0;
    m->tlen = 0;
    if (!ot_un) {
        m->rnext = (!my_un && my_id == ot_id) ? "=" : (ot_id < ix->view.mta_len ? ix->names[ot_id] : "*");
        m->pnext = AS_RC_OFF(*ot) + 1;
    }
    if (!my_un && !ot_un && my_id == ot_id) {
        const int64_t a0 = (int64_t)AS_RC_OFF(*me), a1 = a0 + my_len, b0 = (int64_t)AS_RC_OFF(*ot), b1 = b0 + ot_len;
        const int64_t lo = a0 < b0 ? a0 : b0, hi = a1 > b1 ? a1 : b1;
        const int64_t span = hi - lo;
        /* the leftmost mate carries +span, the other -span; on a tie the first mate is the leftmost */
        const int leftmost = a0 < b0 || (a0 == b0 && first);
        m->tlen = leftmost ? span : -span;
        if (AS_RC_STRAND(*me) != AS_RC_STRAND(*ot) && span <= max_insert) f |= 0x2u;
    }
    m->flag_or = f;
}

/* SAM text of n read pairs, mate 1 then mate 2 of each pair; res1 / res2 and runs1 / runs2 are the results of ONE
 * as_map_batch_packed call per mate file (run bytes back to back).  One buffer, grown as needed; *out / *out_len stay
 * valid until the next call on this emitter. */
int as_emit_pairs(as_emitter *e, const as_host_index *ix, size_t n, const char *const *name1, const uint32_t *name1_len,
                  const char *const *seq1, const char *const *qual1, const uint32_t *len1, const as_result_compact *res1,
                  const uint8_t *runs1, const char *const *name2, const uint32_t *name2_len, const char *const *seq2,
                  const char *const *qual2, const uint32_t *len2, const as_result_compact *res2, const uint8_t *runs2,
                  long max_insert, const char **out, size_t *out_len) {
    if (!e || !ix || !res1 || !res2 || !out || !out_len) return AS_EINVAL;
    uint64_t need = 0;
    for (size_t i = 0; i < n; i++) need += 3ull * len1[i] + name1_len[i] + 3ull * len2[i] + name2_len[i] + 2 * 260;
    outbuf *o = &e->out[0];
    if (o->cap < need + 64) {
        free(o->buf);
        o->cap = need + need / 8 + 4096;
        o->buf = malloc(o->cap);
        if (!o->buf) {
            o->cap = 0;
            return AS_ENOMEM;
        }
    }
    char *p = o->buf;
    uint64_t off1 = 0, off2 = 0;
    for (size_t i = 0; i < n; i++) {
        mate_info m1, m2;
        mate_fields(ix, &res1[i], len1[i], &res2[i], len2[i], 1, max_insert, &m1);
        mate_fields(ix, &res2[i], len2[i], &res1[i], len1[i], 0, max_insert, &m2);
        p = put_record(p, ix, name1[i], name1_len[i], seq1[i], qual1[i], len1[i], &res1[i], runs1 ? runs1 + off1 : NULL, &m1);
        p = put_record(p, ix, name2[i], name2_len[i], seq2[i], qual2[i], len2[i], &res2[i], runs2 ? runs2 + off2 : NULL, &m2);
        off1 += AS_RC_NRUNS(res1[i]);
        off2 += AS_RC_NRUNS(res2[i]);
    }
    o->len = (size_t)(p - o->buf);
    *out = o->buf;
    *out_len = o->len;
    return AS_OK;
}
