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
static inline char *put_record(char *p, const as_host_index *ix, const char *name, uint32_t name_len, const char *seq,
                               const char *qual, uint32_t len, const as_result_compact *c, const uint8_t *runs) {
    const int seq_id = AS_RC_SEQ_ID(*c);
    const unsigned flag = AS_RC_FLAG(*c);
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
    memcpy(p, "\t*\t0\t0\t", 7);
    p += 7;
    if (AS_RC_STRAND(*c) && seq_id >= 0) {                                    /* strand 1: SEQ is printed reverse-complemented */
        for (uint32_t i = 0; i < len; i++) {
            const unsigned char x = COMP[(unsigned char)seq[len - 1 - i]];
            p[i] = x ? (char)x : 'N';
        }
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
                       j->shard_runs[s] ? j->shard_runs[s] + off : NULL);
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
