/*
 * ingest.c -- multi-threaded FASTQ ingest straight into the packed batch as_map_batch_packed takes (host tooling, plain C +
 * the thread pool of pool.c).  Replaces reads_load (accaln.c:45-58, over kseq.h:175-214) + refactor_reads_seq (alnmain.c:87-103) for the
 * common case of 4-line FASTQ: the reference parses one record at a time on one thread and copies every sequence twice;
 * here the file is cut into blocks that end at record boundaries, every thread finds the records of its share of a block
 * (memchr), and then packs them 2 bits per base into pinned memory while the record text stays where it is (the block
 * buffer) for the SAM emitter.
 *
 * A record is recognised the way kseq does for well-formed input: '@' name [space comment] NL, one sequence line, '+' line,
 * one quality line of the same length.  Anything else (wrapped sequences, FASTA, a truncated record) makes the ingest
 * return AS_EIO and the caller falls back to the serial reader of fastq.c, which handles the general grammar.
 */
#define _GNU_SOURCE
#include "accseq_host.h"

#include <fcntl.h>
#include <stdlib.h>
#include <string.h>
#include <sys/stat.h>
#include <unistd.h>
#include <zlib.h>
#include <pthread.h>

struct as_ingest {
    /* plain file: blocks are read with pread, every thread its share (page faults on a shared mapping serialise on the
     * address-space lock; measured 2.5 M reads/s per thread through mmap) */
    int fd;
    size_t file_len, file_pos;
    /* gzip: inflate into blocks */
    gzFile gz;
    char *carry;                /* bytes of an incomplete record left over from the last block */
    size_t carry_len;
    int eof;
    int threads;
    as_pool *pool;
    size_t block_bytes;
    /* Block buffers and record arrays are recycled: a fresh 64 MB allocation is 16 k page faults and as many pages zeroed by
     * the kernel, which cost more than copying the file into it (measured: 0.5 s of 0.95 s per 2 M reads on one thread). */
    pthread_mutex_t mu;
    void *cache[16];
    struct rec_s **parts_keep;  /* per-thread record arrays, kept between blocks */
    size_t *parts_cap;
};

/* a buffer of at least `bytes` with its capacity in a 64-byte header; recycled through in->cache */
static void *buf_get(as_ingest *in, size_t bytes) {
    void *b = NULL;
    pthread_mutex_lock(&in->mu);
    int best = -1;
    for (int i = 0; i < 16; i++) {
        if (!in->cache[i]) continue;
        const size_t cap = *(size_t *)in->cache[i];
        if (cap >= bytes && cap <= 4 * bytes + (1 << 20) && (best < 0 || cap < *(size_t *)in->cache[best])) best = i;
    }
    if (best >= 0) {
        b = in->cache[best];
        in->cache[best] = NULL;
    }
    pthread_mutex_unlock(&in->mu);
    if (!b) {
        const size_t cap = bytes + bytes / 4 + 4096;
        b = malloc(cap + 64);
        if (!b) return NULL;
        *(size_t *)b = cap;
    }
    return (char *)b + 64;
}

static void buf_put(as_ingest *in, const void *p) {
    if (!p) return;
    void *b = (char *)p - 64;
    if (in) {
        pthread_mutex_lock(&in->mu);
        for (int i = 0; i < 16; i++)
            if (!in->cache[i]) {
                in->cache[i] = b;
                b = NULL;
                break;
            }
        pthread_mutex_unlock(&in->mu);
    }
    free(b);
}

as_ingest *as_ingest_open(const char *path, int threads, size_t block_bytes) {
    if (!path) return NULL;
    as_ingest *in = calloc(1, sizeof *in);
    if (!in) return NULL;
    in->threads = threads > 0 ? threads : 1;
    in->pool = as_pool_new(in->threads);
    in->threads = as_pool_threads(in->pool);
    pthread_mutex_init(&in->mu, NULL);
    in->parts_keep = calloc((size_t)in->threads, sizeof *in->parts_keep);
    in->parts_cap = calloc((size_t)in->threads, sizeof *in->parts_cap);
    in->block_bytes = block_bytes ? block_bytes : ((size_t)64 << 20);
    in->fd = -1;
    int fd = open(path, O_RDONLY);
    if (fd < 0) {
        as_pool_free(in->pool);
        free(in);
        return NULL;
    }
    unsigned char magic[2] = {0, 0};
    ssize_t got = read(fd, magic, 2);
    struct stat st;
    if (got == 2 && magic[0] == 0x1f && magic[1] == 0x8b) {
        close(fd);
        in->gz = gzopen(path, "rb");
        if (!in->gz) {
            free(in);
            return NULL;
        }
        gzbuffer(in->gz, 1 << 20);
        return in;
    }
    if (fstat(fd, &st) != 0 || st.st_size == 0) {
        close(fd);
        in->eof = 1;
        return in;                                   /* empty file: no records */
    }
    in->file_len = (size_t)st.st_size;
    in->fd = fd;
    posix_fadvise(fd, 0, 0, POSIX_FADV_SEQUENTIAL);
    return in;
}

void as_ingest_close(as_ingest *in) {
    if (!in) return;
    if (in->fd >= 0) close(in->fd);
    if (in->gz) gzclose(in->gz);
    as_pool_free(in->pool);
    for (int i = 0; i < 16; i++) free(in->cache[i]);
    if (in->parts_keep)
        for (int t = 0; t < in->threads; t++) free(in->parts_keep[t]);
    free(in->parts_keep);
    free(in->parts_cap);
    free(in->carry);
    free(in);
}

/* is `p` (the first byte of a line) the header of a 4-line record?  '@', and the line after next starts with '+' */
static int is_header(const char *p, const char *end) {
    if (p >= end || *p != '@') return 0;
    const char *l1 = memchr(p, '\n', (size_t)(end - p));
    if (!l1) return 0;
    const char *l2 = memchr(l1 + 1, '\n', (size_t)(end - l1 - 1));
    if (!l2 || l2 + 1 >= end) return 0;
    return l2[1] == '+';
}

/* first record header at or after p (p need not be a line start) */
static const char *next_header(const char *p, const char *begin, const char *end) {
    if (p == begin && is_header(p, end)) return p;
    while (p < end) {
        const char *nl = memchr(p, '\n', (size_t)(end - p));
        if (!nl) return end;
        p = nl + 1;
        if (is_header(p, end)) return p;
    }
    return end;
}

/* offset of the last record header in buf[0, have): the text before it is whole records (0: none found) */
static size_t last_header(const char *buf, size_t have) {
    const char *p = buf + have, *last = NULL;
    for (size_t win = (size_t)1 << 16; !last; win <<= 2) {     /* look back 64 KB, then 256 KB, ... for long reads */
        const size_t back = have > win ? have - win : 0;
        const char *q = next_header(buf + back, buf, p);
        while (q < p) {
            last = q;
            const char *nl = memchr(q, '\n', (size_t)(p - q));
            if (!nl) break;
            q = next_header(nl + 1, buf, p);
        }
        if (back == 0) break;
    }
    return last ? (size_t)(last - buf) : 0;
}

typedef struct {
    int fd;
    char *buf;
    size_t pos, len;
    int bad;
} read_job;

static void job_read(int t, int T, void *v) {
    read_job *j = v;
    size_t a = (j->len / (size_t)T) * (size_t)t, b = t + 1 == T ? j->len : (j->len / (size_t)T) * (size_t)(t + 1);
    while (a < b) {
        const ssize_t r = pread(j->fd, j->buf + a, b - a, (off_t)(j->pos + a));
        if (r <= 0) {
            __atomic_store_n(&j->bad, 1, __ATOMIC_RELAXED);
            return;
        }
        a += (size_t)r;
    }
}

/* the text of the next block: [*blk, *blk + *len) ends at a record boundary; *owned = 1 when the caller frees it */
static int next_block(as_ingest *in, const char **blk, size_t *len, int *owned) {
    *blk = NULL;
    *len = 0;
    *owned = 0;
    if (in->fd >= 0) {
        if (in->file_pos >= in->file_len) return AS_OK;
        for (size_t block = in->block_bytes;; block *= 2) {   /* a record larger than the block: read twice as much */
            size_t want = in->file_len - in->file_pos;
            const int tail = want <= block + (block >> 3);    /* the rest of the file fits: take it whole */
            if (!tail) want = block;
            char *buf = buf_get(in, want + 1);
            if (!buf) return AS_ENOMEM;
            read_job rj = {in->fd, buf, in->file_pos, want, 0};
            as_pool_run(in->pool, job_read, &rj);
            if (rj.bad) {
                buf_put(in, buf);
                return AS_EIO;
            }
            const size_t cut = tail ? want : last_header(buf, want);
            if (cut == 0) {
                buf_put(in, buf);
                continue;
            }
            in->file_pos += cut;
            *blk = buf;
            *len = cut;
            *owned = 1;
            return AS_OK;
        }
    }
    if (in->eof && in->carry_len == 0) return AS_OK;
    size_t cap = in->block_bytes + in->carry_len + (1 << 20);
    char *buf = buf_get(in, cap + 1);
    if (!buf) return AS_ENOMEM;
    size_t have = in->carry_len;
    if (have) memcpy(buf, in->carry, have);
    in->carry_len = 0;
    while (!in->eof && have < in->block_bytes) {
        const size_t room = cap - have;
        const int r = gzread(in->gz, buf + have, (unsigned)(room > (1u << 30) ? (1u << 30) : room));
        if (r < 0) {
            buf_put(in, buf);
            return AS_EIO;
        }
        if (r == 0) {
            in->eof = 1;
            break;
        }
        have += (size_t)r;
        if (have + (1 << 16) > cap) break;
    }
    size_t cut = have;
    if (!in->eof) {
        /* keep the tail from the last header on: that record may be incomplete */
        cut = last_header(buf, have);
        if (cut == 0) {                                       /* one record larger than the block: let the serial reader have it */
            buf_put(in, buf);
            return AS_EIO;
        }
        const size_t rest = have - cut;
        char *c2 = realloc(in->carry, rest ? rest : 1);
        if (!c2) {
            buf_put(in, buf);
            return AS_ENOMEM;
        }
        in->carry = c2;
        memcpy(in->carry, buf + cut, rest);
        in->carry_len = rest;
    }
    *blk = buf;
    *len = cut;
    *owned = 1;
    return AS_OK;
}

typedef struct rec_s {
    const char *name, *seq, *qual;
    uint32_t name_len, len;
} rec_t;

/* records of [p, end) into a growing array; returns <0 on malformed input */
static int parse_range(const char *p, const char *end, rec_t **out, size_t *cap_io, size_t *n_out) {
    size_t cap = *cap_io, n = 0;
    rec_t *r = *out;
    while (p < end) {
        if (*p == '\n' || *p == '\r') {                       /* blank line between records */
            p++;
            continue;
        }
        if (*p != '@') goto bad;
        const char *l0 = memchr(p, '\n', (size_t)(end - p));
        if (!l0) goto bad;
        const char *l1 = memchr(l0 + 1, '\n', (size_t)(end - l0 - 1));
        if (!l1 || l1 + 1 >= end || l1[1] != '+') goto bad;
        const char *l2 = memchr(l1 + 1, '\n', (size_t)(end - l1 - 1));
        if (!l2) goto bad;
        const char *l3 = memchr(l2 + 1, '\n', (size_t)(end - l2 - 1));
        const char *qend = l3 ? l3 : end;                     /* the last record may lack the final newline */
        size_t slen = (size_t)(l1 - l0 - 1), qlen = (size_t)(qend - l2 - 1);
        if (slen && l0[slen] == '\r') slen--;
        if (qlen && l2[qlen] == '\r') qlen--;
        if (slen != qlen || slen > 0xFFFFFFF0u) goto bad;
        if (n == cap) {
            cap = cap ? 2 * cap : 4096;
            rec_t *r2 = realloc(r, cap * sizeof *r);
            if (!r2) {
                *out = r;
                return AS_ENOMEM;
            }
            r = r2;
            *cap_io = cap;
        }
        size_t nl = (size_t)(l0 - p - 1);                     /* kseq: the name ends at the first white space */
        if (nl && p[nl] == '\r') nl--;
        const char *sp = memchr(p + 1, ' ', nl);
        const char *tb = memchr(p + 1, '\t', sp ? (size_t)(sp - p - 1) : nl);
        if (tb) sp = tb;
        r[n].name = p + 1;
        r[n].name_len = (uint32_t)(sp ? (size_t)(sp - p - 1) : nl);
        r[n].seq = l0 + 1;
        r[n].qual = l2 + 1;
        r[n].len = (uint32_t)slen;
        n++;
        p = l3 ? l3 + 1 : end;
    }
    *out = r;
    *n_out = n;
    return AS_OK;
bad:
    *out = r;
    return AS_EIO;
}

void as_packed_batch_free(as_packed_batch *b, void (*dealloc)(void *)) {
    if (!b) return;
    if (b->packed) (dealloc ? dealloc : free)(b->packed);
    if (b->lens) (dealloc ? dealloc : free)(b->lens);
    as_ingest *in = b->owner;                             /* the buffers go back to the ingest that made them */
    buf_put(in, b->name);
    buf_put(in, b->name_len);
    buf_put(in, b->seq);
    buf_put(in, b->qual);
    if (b->text_owned) buf_put(in, b->text);
    memset(b, 0, sizeof *b);
}

typedef struct {
    const char *blk;
    size_t len;
    const char **cut;
    rec_t **parts;
    size_t *parts_cap;
    size_t *cnt;
    uint32_t *mx;
    int bad;
    as_packed_batch *b;
} parse_job;

static void job_cut(int t, int T, void *v) {                  /* every thread's share starts at the first header at or after its offset */
    parse_job *j = v;
    if (t > 0) j->cut[t] = next_header(j->blk + (j->len / (size_t)T) * (size_t)t, j->blk, j->blk + j->len);
}

static void job_parse(int t, int T, void *v) {
    (void)T;
    parse_job *j = v;
    size_t n = 0;
    if (parse_range(j->cut[t], j->cut[t + 1], &j->parts[t], &j->parts_cap[t], &n) != AS_OK) __atomic_store_n(&j->bad, 1, __ATOMIC_RELAXED);
    j->cnt[t + 1] = n;
    uint32_t mx = 0;
    for (size_t i = 0; i < n; i++) mx = j->parts[t][i].len > mx ? j->parts[t][i].len : mx;
    j->mx[t] = mx;
}

static void job_fill(int t, int T, void *v) {
    (void)T;
    parse_job *j = v;
    as_packed_batch *b = j->b;
    const size_t base = j->cnt[t], m = j->cnt[t + 1] - j->cnt[t];
    for (size_t i = 0; i < m; i++) {
        const rec_t *r = &j->parts[t][i];
        const size_t k = base + i;
        b->name[k] = r->name;
        b->name_len[k] = r->name_len;
        b->seq[k] = r->seq;
        b->qual[k] = r->qual;
        b->lens[k] = r->len;
        as_pack_read(r->seq, r->len, b->packed + k * b->wstride, b->wstride);
    }
}

int as_ingest_next(as_ingest *in, as_packed_batch *b, void *(*alloc)(size_t), void (*dealloc)(void *)) {
    if (!in || !b) return AS_EINVAL;
    memset(b, 0, sizeof *b);
    const char *blk;
    size_t len;
    int owned;
    int rc = next_block(in, &blk, &len, &owned);
    if (rc != AS_OK || len == 0) {
        if (owned) buf_put(in, blk);
        return rc;
    }
    b->text = blk;
    b->text_len = len;
    b->text_owned = owned;
    b->owner = in;
    const int T = in->threads;
    parse_job j;
    memset(&j, 0, sizeof j);
    j.blk = blk;
    j.len = len;
    j.b = b;
    j.parts = in->parts_keep;
    j.parts_cap = in->parts_cap;
    j.cnt = calloc((size_t)T + 1, sizeof *j.cnt);
    j.cut = calloc((size_t)T + 1, sizeof *j.cut);
    j.mx = calloc((size_t)T, sizeof *j.mx);
    if (!j.parts || !j.cnt || !j.cut || !j.mx) {
        rc = AS_ENOMEM;
        goto done;
    }
    j.cut[0] = blk;
    j.cut[T] = blk + len;
    as_pool_run(in->pool, job_cut, &j);
    for (int t = 1; t <= T; t++)
        if (j.cut[t] < j.cut[t - 1]) j.cut[t] = j.cut[t - 1];
    as_pool_run(in->pool, job_parse, &j);
    if (j.bad) {
        rc = AS_EIO;
        goto done;
    }
    for (int t = 0; t < T; t++) j.cnt[t + 1] += j.cnt[t];
    const size_t n = j.cnt[T];
    b->n = n;
    if (n == 0) goto done;
    uint32_t max_len = 0;
    for (int t = 0; t < T; t++) max_len = j.mx[t] > max_len ? j.mx[t] : max_len;
    b->max_len = max_len;
    b->wstride = ((size_t)max_len + 15) / 16;
    if (b->wstride == 0) b->wstride = 1;
    b->packed = alloc ? alloc(n * b->wstride * sizeof(uint32_t)) : malloc(n * b->wstride * sizeof(uint32_t));
    b->lens = alloc ? alloc(n * sizeof(uint32_t)) : malloc(n * sizeof(uint32_t));
    b->name = buf_get(in, n * sizeof *b->name);
    b->name_len = buf_get(in, n * sizeof *b->name_len);
    b->seq = buf_get(in, n * sizeof *b->seq);
    b->qual = buf_get(in, n * sizeof *b->qual);
    if (!b->packed || !b->lens || !b->name || !b->name_len || !b->seq || !b->qual) {
        rc = AS_ENOMEM;
        goto done;
    }
    as_pool_run(in->pool, job_fill, &j);
done:
    free(j.cnt);
    free(j.cut);
    free(j.mx);
    if (rc != AS_OK) as_packed_batch_free(b, dealloc);
    return rc;
}
