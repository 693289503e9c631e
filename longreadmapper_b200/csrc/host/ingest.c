/*
 * ingest.c -- multi-threaded FASTQ ingest straight into the packed batch as_map_batch_packed takes (host tooling, plain C +
 * OpenMP).  Replaces reads_load (accaln.c:45-58, over kseq.h:175-214) + refactor_reads_seq (alnmain.c:87-103) for the
 * common case of 4-line FASTQ: the reference parses one record at a time on one thread and copies every sequence twice;
 * here the file is cut into blocks that end at record boundaries, every thread finds the records of its share of a block
 * (memchr), and then packs them 2 bits per base into pinned memory while the record text stays where it is (the mapped
 * file, or the inflate buffer of a .gz) for the SAM emitter.
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
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>
#include <zlib.h>

struct as_ingest {
    /* plain file: the whole file mapped */
    const char *map;
    size_t map_len, map_pos;
    /* gzip: inflate into blocks */
    gzFile gz;
    char *carry;                /* bytes of an incomplete record left over from the last block */
    size_t carry_len;
    int eof;
    int threads;
    size_t block_bytes;
};

as_ingest *as_ingest_open(const char *path, int threads, size_t block_bytes) {
    if (!path) return NULL;
    as_ingest *in = calloc(1, sizeof *in);
    if (!in) return NULL;
    in->threads = threads > 0 ? threads : 1;
    in->block_bytes = block_bytes ? block_bytes : ((size_t)256 << 20);
    int fd = open(path, O_RDONLY);
    if (fd < 0) {
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
    in->map_len = (size_t)st.st_size;
    void *m = mmap(NULL, in->map_len, PROT_READ, MAP_PRIVATE, fd, 0);
    close(fd);
    if (m == MAP_FAILED) {
        free(in);
        return NULL;
    }
    madvise(m, in->map_len, MADV_SEQUENTIAL);
    in->map = m;
    return in;
}

void as_ingest_close(as_ingest *in) {
    if (!in) return;
    if (in->map) munmap((void *)in->map, in->map_len);
    if (in->gz) gzclose(in->gz);
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

/* the text of the next block: [*blk, *blk + *len) ends at a record boundary; *owned = 1 when the caller frees it */
static int next_block(as_ingest *in, const char **blk, size_t *len, int *owned) {
    *blk = NULL;
    *len = 0;
    *owned = 0;
    if (in->map) {
        if (in->map_pos >= in->map_len) return AS_OK;
        const char *b = in->map + in->map_pos, *fend = in->map + in->map_len;
        size_t want = in->block_bytes;
        const char *e = fend;
        if ((size_t)(fend - b) > want) {
            e = next_header(b + want, b, fend);              /* cut at the first header past the block size */
        }
        *blk = b;
        *len = (size_t)(e - b);
        in->map_pos += *len;
        return AS_OK;
    }
    if (in->eof && in->carry_len == 0) return AS_OK;
    size_t cap = in->block_bytes + in->carry_len + (1 << 20);
    char *buf = malloc(cap + 1);
    if (!buf) return AS_ENOMEM;
    size_t have = in->carry_len;
    if (have) memcpy(buf, in->carry, have);
    in->carry_len = 0;
    while (!in->eof && have < in->block_bytes + (have ? 0 : 0)) {
        const size_t room = cap - have;
        const int r = gzread(in->gz, buf + have, (unsigned)(room > (1u << 30) ? (1u << 30) : room));
        if (r < 0) {
            free(buf);
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
        const char *p = buf + have;
        const char *last = NULL;
        /* walk back over at most a few lines to find the last complete header boundary */
        size_t back = have > (1u << 22) ? have - (1u << 22) : 0;
        const char *q = next_header(buf + back, buf, p);
        while (q < p) {
            last = q;
            const char *nl = memchr(q, '\n', (size_t)(p - q));
            if (!nl) break;
            q = next_header(nl + 1, buf, p);
        }
        if (last && last > buf) cut = (size_t)(last - buf);
        else if (last == buf) cut = 0;
        if (cut == 0) {                                       /* one record larger than the block: let the serial reader have it */
            free(buf);
            return AS_EIO;
        }
        const size_t rest = have - cut;
        char *c2 = realloc(in->carry, rest ? rest : 1);
        if (!c2) {
            free(buf);
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

typedef struct {
    const char *name, *seq, *qual;
    uint32_t name_len, len;
} rec_t;

/* records of [p, end) into a growing array; returns <0 on malformed input */
static int parse_range(const char *p, const char *end, rec_t **out, size_t *n_out) {
    size_t cap = 0, n = 0;
    rec_t *r = NULL;
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
                free(r);
                return AS_ENOMEM;
            }
            r = r2;
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
    free(r);
    return AS_EIO;
}

void as_packed_batch_free(as_packed_batch *b, void (*dealloc)(void *)) {
    if (!b) return;
    if (b->packed) (dealloc ? dealloc : free)(b->packed);
    if (b->lens) (dealloc ? dealloc : free)(b->lens);
    free(b->name);
    free(b->name_len);
    free(b->seq);
    free(b->qual);
    if (b->text_owned) free((void *)b->text);
    memset(b, 0, sizeof *b);
}

int as_ingest_next(as_ingest *in, as_packed_batch *b, void *(*alloc)(size_t), void (*dealloc)(void *)) {
    if (!in || !b) return AS_EINVAL;
    memset(b, 0, sizeof *b);
    const char *blk;
    size_t len;
    int owned;
    int rc = next_block(in, &blk, &len, &owned);
    if (rc != AS_OK || len == 0) {
        if (owned) free((void *)blk);
        return rc;
    }
    b->text = blk;
    b->text_len = len;
    b->text_owned = owned;
    const int T = in->threads;
    rec_t **parts = calloc((size_t)T, sizeof *parts);
    size_t *cnt = calloc((size_t)T + 1, sizeof *cnt);
    const char **cut = calloc((size_t)T + 1, sizeof *cut);
    int bad = 0;
    if (!parts || !cnt || !cut) {
        rc = AS_ENOMEM;
        goto done;
    }
    /* every thread's share starts at the first record header at or after its byte offset */
    cut[0] = blk;
    cut[T] = blk + len;
#pragma omp parallel for num_threads(T) schedule(static, 1)
    for (int t = 1; t < T; t++) cut[t] = next_header(blk + (len / (size_t)T) * (size_t)t, blk, blk + len);
    for (int t = 1; t <= T; t++)
        if (cut[t] < cut[t - 1]) cut[t] = cut[t - 1];
#pragma omp parallel for num_threads(T) schedule(static, 1) reduction(| : bad)
    for (int t = 0; t < T; t++) {
        size_t n = 0;
        if (parse_range(cut[t], cut[t + 1], &parts[t], &n) != AS_OK) bad |= 1;
        cnt[t + 1] = n;
    }
    if (bad) {
        rc = AS_EIO;
        goto done;
    }
    for (int t = 0; t < T; t++) cnt[t + 1] += cnt[t];
    const size_t n = cnt[T];
    b->n = n;
    if (n == 0) goto done;
    uint32_t max_len = 0;
    for (int t = 0; t < T; t++) {
        const size_t m = cnt[t + 1] - cnt[t];
        uint32_t mx = 0;
        for (size_t i = 0; i < m; i++) mx = parts[t][i].len > mx ? parts[t][i].len : mx;
        max_len = mx > max_len ? mx : max_len;
    }
    b->max_len = max_len;
    b->wstride = ((size_t)max_len + 15) / 16;
    if (b->wstride == 0) b->wstride = 1;
    b->packed = alloc ? alloc(n * b->wstride * sizeof(uint32_t)) : malloc(n * b->wstride * sizeof(uint32_t));
    b->lens = alloc ? alloc(n * sizeof(uint32_t)) : malloc(n * sizeof(uint32_t));
    b->name = malloc(n * sizeof *b->name);
    b->name_len = malloc(n * sizeof *b->name_len);
    b->seq = malloc(n * sizeof *b->seq);
    b->qual = malloc(n * sizeof *b->qual);
    if (!b->packed || !b->lens || !b->name || !b->name_len || !b->seq || !b->qual) {
        rc = AS_ENOMEM;
        goto done;
    }
#pragma omp parallel for num_threads(T) schedule(static, 1)
    for (int t = 0; t < T; t++) {
        const size_t base = cnt[t], m = cnt[t + 1] - cnt[t];
        for (size_t i = 0; i < m; i++) {
            const rec_t *r = &parts[t][i];
            const size_t k = base + i;
            b->name[k] = r->name;
            b->name_len[k] = r->name_len;
            b->seq[k] = r->seq;
            b->qual[k] = r->qual;
            b->lens[k] = r->len;
            as_pack_read(r->seq, r->len, b->packed + k * b->wstride, b->wstride);
        }
    }
done:
    if (parts)
        for (int t = 0; t < T; t++) free(parts[t]);
    free(parts);
    free(cnt);
    free(cut);
    if (rc != AS_OK) as_packed_batch_free(b, dealloc);
    return rc;
}
