/*
 * fastq.c -- FASTQ/FASTA ingest into the batch layout the hot path takes (host tooling, plain C).
 * Replaces reads_load (accaln.c:45-58, over kseq.h:175-214) + refactor_reads_seq (alnmain.c:87-103): records
 * are parsed straight into one zero-filled buffer of stride max_len+1, which is the upload unit of
 * as_map_batch.  Record grammar as kseq reads it: '@'/'>' name [space comment] NL, sequence lines, and for
 * FASTQ '+' NL and quality lines until as many bytes as bases.  Plain or gzip input (zlib gzread).
 */
#include "accseq_host.h"

#include <stdlib.h>
#include <string.h>
#include <zlib.h>

#define FQ_BUF (1 << 20)

struct as_fastq {
    gzFile fp;
    unsigned char *buf;
    int beg, end, eof;
    int pending;             /* a header marker already consumed ('@' or '>'), 0 if none */
};

as_fastq *as_fastq_open(const char *path) {
    if (!path) return NULL;
    gzFile fp = gzopen(path, "rb");
    if (!fp) return NULL;
    gzbuffer(fp, 1 << 18);
    as_fastq *fq = calloc(1, sizeof *fq);
    if (fq) fq->buf = malloc(FQ_BUF);
    if (!fq || !fq->buf) {
        gzclose(fp);
        if (fq) free(fq->buf);
        free(fq);
        return NULL;
    }
    fq->fp = fp;
    return fq;
}

void as_fastq_close(as_fastq *fq) {
    if (!fq) return;
    gzclose(fq->fp);
    free(fq->buf);
    free(fq);
}

static int fq_getc(as_fastq *fq) {
    if (fq->beg >= fq->end) {
        if (fq->eof) return -1;
        fq->beg = 0;
        fq->end = gzread(fq->fp, fq->buf, FQ_BUF);
        if (fq->end <= 0) { fq->eof = 1; fq->end = 0; return -1; }
    }
    return fq->buf[fq->beg++];
}

typedef struct { char *s; size_t l, cap; } dstr;

static int dpush(dstr *d, const unsigned char *p, size_t n) {
    if (d->l + n + 1 > d->cap) {
        size_t nc = d->cap ? d->cap : 256;
        while (nc < d->l + n + 1) nc *= 2;
        char *s2 = realloc(d->s, nc);
        if (!s2) return -1;
        d->s = s2;
        d->cap = nc;
    }
    memcpy(d->s + d->l, p, n);
    d->l += n;
    d->s[d->l] = 0;
    return 0;
}

/* append the rest of the current line (without the terminator) to d; returns the delimiter or -1 at EOF */
static int fq_line(as_fastq *fq, dstr *d) {
    for (;;) {
        if (fq->beg >= fq->end) {
            int c = fq_getc(fq);
            if (c < 0) return -1;
            fq->beg--;
        }
        unsigned char *p = fq->buf + fq->beg;
        unsigned char *nl = memchr(p, '\n', (size_t)(fq->end - fq->beg));
        size_t n = nl ? (size_t)(nl - p) : (size_t)(fq->end - fq->beg);
        if (d && dpush(d, p, n)) return -2;
        fq->beg += (int)n + (nl ? 1 : 0);
        if (nl) {
            if (d && d->l && d->s[d->l - 1] == '\r') d->s[--d->l] = 0;
            return '\n';
        }
    }
}

/* one record; returns 1, 0 at EOF, <0 on error */
static int fq_record(as_fastq *fq, dstr *name, dstr *seq, dstr *qual) {
    int c = fq->pending;
    fq->pending = 0;
    if (!c) {
        while ((c = fq_getc(fq)) >= 0 && c != '>' && c != '@') {}
        if (c < 0) return 0;
    }
    name->l = seq->l = qual->l = 0;
    if (name->s) name->s[0] = 0;
    if (fq_line(fq, name) == -2) return AS_ENOMEM;
    /* kseq: name ends at the first white space */
    for (size_t i = 0; i < name->l; i++)
        if (name->s[i] == ' ' || name->s[i] == '\t') { name->s[i] = 0; name->l = i; break; }
    /* sequence lines */
    for (;;) {
        c = fq_getc(fq);
        if (c < 0 || c == '>' || c == '@' || c == '+') break;
        if (c == '\n' || c == '\r') continue;
        fq->beg--;
        int r = fq_line(fq, seq);
        if (r == -2) return AS_ENOMEM;
        if (r < 0) { c = -1; break; }
    }
    if (c == '>' || c == '@') fq->pending = c;
    if (!seq->s && dpush(seq, (const unsigned char *)"", 0)) return AS_ENOMEM;
    if (c != '+') return 1;                               /* FASTA record */
    if (fq_line(fq, NULL) < 0) return AS_EIO;             /* rest of the '+' line */
    while (qual->l < seq->l) {
        int r = fq_line(fq, qual);
        if (r == -2) return AS_ENOMEM;
        if (r < 0) break;
    }
    if (qual->l != seq->l) return AS_EIO;                 /* kseq returns -2: truncated quality */
    return 1;
}

void as_read_batch_free(as_read_batch *b, void (*dealloc)(void *)) {
    if (!b) return;
    if (b->seqs) (dealloc ? dealloc : free)(b->seqs);
    free(b->lens);
    for (size_t i = 0; i < b->n; i++) {
        if (b->names) free(b->names[i]);
        if (b->quals) free(b->quals[i]);
    }
    free(b->names);
    free(b->quals);
    memset(b, 0, sizeof *b);
}

int as_fastq_next_batch(as_fastq *fq, size_t max_reads, as_read_batch *b, void *(*alloc)(size_t)) {
    if (!fq || !b) return AS_EINVAL;
    memset(b, 0, sizeof *b);
    if (max_reads == 0) return AS_OK;
    dstr name = {0}, seq = {0}, qual = {0};
    size_t cap = max_reads < 4096 ? max_reads : 4096;
    char **seqs = malloc(sizeof(char *) * cap);
    b->names = malloc(sizeof(char *) * cap);
    b->quals = malloc(sizeof(char *) * cap);
    b->lens = malloc(sizeof(uint32_t) * cap);
    int rc = AS_OK;
    if (!seqs || !b->names || !b->quals || !b->lens) { rc = AS_ENOMEM; goto done; }
    while (b->n < max_reads) {
        int r = fq_record(fq, &name, &seq, &qual);
        if (r == 0) break;
        if (r < 0) { rc = r; break; }
        if (b->n == cap) {
            size_t nc = cap * 2 < max_reads ? cap * 2 : max_reads;
            char **s2 = realloc(seqs, sizeof(char *) * nc);
            if (s2) seqs = s2;
            char **n2 = realloc(b->names, sizeof(char *) * nc);
            if (n2) b->names = n2;
            char **q2 = realloc(b->quals, sizeof(char *) * nc);
            if (q2) b->quals = q2;
            uint32_t *l2 = realloc(b->lens, sizeof(uint32_t) * nc);
            if (l2) b->lens = l2;
            if (!s2 || !n2 || !q2 || !l2) { rc = AS_ENOMEM; break; }
            cap = nc;
        }
        size_t i = b->n;
        seqs[i] = malloc(seq.l + 1);
        b->names[i] = malloc(name.l + 1);
        b->quals[i] = malloc(qual.l + 1);
        if (!seqs[i] || !b->names[i] || !b->quals[i]) {
            free(seqs[i]); free(b->names[i]); free(b->quals[i]);
            rc = AS_ENOMEM;
            break;
        }
        memcpy(seqs[i], seq.s, seq.l + 1);
        memcpy(b->names[i], name.s ? name.s : "", name.l + 1);
        memcpy(b->quals[i], qual.s ? qual.s : "", qual.l);
        b->quals[i][qual.l] = 0;
        b->lens[i] = (uint32_t)seq.l;
        if (seq.l > b->max_len) b->max_len = (uint32_t)seq.l;
        b->n++;
    }
    if (rc == AS_OK && b->n) {
        /* refactor_reads_seq: one zero-filled buffer, stride max_len+1 (alnmain.c:87-103) */
        b->stride = (size_t)b->max_len + 1;
        size_t bytes = b->stride * b->n;
        b->seqs = alloc ? alloc(bytes) : malloc(bytes);
        if (!b->seqs) rc = AS_ENOMEM;
        else {
            memset(b->seqs, 0, bytes);
            for (size_t i = 0; i < b->n; i++) memcpy(b->seqs + i * b->stride, seqs[i], b->lens[i]);
        }
    }
done:
    if (seqs) {
        for (size_t i = 0; i < b->n; i++) free(seqs[i]);
        free(seqs);
    }
    free(name.s); free(seq.s); free(qual.s);
    if (rc != AS_OK) {
        as_read_batch_free(b, NULL);
        return rc;
    }
    return AS_OK;
}

/* every record of a FASTA (or FASTQ) file as separate heap strings: what create_meta (asindex.c:78-117) iterates over */
int as_fasta_load(const char *path, int *nseq, char ***names, char ***seqs, uint64_t **lens) {
    if (!path || !nseq || !names || !seqs || !lens) return AS_EINVAL;
    as_fastq *fq = as_fastq_open(path);
    if (!fq) return AS_EIO;
    dstr name = {0}, seq = {0}, qual = {0};
    int n = 0, cap = 0, rc = AS_OK;
    char **nm = NULL, **sq = NULL;
    uint64_t *ln = NULL;
    for (;;) {
        int r = fq_record(fq, &name, &seq, &qual);
        if (r == 0) break;
        if (r < 0) { rc = r; break; }
        if (n == cap) {
            cap = cap ? 2 * cap : 16;
            nm = realloc(nm, sizeof(char *) * (size_t)cap);
            sq = realloc(sq, sizeof(char *) * (size_t)cap);
            ln = realloc(ln, sizeof(uint64_t) * (size_t)cap);
            if (!nm || !sq || !ln) { rc = AS_ENOMEM; break; }
        }
        nm[n] = malloc(name.l + 1);
        if (!nm[n]) { rc = AS_ENOMEM; break; }
        memcpy(nm[n], name.s ? name.s : "", name.l);
        nm[n][name.l] = 0;
        sq[n] = seq.s;                    /* hand the buffer over instead of copying a chromosome */
        ln[n] = seq.l;
        seq.s = NULL;
        seq.l = seq.cap = 0;
        n++;
    }
    free(name.s); free(seq.s); free(qual.s);
    as_fastq_close(fq);
    if (rc != AS_OK) {
        for (int i = 0; i < n; i++) { free(nm[i]); free(sq[i]); }
        free(nm); free(sq); free(ln);
        return rc;
    }
    *nseq = n; *names = nm; *seqs = sq; *lens = ln;
    return AS_OK;
}
