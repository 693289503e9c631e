/*
 * accseq_compat.c -- libaccseq_compat.so: the reference's scalar C surface (include/accseq_compat.h) as thin host
 * wrappers over the batched C ABI of libaccseq_b200.so.  See the header for what runs on the GPU and what is plain
 * container / file code.  Written from the reference's headers and call sites (cited per function), not from its bodies.
 */
#include "accseq_compat.h"
#include "accseq_b200.h"
#include "accseq_host.h"

#include <stdio.h>
#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------------------------------------
 * device copies: one as_gpu per dna_fmi, keyed by the bwt pointer (the one member every dna_fmi owns)
 * ---------------------------------------------------------------------------------------------- */
#define MAX_SLOTS 16
typedef struct {
    const char *bwt;            /* key */
    const uint64_t *lc;         /* lc table the handle was opened with (NULL: none) */
    const void *sa;             /* sa_buf->mem the handle was opened with (NULL: none) */
    const uint64_t *csa;        /* csa attached */
    int built;                  /* the handle fmi_build kept: it owns every table of the index, lc at hlen 12 included */
    as_gpu *h;
} slot;
static slot g_slot[MAX_SLOTS];
static int g_device = 0;
static as_gpu *g_align = NULL;  /* index-less handle: cigar_align / simple_gact, and sa_access on the global sa_buf */
static const void *g_align_sa = NULL;

sa_mem *sa_buf = NULL;          /* fmidx.c:16 */

void as_compat_set_device(int device) { g_device = device; }

static void die(const char *what, int rc) {
    fprintf(stderr, "libaccseq_compat: %s failed (%d): %s -- this library has no CPU fallback\n", what, rc, as_last_error());
    abort();
}

void as_compat_release(void) {
    for (int i = 0; i < MAX_SLOTS; i++) {
        if (g_slot[i].h) as_gpu_close(g_slot[i].h);
        memset(&g_slot[i], 0, sizeof g_slot[i]);
    }
    if (g_align) as_gpu_close(g_align);
    g_align = NULL;
    g_align_sa = NULL;
}

static slot *slot_find(const char *bwt) {
    for (int i = 0; i < MAX_SLOTS; i++)
        if (g_slot[i].h && g_slot[i].bwt == bwt) return &g_slot[i];
    return NULL;
}

static slot *slot_new(const char *bwt) {
    for (int i = 0; i < MAX_SLOTS; i++)
        if (!g_slot[i].h) {
            memset(&g_slot[i], 0, sizeof g_slot[i]);
            g_slot[i].bwt = bwt;
            return &g_slot[i];
        }
    /* table full: recycle the first slot */
    as_gpu_close(g_slot[0].h);
    memset(&g_slot[0], 0, sizeof g_slot[0]);
    g_slot[0].bwt = bwt;
    return &g_slot[0];
}

/* the handle of `fmi`, (re)opened so that it carries `lc` and the global sa_buf when they are wanted */
static as_gpu *handle_for(const dna_fmi *fmi, const lc_hash *lch, int want_sa) {
    const void *sa = (want_sa && sa_buf && sa_buf->mem && sa_buf->len >= fmi->length) ? (const void *)sa_buf->mem : NULL;
    slot *s = slot_find(fmi->bwt);
    if (s && (!lch || s->lc == lch->lc) && (!sa || s->sa == sa || s->built)) return s->h;
    const uint64_t *lc = lch ? lch->lc : (s ? s->lc : NULL);
    if (!sa && s) sa = s->sa;
    if (s) {
        as_gpu_close(s->h);
        s->h = NULL;
    } else {
        s = slot_new(fmi->bwt);
    }
    as_index_view v;
    memset(&v, 0, sizeof v);
    v.length = fmi->length;
    v.C = fmi->c;
    v.O = fmi->o;
    v.o_len = fmi->o_len;
    v.o_ratio = fmi->o_ratio;
    v.bwt = fmi->bwt;
    if (lc) {
        v.lc = lc;
        v.hlen = lch ? lch->hlen : 0;
        v.lc_len = lch ? lch->len : 0;
    }
    if (sa) {
        v.sa = sa;
        v.sa_entry_bytes = (int)sizeof(ui40_t);         /* 8: the padded struct (sa_use.h:17-20) */
    }
    as_open_opts o;
    memset(&o, 0, sizeof o);
    o.lc2_hlen = 1;                                     /* scalar lookups never use the deep table */
    o.no_sax = 1;
    int rc = as_gpu_open_opts(&s->h, g_device, &v, &o);
    if (rc != AS_OK) die("as_gpu_open", rc);
    s->bwt = fmi->bwt;
    s->lc = lc;
    s->sa = sa;
    s->csa = NULL;
    s->built = 0;
    return s->h;
}

/* ------------------------------------------------------------------------------------------------
 * fmidx.h
 * ---------------------------------------------------------------------------------------------- */
static char *cat2(const char *a, const char *b) {
    char *p = malloc(strlen(a) + strlen(b) + 1);
    if (!p) abort();
    strcpy(p, a);
    strcat(p, b);
    return p;
}

static FILE *open2(const char *a, const char *b, const char *mode) {
    char *p = cat2(a, b);
    FILE *fp = fopen(p, mode);
    if (!fp) {
        fprintf(stderr, "libaccseq_compat: cannot open %s\n", p);
        abort();
    }
    free(p);
    return fp;
}

static char *slurp(const char *path, uint64_t *len) {
    FILE *fp = open2(path, "", "rb");
    fseek(fp, 0, SEEK_END);
    long n = ftell(fp);
    fseek(fp, 0, SEEK_SET);
    char *buf = malloc((size_t)n + 1);
    if (!buf || fread(buf, 1, (size_t)n, fp) != (size_t)n) abort();
    buf[n] = 0;
    fclose(fp);
    *len = (uint64_t)n;
    return buf;
}

/* fmidx.c:76-199 fmi_build: suffix array -> <prefix>.sa5, C, BWT, O at o_ratio, CSA at ratio 4 -- all on the GPU
 * (as_gpu_build), returned in the reference's layouts (as_gpu_export).  ram_use is pSAscan's budget: unused. */
void fmi_build(const char *prefix, dna_fmi *idx, int o_ratio, long ram_use) {
    (void)ram_use;
    uint64_t n = 0;
    char *text = slurp(prefix, &n);
    as_gpu *h = NULL;
    as_open_opts o;
    memset(&o, 0, sizeof o);
    o.lc2_hlen = 1;
    o.no_sax = 1;
    int rc = as_gpu_build_opts(&h, g_device, text, n, NULL, 0, 12, &o);
    if (rc != AS_OK) die("as_gpu_build", rc);
    memset(idx, 0, sizeof *idx);
    idx->length = n;
    idx->o_ratio = o_ratio;
    idx->o_len = 4 * (n / (uint64_t)o_ratio + 1);
    idx->csa_ratio = 4;                                 /* fmidx.c:194 */
    idx->csa_len = n / 4 + 1;
    idx->c = calloc(256, sizeof(uint64_t));
    idx->o = calloc(idx->o_len, sizeof(uint64_t));
    idx->csa = calloc(idx->csa_len, sizeof(uint64_t));
    idx->bwt = calloc(n + 1, 1);
    uint8_t *sa5 = malloc(5 * n);
    if (!idx->c || !idx->o || !idx->csa || !idx->bwt || !sa5) abort();
    as_index_out out;
    memset(&out, 0, sizeof out);
    out.C = idx->c; out.O = idx->o; out.o_ratio = o_ratio; out.bwt = idx->bwt;
    out.sa = sa5; out.sa_entry_bytes = 5; out.csa = idx->csa; out.csa_ratio = 4;
    rc = as_gpu_export(h, &out);
    if (rc != AS_OK) die("as_gpu_export", rc);
    FILE *fp = open2(prefix, ".sa5", "wb");             /* psascan/sa_use.cc:8-18 */
    if (fwrite(sa5, 5, n, fp) != n) abort();
    fclose(fp);
    free(sa5);
    free(text);
    /* the built handle already holds this index (with its SA and lc table): keep it for the lookups that follow */
    slot *s = slot_new(idx->bwt);
    s->h = h;
    s->built = 1;
}

void sa_build(const char *fname, long ram_use) {        /* psascan/sa_use.cc:8-18: only <fname>.sa5 */
    dna_fmi tmp;
    fmi_build(fname, &tmp, 32, ram_use);
    fmi_destroy(&tmp);
}

uint64_t fmi_aln(const dna_fmi *idx, const char *qry, int len, uint64_t *k, uint64_t *l) {       /* fmidx.c:295-313 */
    uint64_t r = 0;
    if (len <= 0) return *k > *l ? 0 : *l - *k + 1;
    int rc = as_fmi_aln_from_batch(handle_for(idx, NULL, 0), qry, len, 1, k, l, &r);
    if (rc != AS_OK) die("as_fmi_aln_from_batch", rc);
    return r;
}

/* fmidx.c:18-33: SA[loc] out of the global sa_buf, which the first call fills from <prefix>.sa5 (cache_sz entries).
 * The gather itself is served by the device copy of the array. */
uint64_t sa_access(const char *prefix, uint64_t cache_sz, uint64_t loc) {
    if (sa_buf) {
        if (loc > sa_buf->len) {
            printf("Accessing sa element out of scope");
            return 0;
        }
    } else {
        sa_buf = calloc(1, sizeof *sa_buf);
        sa_buf->mem = malloc(sizeof(ui40_t) * (cache_sz ? cache_sz : 1));
        FILE *fp = open2(prefix, ".sa5", "rb");
        uint8_t e[5];
        uint64_t i = 0;
        for (; i < cache_sz && fread(e, 1, 5, fp) == 5; i++) {
            memset(&sa_buf->mem[i], 0, sizeof(ui40_t));
            memcpy(&sa_buf->mem[i].low, e, 4);
            sa_buf->mem[i].high = e[4];
        }
        sa_buf->len = i;
        fclose(fp);
    }
    /* the gather is served from device memory: by a handle that already carries this array, else by the index-less
     * handle, to which the array is attached once */
    uint64_t pos = 0;
    for (int i = 0; i < MAX_SLOTS; i++)
        if (g_slot[i].h && g_slot[i].sa == (const void *)sa_buf->mem) {
            int rc = as_sa_access_batch(g_slot[i].h, &loc, 1, &pos);
            if (rc != AS_OK) die("as_sa_access_batch", rc);
            return pos;
        }
    if (loc >= sa_buf->len) return 0;                   /* loc == len passes the reference's check and reads past its array */
    if (!g_align) {
        int rc = as_gpu_open_aligner(&g_align, g_device);
        if (rc != AS_OK) die("as_gpu_open_aligner", rc);
    }
    if (g_align_sa != (const void *)sa_buf->mem) {
        int rc = as_gpu_attach_sa(g_align, sa_buf->mem, sa_buf->len, (int)sizeof(ui40_t));
        if (rc != AS_OK) die("as_gpu_attach_sa", rc);
        g_align_sa = sa_buf->mem;
    }
    int rc = as_sa_access_batch(g_align, &loc, 1, &pos);
    if (rc != AS_OK) die("as_sa_access_batch", rc);
    return pos;
}

void sa_done_access(void) {                             /* fmidx.c:35-38 */
    if (!sa_buf) return;
    g_align_sa = NULL;
    free(sa_buf->mem);
    free(sa_buf);
    sa_buf = NULL;
}

/* fmidx.c:315-331, the statement as the reference wrote it (`- 1`, give up after 5*ratio steps): the LF walk runs on
 * the device over the FM sectors, the sampled array idx->csa is attached to the handle once */
uint64_t csa_access(dna_fmi *fmi, uint64_t loc) {
    as_gpu *h = handle_for(fmi, NULL, 0);
    slot *s = slot_find(fmi->bwt);
    if (s->csa != fmi->csa) {
        int rc = as_gpu_attach_csa(h, fmi->csa, fmi->csa_len, fmi->csa_ratio);
        if (rc != AS_OK) die("as_gpu_attach_csa", rc);
        s->csa = fmi->csa;
    }
    uint64_t pos = 0;
    int rc = as_csa_access_batch(h, fmi->csa_ratio, 1, &loc, 1, &pos);
    if (rc != AS_OK) die("as_csa_access_batch", rc);
    return pos;
}

void fmi_write(dna_fmi idx, const char *prefix) {       /* .mfi layout, Notes.txt:6-25 / fmidx.c:221-244 */
    FILE *fp = open2(prefix, ".mfi", "wb");
    fwrite(idx.c, 8, 256, fp);
    fwrite(&idx.o_ratio, sizeof(int), 1, fp);
    fwrite(&idx.o_len, 8, 1, fp);
    fwrite(idx.o, 8, idx.o_len, fp);
    fwrite(&idx.length, 8, 1, fp);
    fwrite(idx.bwt, 1, idx.length, fp);
    fwrite(&idx.csa_ratio, sizeof(int), 1, fp);
    fwrite(&idx.csa_len, 8, 1, fp);
    fwrite(idx.csa, 8, idx.csa_len, fp);
    fclose(fp);
}

void fmi_read(dna_fmi *idx, const char *prefix) {       /* fmidx.c:246-275 */
    FILE *fp = open2(prefix, ".mfi", "rb");
    memset(idx, 0, sizeof *idx);
    idx->c = malloc(256 * 8);
    int ok = fread(idx->c, 8, 256, fp) == 256;
    ok &= fread(&idx->o_ratio, sizeof(int), 1, fp) == 1;
    ok &= fread(&idx->o_len, 8, 1, fp) == 1;
    idx->o = malloc(idx->o_len * 8);
    ok &= fread(idx->o, 8, idx->o_len, fp) == idx->o_len;
    ok &= fread(&idx->length, 8, 1, fp) == 1;
    idx->bwt = malloc(idx->length + 1);
    ok &= fread(idx->bwt, 1, idx->length, fp) == idx->length;
    idx->bwt[idx->length] = 0;
    ok &= fread(&idx->csa_ratio, sizeof(int), 1, fp) == 1;
    ok &= fread(&idx->csa_len, 8, 1, fp) == 1;
    idx->csa = malloc(idx->csa_len * 8);
    ok &= fread(idx->csa, 8, idx->csa_len, fp) == idx->csa_len;
    fclose(fp);
    if (!ok) {
        fprintf(stderr, "libaccseq_compat: %s.mfi is truncated\n", prefix);
        abort();
    }
}

void fmi_destroy(dna_fmi *idx) {                        /* fmidx.c:201-219: c, o, bwt (csa is left to the caller) */
    slot *s = idx->bwt ? slot_find(idx->bwt) : NULL;
    if (s) {
        as_gpu_close(s->h);
        memset(s, 0, sizeof *s);
    }
    free(idx->c);
    free(idx->o);
    free(idx->bwt);
    idx->c = idx->o = NULL;
    idx->bwt = NULL;
    idx->length = 0;
}

/* ------------------------------------------------------------------------------------------------
 * lchash.h
 * ---------------------------------------------------------------------------------------------- */
uint64_t lc_access(lc_hash h, uint64_t loc, uint64_t *k, uint64_t *l) {       /* lchash.c:12-16 */
    *k = h.lc[2 * loc];
    *l = h.lc[2 * loc + 1];
    return *l - *k;
}

uint64_t lc_length(int hlen) { return 1U << (2 * hlen); }                      /* lchash.c:75-77 */

lc_hash lc_build(const dna_fmi *fmi, int hlen) {                               /* lchash.c:52-73, 4^hlen searches at once */
    lc_hash hs;
    hs.hlen = hlen;
    hs.len = 2 * lc_length(hlen);
    hs.lc = malloc(hs.len * sizeof(uint64_t));
    if (!hs.lc) abort();
    int rc = as_lc_build(handle_for(fmi, NULL, 0), hlen, hs.lc);
    if (rc != AS_OK) die("as_lc_build", rc);
    return hs;
}

void lc_destroy(lc_hash hash) { free(hash.lc); }                               /* lchash.c:79-87 */

uint64_t lc_aln(const char *qry, int qlen, uint64_t *k, uint64_t *l, const dna_fmi *fmi, const lc_hash *hash) {   /* lchash.c:89-104 */
    uint64_t r = 0;
    int rc = as_lc_aln_batch(handle_for(fmi, hash, 0), qry, qlen, 1, k, l, &r);
    if (rc != AS_OK) die("as_lc_aln_batch", rc);
    return r;
}

void lc_write(const char *path, const lc_hash *h) {                            /* lchash.c:106-112 */
    FILE *fp = open2(path, "", "wb");
    fwrite(&h->hlen, sizeof(int), 1, fp);
    fwrite(&h->len, 8, 1, fp);
    fwrite(h->lc, 8, h->len, fp);
    fclose(fp);
}

void lc_read(const char *path, lc_hash *h) {                                   /* lchash.c:114-127 */
    FILE *fp = open2(path, "", "rb");
    int ok = fread(&h->hlen, sizeof(int), 1, fp) == 1;
    ok &= fread(&h->len, 8, 1, fp) == 1;
    h->lc = malloc(h->len * 8);
    ok &= h->lc && fread(h->lc, 8, h->len, fp) == h->len;
    fclose(fp);
    if (!ok) {
        fprintf(stderr, "libaccseq_compat: %s is truncated\n", path);
        abort();
    }
}

/* ------------------------------------------------------------------------------------------------
 * histo.h: the caller-owned vote list (histo.c:9-96).  Kept for callers of the scalar API; as_map_batch votes in its
 * own shared-memory table and never touches this.
 * ---------------------------------------------------------------------------------------------- */
histo *histo_init(u32 cap) {
    histo *h = malloc(sizeof *h);
    if (!h) abort();
    h->cap = cap;
    h->size = h->maxi = h->smaxi = 0;
    h->entries = malloc(sizeof(entry) * (cap ? cap : 1));
    return h;
}

void histo_destroy(histo *h) {
    free(h->entries);
    free(h);
}

void histo_add(histo *h, u64 key) {
    const u64 bucket = key >> 4;                        /* histo_key_hash */
    int hit = 0;
    for (u32 i = 0; i < h->size; i++) {
        entry *e = &h->entries[i];
        if (e->bucket != bucket) continue;
        hit = 1;
        e->val++;
        if (key < e->key) e->key = key;
    }
    if (hit) return;
    if (h->size + 1 == h->cap) {                        /* the list doubles one slot early, as the reference's does */
        h->entries = realloc(h->entries, sizeof(entry) * 2 * h->cap);
        if (!h->entries) abort();
        h->cap *= 2;
    }
    entry *e = &h->entries[h->size++];
    e->key = key;
    e->val = 1;
    e->bucket = bucket;
}

u64 histo_get(histo *h, u64 key) {                      /* votes of the key's bucket (the reference's own body reads past its
                                                         * list here, histo.c:58-72: what it means to return is this) */
    for (u32 i = 0; i < h->size; i++)
        if (h->entries[i].bucket == key >> 4) return h->entries[i].val;
    return 0;
}

u64 histo_find_max(histo *h) {
    u64 best = 0, val = 0;
    for (u32 i = 0; i < h->size; i++)
        if (h->entries[i].val > val) {
            val = h->entries[i].val;
            best = i;
        }
    return h->size ? h->entries[best].key : 0;
}

u64 histo_find_2_max(histo *h, entry *store) {          /* strict compares: the earlier entry wins a tie */
    memset(store, 0, 2 * sizeof(entry));
    for (u32 i = 0; i < h->size; i++) {
        const entry e = h->entries[i];
        if (e.val > store[0].val) {
            store[1] = store[0];
            store[0] = e;
        } else if (e.val > store[1].val) {
            store[1] = e;
        }
    }
    return store[0].val + store[1].val;
}

/* ------------------------------------------------------------------------------------------------
 * gact.h / mutils.h: the aligner names
 * ---------------------------------------------------------------------------------------------- */
mmstring ms_borrow(char *s, size_t l) {
    mmstring m;
    m.l = l;
    m.s = s;
    return m;
}

static as_gpu *any_handle(void) {
    for (int i = 0; i < MAX_SLOTS; i++)
        if (g_slot[i].h) return g_slot[i].h;
    if (!g_align) {
        int rc = as_gpu_open_aligner(&g_align, g_device);
        if (rc != AS_OK) die("as_gpu_open_aligner", rc);
    }
    return g_align;
}

void simple_gact(mmstring query, mmstring target, cigar *result) {              /* mutils.c:97-102 */
    as_params p;
    as_params_default(&p);
    uint32_t ql = (uint32_t)query.l, tl = (uint32_t)target.l;
    int32_t n_ops = 0, score = -1;
    int rc = as_cigar_align_batch(any_handle(), &p, query.s, ql ? ql : 1, &ql, target.s, tl ? tl : 1, &tl, 1, result->cigar,
                                  (size_t)ql + tl, &n_ops, &score);
    if (rc != AS_OK) die("as_cigar_align_batch", rc);
    result->n_cigar_op = n_ops;
    result->score = score;
}

cigar cigar_align(const char *qry, int qlen, const char *target, int tlen, int *limit, uint8_t *cigar_result) {   /* mutils.c:94-105 */
    cigar c;
    c.cigar = cigar_result;
    c.n_cigar_op = 0;
    c.score = -1;
    simple_gact(ms_borrow((char *)qry, (size_t)qlen), ms_borrow((char *)target, (size_t)tlen), &c);
    *limit = c.score;                                   /* the incoming limit is ignored and overwritten, mutils.c:103 */
    return c;
}

/* alnmain.c:497-498: run-length text of the column ops into a buffer of 2*len bytes, STANDARD alphabet M/I/D
 * (edlib.cpp:262-311); `*` when there is nothing to print or it does not fit */
void parse_cigar(cigar *c, size_t len, char *buf) {
    if (!c || !buf || len == 0) return;
    const size_t cap = 2 * len;
    int w = c->n_cigar_op > 0 && c->cigar ? as_cigar_text(c->cigar, c->n_cigar_op, buf, cap) : -1;
    if (w < 0) {
        buf[0] = '*';
        if (cap > 1) buf[1] = 0;
    }
}
