/*
 * oracle/oracle.c -- CPU restatement of the AccSeqV9 per-read mapping hot path.
 * TEST INFRASTRUCTURE ONLY (see oracle.h).  GACT parity: UNPINNED (submodule absent).
 * Citations are into /root/reference.
 */
#define _GNU_SOURCE
#include "oracle.h"

#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <stdbool.h>

/* ------------------------------------------------------------------------- */
/* Suffix array: SA-IS (Nong, Zhang, Chan 2009), 64-bit indices.              */
/* The reference obtains the same (unique) array from pSAscan                 */
/* (psascan/sa_use.cc:8-18); only the resulting .sa5 bytes matter.            */
/* ------------------------------------------------------------------------- */
typedef int64_t sidx;

#define TGET(i) ((t[(i) >> 3] >> ((i) & 7)) & 1)
#define TSET(i, b) (t[(i) >> 3] = (uint8_t)((b) ? (t[(i) >> 3] | (1u << ((i) & 7))) : (t[(i) >> 3] & ~(1u << ((i) & 7)))))
#define CHR(i) (cs == 1 ? (sidx)((const uint8_t *)s)[i] : ((const sidx *)s)[i])
#define IS_LMS(i) ((i) > 0 && TGET(i) && !TGET((i) - 1))

static void sais_buckets(const void *s, sidx *bkt, sidx n, sidx K, int cs, bool end) {
    sidx sum = 0;
    for (sidx i = 0; i <= K; i++) bkt[i] = 0;
    for (sidx i = 0; i < n; i++) bkt[CHR(i)]++;
    for (sidx i = 0; i <= K; i++) {
        sum += bkt[i];
        bkt[i] = end ? sum : sum - bkt[i];
    }
}

static void sais_induce_l(const uint8_t *t, sidx *SA, const void *s, sidx *bkt, sidx n, sidx K, int cs) {
    sais_buckets(s, bkt, n, K, cs, false);
    for (sidx i = 0; i < n; i++) {
        sidx j = SA[i] - 1;
        if (SA[i] > 0 && !TGET(j)) SA[bkt[CHR(j)]++] = j;
    }
}

static void sais_induce_s(const uint8_t *t, sidx *SA, const void *s, sidx *bkt, sidx n, sidx K, int cs) {
    sais_buckets(s, bkt, n, K, cs, true);
    for (sidx i = n - 1; i >= 0; i--) {
        sidx j = SA[i] - 1;
        if (SA[i] > 0 && TGET(j)) SA[--bkt[CHR(j)]] = j;
    }
}

static void sais_main(const void *s, sidx *SA, sidx n, sidx K, int cs) {
    uint8_t *t = calloc((size_t)n / 8 + 1, 1);
    sidx i, j;
    if (n == 1) { SA[0] = 0; free(t); return; }
    TSET(n - 2, 0);
    TSET(n - 1, 1);
    for (i = n - 3; i >= 0; i--)
        TSET(i, (CHR(i) < CHR(i + 1) || (CHR(i) == CHR(i + 1) && TGET(i + 1) == 1)) ? 1 : 0);

    sidx *bkt = malloc(sizeof(sidx) * (size_t)(K + 1));
    sais_buckets(s, bkt, n, K, cs, true);
    for (i = 0; i < n; i++) SA[i] = -1;
    for (i = 1; i < n; i++)
        if (IS_LMS(i)) SA[--bkt[CHR(i)]] = i;
    sais_induce_l(t, SA, s, bkt, n, K, cs);
    sais_induce_s(t, SA, s, bkt, n, K, cs);
    free(bkt);

    sidx n1 = 0;
    for (i = 0; i < n; i++)
        if (IS_LMS(SA[i])) SA[n1++] = SA[i];
    for (i = n1; i < n; i++) SA[i] = -1;
    sidx name = 0, prev = -1;
    for (i = 0; i < n1; i++) {
        sidx pos = SA[i];
        bool diff = false;
        for (sidx d = 0; d < n; d++) {
            if (prev == -1 || CHR(pos + d) != CHR(prev + d) || TGET(pos + d) != TGET(prev + d)) {
                diff = true;
                break;
            } else if (d > 0 && (IS_LMS(pos + d) || IS_LMS(prev + d)))
                break;
        }
        if (diff) { name++; prev = pos; }
        pos = pos / 2;
        SA[n1 + pos] = name - 1;
    }
    for (i = n - 1, j = n - 1; i >= n1; i--)
        if (SA[i] >= 0) SA[j--] = SA[i];

    sidx *SA1 = SA, *s1 = SA + n - n1;
    if (name < n1)
        sais_main(s1, SA1, n1, name - 1, (int)sizeof(sidx));
    else
        for (i = 0; i < n1; i++) SA1[s1[i]] = i;

    bkt = malloc(sizeof(sidx) * (size_t)(K + 1));
    sais_buckets(s, bkt, n, K, cs, true);
    for (i = 1, j = 0; i < n; i++)
        if (IS_LMS(i)) s1[j++] = i;
    for (i = 0; i < n1; i++) SA1[i] = s1[SA1[i]];
    for (i = n1; i < n; i++) SA[i] = -1;
    for (i = n1 - 1; i >= 0; i--) {
        j = SA[i];
        SA[i] = -1;
        SA[--bkt[CHR(j)]] = j;
    }
    sais_induce_l(t, SA, s, bkt, n, K, cs);
    sais_induce_s(t, SA, s, bkt, n, K, cs);
    free(bkt);
    free(t);
}

int orc_suffix_array(const uint8_t *text, uint64_t n, uint64_t *sa) {
    if (n == 0) return 0;
    for (uint64_t i = 0; i + 1 < n; i++)
        if (text[i] <= text[n - 1]) return -1;   /* last byte must be the unique smallest */
    sais_main(text, (sidx *)sa, (sidx)n, 255, 1);
    return 0;
}

/* ------------------------------------------------------------------------- */
/* Index construction and on-disk formats                                     */
/* ------------------------------------------------------------------------- */
static int base_code(char c) {   /* fmidx.c:278-282 / lchash.c:38-42 mapper[] */
    switch (c) {
    case 'A': case 'a': return 0;
    case 'C': case 'c': return 1;
    case 'G': case 'g': return 2;
    case 'T': case 't': return 3;
    default: return 0;           /* reference: uninitialised stack slot (UB); excluded from parity */
    }
}

static char comp_upper(char c) { /* asindex.c:119-123 revc_mapper */
    switch (c) {
    case 'A': return 'T';
    case 'C': return 'G';
    case 'G': return 'C';
    case 'T': return 'A';
    default: return 'N';
    }
}

static void build_tables(orc_index *ix, int o_ratio, int hlen) {
    uint64_t n = ix->n;
    /* C table: exclusive prefix counts by ASCII over text[0..n-2]  (fmidx.c:101-125) */
    memset(ix->C, 0, sizeof ix->C);
    for (uint64_t i = 0; i + 1 < n; i++) ix->C[(uint8_t)ix->cat[i]]++;
    uint64_t sum = 0;
    for (int i = 0; i < 256; i++) { uint64_t tmp = sum + ix->C[i]; ix->C[i] = sum; sum = tmp; }
    /* BWT from SA (fmidx.c:76-98) */
    ix->bwt = malloc(n + 1);
    for (uint64_t i = 0; i < n; i++) ix->bwt[i] = ix->sa[i] == 0 ? '$' : ix->cat[ix->sa[i] - 1];
    ix->bwt[n] = 0;
    /* O table (fmidx.c:128-150): block-major, counts of A,C,G,T in bwt[0 .. ratio*blk) */
    ix->o_ratio = o_ratio;
    ix->o_len = 4 * (n / (uint64_t)o_ratio + 1);
    ix->O = calloc(ix->o_len, sizeof(uint64_t));
    uint64_t tmp[4] = {0, 0, 0, 0};
    for (uint64_t i = 0; i < n; i++) {
        if (i % (uint64_t)o_ratio == 0) memcpy(ix->O + 4 * (i / (uint64_t)o_ratio), tmp, sizeof tmp);
        char c = ix->bwt[i];
        if (c != '\0' && c != '$') tmp[base_code(c)]++;
    }
    /* CSA (fmidx.c:153-163): SA[4i] */
    ix->csa_ratio = 4;
    ix->csa_len = n / 4 + 1;
    ix->csa = calloc(ix->csa_len, sizeof(uint64_t));
    for (uint64_t i = 0; i < ix->csa_len; i++)
        if (i * 4 < n) ix->csa[i] = ix->sa[i * 4];
    /* lchash (lchash.c:52-73): (k,l) inclusive SA interval of every hlen-mer, (0,0) if absent.
     * The reference runs 4^hlen backward searches; the interval of a pattern is the run of SA rows whose
     * suffix starts with it, so one pass over the SA yields the same table (tests check both ways). */
    ix->hlen = hlen;
    uint64_t upper = 1ULL << (2 * hlen);
    ix->lc_len = 2 * upper;
    ix->lc = calloc(ix->lc_len, sizeof(uint64_t));
    uint64_t run_code = ~0ULL;
    for (uint64_t r = 1; r < n; r++) {
        uint64_t p = ix->sa[r];
        /* skip suffixes that reach '$' within hlen bases, and also the occurrence that ENDS on the last real
         * base: the reference's backward search starts from k=1 (alnmain.c:354, lchash.c:57), so occ(c,0)
         * counts bwt[0] and the row of that occurrence (always the first of its run) is never included. */
        if (p + (uint64_t)hlen >= n - 1) { run_code = ~0ULL; continue; }
        uint64_t code = 0;
        for (int t = 0; t < hlen; t++) code = (code << 2) | (uint64_t)base_code(ix->cat[p + t]);
        if (code != run_code) { ix->lc[2 * code] = r; run_code = code; }
        ix->lc[2 * code + 1] = r;
    }
}

orc_index *orc_index_build(int nseq, const char *const *names, const char *const *seqs,
                           const uint64_t *lens, int o_ratio, int hlen) {
    orc_index *ix = calloc(1, sizeof *ix);
    uint64_t total = 0;
    for (int i = 0; i < nseq; i++) total += lens[i];
    ix->n = 2 * total + 1;
    ix->cat = malloc(ix->n + 1);
    ix->n_mta = nseq;
    ix->mta = calloc((size_t)nseq, sizeof(orc_mta));
    uint64_t off = 0;
    for (int i = 0; i < nseq; i++) {                       /* asindex.c:78-117 create_meta */
        ix->mta[i].name = strdup(names[i]);
        ix->mta[i].name_len = strlen(names[i]);
        ix->mta[i].offset = off;
        ix->mta[i].seq_len = lens[i];
        char *fwd = ix->cat + off;
        for (uint64_t j = 0; j < lens[i]; j++) {
            char c = seqs[i][j];
            if (c == 'n' || c == 'N') c = "ACGT"[j & 3];   /* reference: lrand48 seeded by time(); synthetic inputs have no N */
            if (c > 0x60) c -= 0x20;                       /* asindex.c:63-67 */
            fwd[j] = c;
        }
        char *rev = fwd + lens[i];
        for (uint64_t j = 0; j < lens[i]; j++) rev[j] = comp_upper(fwd[lens[i] - 1 - j]);
        off += 2 * lens[i];
    }
    ix->cat[ix->n - 1] = '$';
    ix->cat[ix->n] = 0;
    ix->sa = malloc(ix->n * sizeof(uint64_t));
    if (orc_suffix_array((const uint8_t *)ix->cat, ix->n, ix->sa) != 0) { orc_index_free(ix); return NULL; }
    build_tables(ix, o_ratio, hlen);
    return ix;
}

static char *path_cat(const char *a, const char *b) {
    char *s = malloc(strlen(a) + strlen(b) + 1);
    strcpy(s, a);
    strcat(s, b);
    return s;
}

int orc_index_write(const orc_index *ix, const char *fa) {
    char *p;
    FILE *fp;
    p = path_cat(fa, ".mta"); fp = fopen(p, "wb"); free(p);
    if (!fp) return -1;
    for (int i = 0; i < ix->n_mta; i++) {                  /* asindex.c:89-93, mutils.c:53-56 */
        fwrite(&ix->mta[i].name_len, 8, 1, fp);
        fwrite(ix->mta[i].name, 1, ix->mta[i].name_len, fp);
        fwrite(&ix->mta[i].offset, 8, 1, fp);
        fwrite(&ix->mta[i].seq_len, 8, 1, fp);
    }
    fclose(fp);
    p = path_cat(fa, ".cat"); fp = fopen(p, "wb"); free(p);
    if (!fp) return -1;
    fwrite(ix->cat, 1, ix->n, fp);
    fclose(fp);
    p = path_cat(fa, ".cat.sa5"); fp = fopen(p, "wb"); free(p);   /* sa_use.h:17-29: u32 low, u8 high */
    if (!fp) return -1;
    {
        size_t chunk = 1 << 20;
        uint8_t *buf = malloc(chunk * 5);
        for (uint64_t i = 0; i < ix->n; i += chunk) {
            size_t m = (size_t)((ix->n - i < chunk) ? ix->n - i : chunk);
            for (size_t j = 0; j < m; j++) {
                uint64_t v = ix->sa[i + j];
                uint32_t low = (uint32_t)v;
                memcpy(buf + 5 * j, &low, 4);
                buf[5 * j + 4] = (uint8_t)(v >> 32);
            }
            fwrite(buf, 5, m, fp);
        }
        free(buf);
    }
    fclose(fp);
    p = path_cat(fa, ".cat.mfi"); fp = fopen(p, "wb"); free(p);   /* fmidx.c:221-244 */
    if (!fp) return -1;
    fwrite(ix->C, 8, 256, fp);
    fwrite(&ix->o_ratio, sizeof(int), 1, fp);
    fwrite(&ix->o_len, 8, 1, fp);
    fwrite(ix->O, 8, ix->o_len, fp);
    fwrite(&ix->n, 8, 1, fp);
    fwrite(ix->bwt, 1, ix->n, fp);
    fwrite(&ix->csa_ratio, sizeof(int), 1, fp);
    fwrite(&ix->csa_len, 8, 1, fp);
    fwrite(ix->csa, 8, ix->csa_len, fp);
    fclose(fp);
    p = path_cat(fa, ".cat.lch"); fp = fopen(p, "wb"); free(p);   /* lchash.c:106-112 */
    if (!fp) return -1;
    fwrite(&ix->hlen, sizeof(int), 1, fp);
    fwrite(&ix->lc_len, 8, 1, fp);
    fwrite(ix->lc, 8, ix->lc_len, fp);
    fclose(fp);
    return 0;
}

#define RD(ptr, sz, cnt, fp) do { if (fread((ptr), (sz), (cnt), (fp)) != (size_t)(cnt)) goto fail; } while (0)

orc_index *orc_index_load(const char *fa) {
    orc_index *ix = calloc(1, sizeof *ix);
    char *p;
    FILE *fp = NULL;
    /* .mta  (alnmain.c:125-140) */
    p = path_cat(fa, ".mta"); fp = fopen(p, "rb"); free(p);
    if (!fp) goto fail;
    ix->mta = calloc(65535, sizeof(orc_mta));
    for (ix->n_mta = 0; ix->n_mta < 65535; ix->n_mta++) {
        orc_mta *m = &ix->mta[ix->n_mta];
        if (fread(&m->name_len, 8, 1, fp) != 1) break;
        m->name = malloc(m->name_len + 1);
        RD(m->name, 1, m->name_len, fp);
        m->name[m->name_len] = 0;
        RD(&m->offset, 8, 1, fp);
        RD(&m->seq_len, 8, 1, fp);
    }
    fclose(fp); fp = NULL;
    /* .cat.mfi  (fmidx.c:246-275) */
    p = path_cat(fa, ".cat.mfi"); fp = fopen(p, "rb"); free(p);
    if (!fp) goto fail;
    RD(ix->C, 8, 256, fp);
    RD(&ix->o_ratio, sizeof(int), 1, fp);
    RD(&ix->o_len, 8, 1, fp);
    ix->O = malloc(ix->o_len * 8);
    RD(ix->O, 8, ix->o_len, fp);
    RD(&ix->n, 8, 1, fp);
    ix->bwt = malloc(ix->n + 1);
    RD(ix->bwt, 1, ix->n, fp);
    ix->bwt[ix->n] = 0;
    RD(&ix->csa_ratio, sizeof(int), 1, fp);
    RD(&ix->csa_len, 8, 1, fp);
    ix->csa = malloc(ix->csa_len * 8);
    RD(ix->csa, 8, ix->csa_len, fp);
    fclose(fp); fp = NULL;
    /* .cat.lch  (lchash.c:114-127) */
    p = path_cat(fa, ".cat.lch"); fp = fopen(p, "rb"); free(p);
    if (!fp) goto fail;
    RD(&ix->hlen, sizeof(int), 1, fp);
    RD(&ix->lc_len, 8, 1, fp);
    ix->lc = malloc(ix->lc_len * 8);
    RD(ix->lc, 8, ix->lc_len, fp);
    fclose(fp); fp = NULL;
    /* .cat  (alnmain.c:237-240) */
    p = path_cat(fa, ".cat"); fp = fopen(p, "rb"); free(p);
    if (!fp) goto fail;
    ix->cat = malloc(ix->n + 1);
    RD(ix->cat, 1, ix->n, fp);
    ix->cat[ix->n] = 0;
    fclose(fp); fp = NULL;
    /* .cat.sa5  (alnmain.c:243-252, sa_use.h:31-46) */
    p = path_cat(fa, ".cat.sa5"); fp = fopen(p, "rb"); free(p);
    if (!fp) goto fail;
    ix->sa = malloc(ix->n * 8);
    {
        size_t chunk = 1 << 20;
        uint8_t *buf = malloc(chunk * 5);
        for (uint64_t i = 0; i < ix->n; i += chunk) {
            size_t m = (size_t)((ix->n - i < chunk) ? ix->n - i : chunk);
            if (fread(buf, 5, m, fp) != m) { free(buf); goto fail; }
            for (size_t j = 0; j < m; j++) {
                uint32_t low;
                memcpy(&low, buf + 5 * j, 4);
                ix->sa[i + j] = ((uint64_t)buf[5 * j + 4] << 32) | low;   /* ui40_convert */
            }
        }
        free(buf);
    }
    fclose(fp);
    return ix;
fail:
    if (fp) fclose(fp);
    orc_index_free(ix);
    return NULL;
}

void orc_index_free(orc_index *ix) {
    if (!ix) return;
    free(ix->cat); free(ix->O); free(ix->bwt); free(ix->csa); free(ix->lc); free(ix->sa);
    if (ix->mta) {
        for (int i = 0; i < ix->n_mta; i++) free(ix->mta[i].name);
        free(ix->mta);
    }
    free(ix);
}

/* ------------------------------------------------------------------------- */
/* Lookups                                                                     */
/* ------------------------------------------------------------------------- */
uint64_t orc_occ(const orc_index *ix, char c, uint64_t loc) {         /* fmidx.c:277-293 */
    uint64_t ratio = (uint64_t)ix->o_ratio;
    uint64_t id = loc / ratio, count = 0;
    for (uint64_t i = id * ratio; i <= loc; i++)
        if (ix->bwt[i] == c) count++;
    return ix->O[4 * id + (uint64_t)base_code(c)] + count;
}

static uint64_t fmi_aln_count(const orc_index *ix, const char *qry, int len, uint64_t *k, uint64_t *l,
                              uint64_t *steps) {                      /* fmidx.c:295-313 */
    uint64_t kk = *k, ll = *l;
    for (int i = len - 1; i >= 0; i--) {
        char c = qry[i];
        kk = ix->C[(uint8_t)c] + orc_occ(ix, c, kk - 1) + 1;
        ll = ix->C[(uint8_t)c] + orc_occ(ix, c, ll);
        if (steps) (*steps)++;
        if (kk > ll) break;
    }
    *k = kk;
    *l = ll;
    return kk > ll ? 0 : ll - kk + 1;
}

uint64_t orc_fmi_aln(const orc_index *ix, const char *qry, int len, uint64_t *k, uint64_t *l) {
    return fmi_aln_count(ix, qry, len, k, l, NULL);
}

static uint64_t lc_aln_count(const orc_index *ix, const char *qry, int qlen, uint64_t *k, uint64_t *l,
                             uint64_t *steps) {                       /* lchash.c:89-104 */
    int left = qlen - ix->hlen;
    if (qlen >= ix->hlen) {
        uint64_t num = 0;                                             /* lchash.c:36-49 */
        for (int i = 0; i < ix->hlen; i++) num = (num << 2) | (uint64_t)base_code(qry[left + i]);
        *k = ix->lc[2 * num];                                         /* lchash.c:12-16 */
        *l = ix->lc[2 * num + 1];
    } else {
        *k = 1;
        *l = ix->n - 1;
    }
    if (*k == 0 && *l == 0) return 0;
    return fmi_aln_count(ix, qry, left, k, l, steps);
}

uint64_t orc_lc_aln(const orc_index *ix, const char *qry, int qlen, uint64_t *k, uint64_t *l) {
    return lc_aln_count(ix, qry, qlen, k, l, NULL);
}

/* SA[row]: the expanded array, or -- for borrowed human-scale indexes (tests) -- the raw .sa5 bytes (sa_use.h:17-46) */
static inline uint64_t sa_at(const orc_index *ix, uint64_t row) {
    if (ix->sa) return ix->sa[row];
    const uint8_t *e = ix->sa5 + 5 * row;
    return (uint64_t)e[0] | (uint64_t)e[1] << 8 | (uint64_t)e[2] << 16 | (uint64_t)e[3] << 24 | (uint64_t)e[4] << 32;
}
uint64_t orc_sa_access(const orc_index *ix, uint64_t row) { return sa_at(ix, row); }  /* fmidx.c:18-33 */

/* csa_access, fmidx.c:315-331: locate through the sampled suffix array csa[i] = SA[ratio*i] (_build_csa,
 * fmidx.c:152-163) by walking LF until a sampled row.  literal != 0 follows the reference's statement as written,
 * `loc = c[ch] + occ(ch, loc) - 1` with its `counter > 5*ratio -> 0` cut-off; with this index's C table (text
 * characters before '$' only, fmidx.c:110) that lands one row below LF(loc), so the literal walk does not return
 * SA[loc] in general.  literal == 0 is the walk the reference's own test expects (`csa_access(i) == sa[i]`,
 * test/test-fmidx.cc:104-106): LF(loc) = c[ch] + occ(ch, loc), no cut-off. */
uint64_t orc_csa_access(const orc_index *ix, int ratio, uint64_t loc, int literal) {
    uint64_t r = (uint64_t)ratio, counter = 0;
    if (ratio <= 0 || loc >= ix->n) return 0;
    while (loc % r != 0) {
        char c = (char)ix->bwt[loc];
        if (c == '$') return counter;                                  /* fmidx.c:320-322: SA of that row is 0 */
        loc = ix->C[(uint8_t)c] + orc_occ(ix, c, loc) - (literal ? 1 : 0);
        counter++;
        if (literal && counter > 5 * r) return 0;                      /* fmidx.c:325-327 */
    }
    return sa_at(ix, loc) + counter;                                      /* csa[loc / ratio], fmidx.c:330 */
}

void orc_csa_access_batch(const orc_index *ix, int ratio, const uint64_t *rows, size_t n, int literal, uint64_t *pos) {
    for (size_t i = 0; i < n; i++) pos[i] = orc_csa_access(ix, ratio, rows[i], literal);
}

void orc_lc_aln_batch(const orc_index *ix, const char *seeds, int qlen, size_t n,
                      uint64_t *k, uint64_t *l, uint64_t *r) {
    for (size_t i = 0; i < n; i++) {
        k[i] = 1;
        l[i] = ix->n - 1;                                             /* alnmain.c:354 */
        r[i] = orc_lc_aln(ix, seeds + i * (size_t)qlen, qlen, &k[i], &l[i]);
    }
}

/* ------------------------------------------------------------------------- */
/* Voting                                                                      */
/* ------------------------------------------------------------------------- */
void orc_histo_init(orc_histo *h, uint32_t cap) {                    /* histo.c:9-15 */
    h->cap = cap ? cap : 1;
    h->size = 0;
    h->e = malloc(sizeof(orc_entry) * h->cap);
}

void orc_histo_free(orc_histo *h) { free(h->e); h->e = NULL; h->cap = h->size = 0; }

void orc_histo_add(orc_histo *h, uint64_t key) {                     /* histo.c:26-56 */
    uint64_t b = key >> 4;
    bool found = false;
    for (uint32_t i = 0; i < h->size; i++) {
        if (h->e[i].bucket == b) {
            found = true;
            h->e[i].val += 1;
            if (key < h->e[i].key) h->e[i].key = key;
        }
    }
    if (!found) {
        if (h->size + 1 >= h->cap) {
            h->cap *= 2;
            h->e = realloc(h->e, sizeof(orc_entry) * h->cap);
        }
        h->e[h->size].key = key;
        h->e[h->size].val = 1;
        h->e[h->size].bucket = b;
        h->size++;
    }
}

uint64_t orc_histo_top2(const orc_histo *h, orc_entry s[2]) {        /* histo.c:84-96 */
    memset(s, 0, 2 * sizeof(orc_entry));
    for (uint32_t i = 0; i < h->size; i++) {
        orc_entry e = h->e[i];
        if (s[1].val < e.val && s[0].val < e.val) { s[1] = s[0]; s[0] = e; }
        else if (s[1].val < e.val && s[0].val >= e.val) s[1] = e;
    }
    return s[0].val + s[1].val;
}

/* ------------------------------------------------------------------------- */
/* PART 1: seeding + voting + phase control (alnmain.c:333-405)               */
/* ------------------------------------------------------------------------- */
void orc_vote_read(const orc_index *ix, const char *seq, uint32_t len,
                   uint32_t seed_len, uint32_t thres, orc_vote *out) {
    const int sl = (int)seed_len, gl = 1, P = sl + gl;
    orc_histo outer, inner;
    orc_entry cand[2], best;
    memset(&best, 0, sizeof best);
    memset(out, 0, sizeof *out);
    orc_histo_init(&outer, thres);
    int iter;
    bool have_best = false;
    for (iter = 0; iter < P; ++iter) {
        orc_histo_init(&inner, thres);
        /* `j < r.len - sl` is an unsigned compare in the reference (alnmain.c:353); len <= sl would wrap
         * and read out of bounds (UB) -- treated as "no seeds" and excluded from parity. */
        if (len > (uint32_t)sl) {
            for (uint32_t j = (uint32_t)iter; j < len - (uint32_t)sl; j += (uint32_t)P) {
                uint64_t kk = 1, ll = ix->n - 1;
                uint64_t rr = lc_aln_count(ix, seq + j, sl, &kk, &ll, &out->n_steps);
                out->n_seeds++;
                if (rr > 0 && rr < thres) {
                    for (uint64_t k = kk; k <= ll; ++k) {
                        orc_histo_add(&inner, sa_at(ix, k) - j);          /* u64 wrap-around kept */
                        out->n_hits++;
                    }
                }
            }
        }
        int num_seeds = (int)(len / (uint32_t)P);
        if (num_seeds > 0) {
            uint64_t v = orc_histo_top2(&inner, cand);
            double score = (double)v / num_seeds;
            if (score > 0.6) {
                best = cand[0];
                have_best = true;
                orc_histo_free(&inner);
                break;
            } else if (cand[0].val != 0) {
                orc_histo_add(&outer, cand[0].key);
            }
        }
        orc_histo_free(&inner);
    }
    if (iter >= P - 1) {                                               /* alnmain.c:400-403 */
        orc_histo_top2(&outer, cand);
        best = cand[0];
        have_best = true;
        out->from_outer = 1;
    }
    (void)have_best;
    orc_histo_free(&outer);
    out->best_key = best.key;
    out->best_val = best.val;
    out->phases = iter;
}

/* ------------------------------------------------------------------------- */
/* PART 2: locus lookup, strand                                                */
/* ------------------------------------------------------------------------- */
int orc_seq_lookup(const orc_index *ix, uint64_t loc, uint32_t qlen,
                   int *seq_id, int *strand, uint64_t *off, uint64_t *fwd_loc) {   /* alnmain.c:151-176 */
    for (int i = 0; i < ix->n_mta; i++) {
        uint64_t sl = ix->mta[i].seq_len;
        uint64_t start = ix->mta[i].offset;
        uint64_t end = start + sl * 2;
        if (loc >= start && loc + qlen <= start + sl) {
            *strand = 0; *seq_id = i; *fwd_loc = loc; *off = loc - start;
            return 1;
        } else if (loc >= start + sl && loc + qlen <= end) {
            *strand = 1; *seq_id = i; *off = end - loc - qlen; *fwd_loc = *off + start;
            return 1;
        }
    }
    return 0;
}

void orc_revcomp(char *seq, uint32_t len) {                            /* alnmain.c:27-60 */
    for (uint32_t i = 0; i < len; i++) {
        switch (seq[i]) {
        case 'A': case 'a': seq[i] = 'T'; break;
        case 'C': case 'c': seq[i] = 'G'; break;
        case 'G': case 'g': seq[i] = 'C'; break;
        case 'T': case 't': seq[i] = 'A'; break;
        default: seq[i] = 'N'; break;
        }
    }
    for (uint32_t i = 0; i < len / 2; i++) {
        char c = seq[i];
        seq[i] = seq[len - 1 - i];
        seq[len - 1 - i] = c;
    }
}

/* ------------------------------------------------------------------------- */
/* GACT stand-in (PARITY UNPINNED; spec in DESIGN.md)                          */
/*                                                                             */
/* Call contract restated from mutils.c:94-105 / alnmain.c:440-446: global     */
/* alignment of q[0,qlen) against t[0,tlen), one op byte per column written    */
/* to caller storage, integer score (= number of non-match columns), -1 on     */
/* failure.  Tiling follows Darwin's GACT (README.md:5): T x T tiles, overlap  */
/* O, tile DP + traceback, extension restarted where the committed path ends.  */
/* ------------------------------------------------------------------------- */
int orc_gact(const char *q, int qlen, const char *t, int tlen, int T, int O,
             uint8_t *ops, int *n_ops, uint64_t *cells) {
    if (T <= 0 || O < 0 || O >= T || qlen < 0 || tlen < 0) { *n_ops = 0; return -1; }
    const int W = T + 1;
    int32_t *D = malloc(sizeof(int32_t) * (size_t)W * (size_t)W);
    uint8_t *ptr = malloc((size_t)W * (size_t)W);
    uint8_t *tmp = malloc((size_t)2 * (size_t)T + 2);
    int i0 = 0, j0 = 0, n = 0, score = 0;
    const int inner = T - O;
    while (i0 < qlen || j0 < tlen) {
        int Tq = qlen - i0 < T ? qlen - i0 : T;
        int Tt = tlen - j0 < T ? tlen - j0 : T;
        bool q_done = (i0 + Tq == qlen), t_done = (j0 + Tt == tlen);
        bool last = q_done && t_done;
        /* tile DP: unit costs, global start at the anchor corner */
        for (int j = 0; j <= Tt; j++) { D[j] = j; ptr[j] = 2; }
        for (int i = 1; i <= Tq; i++) {
            D[i * W] = i; ptr[i * W] = 1;
            char qc = q[i0 + i - 1];
            for (int j = 1; j <= Tt; j++) {
                int d = D[(i - 1) * W + (j - 1)] + (qc != t[j0 + j - 1]);
                int u = D[(i - 1) * W + j] + 1;
                int l = D[i * W + (j - 1)] + 1;
                int best = d, p = 0;              /* priority: diag, then up, then left */
                if (u < best) { best = u; p = 1; }
                if (l < best) { best = l; p = 2; }
                D[i * W + j] = best;
                ptr[i * W + j] = (uint8_t)p;
            }
        }
        if (cells) *cells += (uint64_t)Tq * (uint64_t)Tt;
        /* end cell */
        int ei = Tq, ej = Tt;
        if (!last) {
            /* candidates: tile edges interior to the full matrix; minimise (D, distance from corner, edge) */
            int64_t bestkey = INT64_MAX;
            if (!q_done) {                       /* bottom row i == Tq == T */
                for (int j = 0; j <= Tt; j++) {
                    int64_t key = ((int64_t)D[Tq * W + j] << 24) | ((int64_t)(Tt - j) << 1) | 0;
                    if (key < bestkey) { bestkey = key; ei = Tq; ej = j; }
                }
            }
            if (!t_done) {                       /* right column j == Tt == T */
                for (int i = 0; i <= Tq; i++) {
                    int64_t key = ((int64_t)D[i * W + Tt] << 24) | ((int64_t)(Tq - i) << 1) | 1;
                    if (key < bestkey) { bestkey = key; ei = i; ej = Tt; }
                }
            }
        }
        /* traceback from the end cell to the anchor; an op is committed iff its source point lies strictly
         * inside the (T-O) x (T-O) corner box (every op of the last tile is committed) */
        int ci = ei, cj = ej, m = 0, ai = 0, aj = 0;
        bool anchored = false;
        while (ci > 0 || cj > 0) {
            int p = ptr[ci * W + cj];
            int si = ci, sj = cj;
            uint8_t op;
            if (p == 0) { si--; sj--; op = (q[i0 + si] == t[j0 + sj]) ? 0 : 3; }
            else if (p == 1) { si--; op = 1; }
            else { sj--; op = 2; }
            if (last || (si < inner && sj < inner)) {
                if (!anchored) { ai = ci; aj = cj; anchored = true; }
                tmp[m++] = op;
            }
            ci = si; cj = sj;
        }
        if (!anchored) { ai = ei; aj = ej; }    /* only when the tile is empty; cannot happen for T-O >= 1 */
        if (m == 0 && !last) { score = -1; break; }
        for (int x = m - 1; x >= 0; x--) {
            ops[n++] = tmp[x];
            if (tmp[x] != 0) score++;
        }
        i0 += ai; j0 += aj;
        if (last) break;
    }
    free(D); free(ptr); free(tmp);
    *n_ops = n;
    return score;
}

/* edlib-equivalent global edit distance (edlib.cpp:694-895 myersCalcEditDistanceNW semantics:
 * unit costs, NW, returns -1 when k >= 0 and the distance exceeds k).  Bit-vector recurrence after
 * Myers 1999 / Hyyro 2001 as used by edlib.cpp:373-408 (calculateBlock); full matrix, no band, so the
 * value is exact and k only post-filters. */
int orc_edit_distance_nw(const char *q, int qlen, const char *t, int tlen, int k) {
    int d;
    if (qlen == 0) d = tlen;
    else if (tlen == 0) d = qlen;
    else {
        int nb = (qlen + 63) / 64;
        uint64_t *peq = calloc((size_t)nb * 256, sizeof(uint64_t));
        uint64_t *Pv = malloc(sizeof(uint64_t) * (size_t)nb), *Mv = calloc((size_t)nb, sizeof(uint64_t));
        for (int i = 0; i < qlen; i++) peq[(size_t)(uint8_t)q[i] * (size_t)nb + (size_t)(i >> 6)] |= 1ULL << (i & 63);
        for (int b = 0; b < nb; b++) Pv[b] = ~0ULL;
        int score = qlen;
        int lastbit = (qlen - 1) & 63;
        for (int j = 0; j < tlen; j++) {
            const uint64_t *eqc = peq + (size_t)(uint8_t)t[j] * (size_t)nb;
            int hin = 1;                               /* D[0][j] - D[0][j-1] = +1 */
            for (int b = 0; b < nb; b++) {
                uint64_t Eq = eqc[b], pv = Pv[b], mv = Mv[b];
                uint64_t hneg = hin < 0 ? 1ULL : 0ULL;
                uint64_t Xv = Eq | mv;
                Eq |= hneg;
                uint64_t Xh = (((Eq & pv) + pv) ^ pv) | Eq;
                uint64_t Ph = mv | ~(Xh | pv);
                uint64_t Mh = pv & Xh;
                int bit = (b == nb - 1) ? lastbit : 63;
                int hout = (int)((Ph >> bit) & 1) - (int)((Mh >> bit) & 1);
                Ph <<= 1; Mh <<= 1;
                Mh |= hneg;
                if (hin > 0) Ph |= 1ULL;
                Pv[b] = Mh | ~(Xv | Ph);
                Mv[b] = Ph & Xv;
                hin = hout;
            }
            score += hin;
        }
        d = score;
        free(peq); free(Pv); free(Mv);
    }
    if (k >= 0 && d > k) return -1;
    return d;
}

/* edlib's alignment path, EDLIB_MODE_NW + EDLIB_TASK_PATH (edlib/edlib.cpp:141-260 edlibAlign -> :1110-1150
 * obtainAlignment -> :909-1108 obtainAlignmentTraceback).  The traceback starts at the last cell and asks, in this
 * order: is the cell above one less (-> EDLIB_EDOP_INSERT 1, consumes a query base), is the cell to the left one
 * less (-> EDLIB_EDOP_DELETE 2), else it takes the diagonal (EDLIB_EDOP_MATCH 0 when the diagonal cell is equal,
 * EDLIB_EDOP_MISMATCH 3 otherwise); on the matrix borders it runs out along the border.  edlib reads those
 * differences from the stored Pv/Mv words of its block band; cells the band leaves out cost more than the best
 * score and never satisfy "one less", so the rule is the same on the full matrix, which is what this restatement
 * fills (plain O(qlen*tlen) DP: test infrastructure for small pairs).  Valid while edlib itself takes this
 * function, i.e. while its alignment data stays under 1 MiB (:1131-1133: 20 bytes * ceil(qlen/64) * tlen +
 * 8 * tlen); above that it switches to Hirschberg's divide and conquer (:1143), whose tie-breaks are not restated.
 * ops: forward order, one code per alignment column; returns the distance, or -1 (no alignment) when k >= 0 and
 * the distance exceeds k (edlib.cpp:705-708,:196).  An empty sequence is outside what edlibAlign handles. */
int orc_edit_alignment_nw(const char *q, int qlen, const char *t, int tlen, int k, uint8_t *ops, int *n_ops) {
    *n_ops = 0;
    const size_t W = (size_t)tlen + 1;
    int32_t *D = malloc(sizeof(int32_t) * ((size_t)qlen + 1) * W);
    if (!D) return -2;
    for (int j = 0; j <= tlen; j++) D[j] = j;
    for (int i = 1; i <= qlen; i++) {
        D[(size_t)i * W] = i;
        for (int j = 1; j <= tlen; j++) {
            int d = D[(size_t)(i - 1) * W + (j - 1)] + (q[i - 1] != t[j - 1]);
            int u = D[(size_t)(i - 1) * W + j] + 1, l = D[(size_t)i * W + (j - 1)] + 1;
            int m = d < u ? d : u;
            D[(size_t)i * W + j] = m < l ? m : l;
        }
    }
    const int dist = D[(size_t)qlen * W + tlen];
    if (k >= 0 && dist > k) { free(D); return -1; }
    int i = qlen, j = tlen, n = 0;
    while (i > 0 || j > 0) {
        uint8_t op;
        if (i == 0) op = 2;
        else if (j == 0) op = 1;
        else {
            const int cur = D[(size_t)i * W + j];
            if (D[(size_t)(i - 1) * W + j] + 1 == cur) op = 1;                 /* edlib.cpp:975 */
            else if (D[(size_t)i * W + (j - 1)] + 1 == cur) op = 2;            /* :1005 */
            else op = D[(size_t)(i - 1) * W + (j - 1)] == cur ? 0 : 3;         /* :1038-1039 */
        }
        ops[n++] = op;
        i -= (op != 2);
        j -= (op != 1);
    }
    for (int a = 0, b = n - 1; a < b; a++, b--) { uint8_t x = ops[a]; ops[a] = ops[b]; ops[b] = x; }   /* :1106 */
    *n_ops = n;
    free(D);
    return dist;
}

int orc_parse_cigar(const uint8_t *ops, int n_ops, char *buf, size_t buflen) {
    static const char code[4] = {'M', 'I', 'D', 'M'};   /* EDLIB_CIGAR_STANDARD, edlib.cpp:270-274 */
    size_t w = 0;
    int i = 0;
    if (buflen == 0) return -1;
    while (i < n_ops) {
        if (ops[i] > 3) return -1;
        char c = code[ops[i]];
        int run = 0;
        while (i < n_ops && ops[i] <= 3 && code[ops[i]] == c) { run++; i++; }
        int wrote = snprintf(buf + w, buflen - w, "%d%c", run, c);
        if (wrote < 0 || (size_t)wrote >= buflen - w) return -1;
        w += (size_t)wrote;
    }
    buf[w] = 0;
    return (int)w;
}

/* ------------------------------------------------------------------------- */
/* Whole path for a batch (alnmain.c:330-479)                                  */
/* ------------------------------------------------------------------------- */
void orc_map_batch(const orc_index *ix, char *reads, size_t stride, const uint32_t *lens, size_t n,
                   uint32_t seed_len, uint32_t thres, int gact_T, int gact_O, int do_myers,
                   orc_result *res, uint8_t *ops_store, size_t ops_stride,
                   uint64_t *total_cells) {
    for (size_t r = 0; r < n; r++) {
        char *seq = reads + r * stride;
        uint32_t len = lens[r];
        orc_vote v;
        orc_result *o = &res[r];
        memset(o, 0, sizeof *o);
        orc_vote_read(ix, seq, len, seed_len, thres, &v);              /* PART 1 */
        o->loc = v.best_key;
        o->phases = v.phases;
        o->from_outer = v.from_outer;
        int limit = (int)(0.05 * len * 2);                             /* alnmain.c:428 (ERROR_RATE) */
        int seq_id = -1, strand = 0;
        uint64_t off = 0, fwd = 0;
        int meta_r = orc_seq_lookup(ix, v.best_key, len, &seq_id, &strand, &off, &fwd);   /* PART 2 */
        o->myers = -2;
        if (!meta_r) {
            /* reference: m[] uninitialised, alignment runs on garbage (UB, alnmain.c:409,430-446).
             * Defined here as: unmapped, no alignment. */
            o->seq_id = -1; o->strand = 0; o->off = 0; o->ed = -1; o->n_ops = 0;
            o->flag = 4; o->mapq = 0; o->valid = 0;
            continue;
        }
        if (strand == 1) orc_revcomp(seq, len);                        /* alnmain.c:433-438 */
        int n_ops = 0;
        uint8_t *ops = ops_store + r * ops_stride;
        int score = orc_gact(seq, (int)len, ix->cat + fwd, (int)len, gact_T, gact_O, ops, &n_ops, total_cells);
        if (do_myers) o->myers = orc_edit_distance_nw(seq, (int)len, ix->cat + fwd, (int)len, limit);
        o->seq_id = seq_id; o->strand = strand; o->off = off; o->n_ops = n_ops;
        o->ed = score;                                                 /* mutils.c:103: *limit = score */
        o->flag = 0; o->mapq = 255; o->valid = score >= 0;             /* PART 3, alnmain.c:458-477 */
        if (score == -1) { o->valid = 0; o->flag += 4; o->mapq = 0; }
        else if (strand == 1) o->flag += 16;
    }
}
