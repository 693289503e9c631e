/*
 * index_io.c -- reader of the reference's on-disk index (host tooling, plain C).
 * Formats: Notes.txt:6-25; writers asindex.c:89-110, fmidx.c:221-244, lchash.c:106-112, psascan .sa5
 * (sa_use.h:17-29); the reference reads them in init() (alnmain.c:179-256): fmi_read fmidx.c:246-275,
 * lc_read lchash.c:114-127, load_mta alnmain.c:125-140, load_file mutils.c:84, ui40_fread sa_use.h:31-46.
 * Differences: every fopen/fread is checked, and the .sa5 entries stay 5 bytes (the reference expands them
 * to 8-byte structs: 49.6 GB instead of 31 GB of host memory at human scale); as_gpu_open takes them as is.
 */
#include "accseq_host.h"

#include <errno.h>
#include <stdlib.h>
#include <string.h>

static char *path2(const char *a, const char *b) {
    size_t la = strlen(a), lb = strlen(b);
    char *s = malloc(la + lb + 1);
    if (!s) return NULL;
    memcpy(s, a, la);
    memcpy(s + la, b, lb + 1);
    return s;
}

static FILE *open2(const char *a, const char *b) {
    char *p = path2(a, b);
    if (!p) return NULL;
    FILE *fp = fopen(p, "rb");
    free(p);
    return fp;
}

static int rd(FILE *fp, void *dst, size_t bytes) { return fread(dst, 1, bytes, fp) == bytes ? 0 : -1; }

void as_index_free(as_host_index *ix) {
    if (!ix) return;
    free(ix->C); free(ix->O); free(ix->lc); free(ix->csa); free(ix->bwt); free(ix->content); free(ix->sa5);
    free(ix->mta);
    if (ix->names) {
        for (int i = 0; i < ix->view.mta_len; i++) free(ix->names[i]);
        free(ix->names);
    }
    free(ix);
}

int as_index_load(const char *fa, as_host_index **out) {
    if (!fa || !out) return AS_EINVAL;
    *out = NULL;
    as_host_index *ix = calloc(1, sizeof *ix);
    if (!ix) return AS_ENOMEM;
    FILE *fp = NULL;
    int rc = AS_EIO;
    uint64_t n = 0;

    /* <fa>.mta: {u64 name_len; name; u64 offset; u64 seq_len}*  (alnmain.c:125-140) */
    fp = open2(fa, ".mta");
    if (!fp) goto fail;
    {
        int cap = 16, cnt = 0;
        ix->mta = malloc(sizeof(as_mta) * (size_t)cap);
        ix->names = calloc((size_t)cap, sizeof(char *));
        if (!ix->mta || !ix->names) { rc = AS_ENOMEM; goto fail; }
        for (;;) {
            uint64_t nl;
            if (rd(fp, &nl, 8)) break;                       /* clean EOF ends the list */
            if (nl > (1u << 20)) goto fail;
            if (cnt == cap) {
                cap *= 2;
                as_mta *m2 = realloc(ix->mta, sizeof(as_mta) * (size_t)cap);
                char **n2 = realloc(ix->names, sizeof(char *) * (size_t)cap);
                if (m2) ix->mta = m2;
                if (n2) ix->names = n2;
                if (!m2 || !n2) { ix->view.mta_len = cnt; rc = AS_ENOMEM; goto fail; }
                memset(ix->names + cnt, 0, sizeof(char *) * (size_t)(cap - cnt));
            }
            char *nm = malloc(nl + 1);
            if (!nm) { ix->view.mta_len = cnt; rc = AS_ENOMEM; goto fail; }
            ix->names[cnt] = nm;
            ix->view.mta_len = cnt + 1;
            if (rd(fp, nm, nl) || rd(fp, &ix->mta[cnt].offset, 8) || rd(fp, &ix->mta[cnt].seq_len, 8)) goto fail;
            nm[nl] = 0;
            cnt++;
        }
        ix->view.mta = ix->mta;
    }
    fclose(fp);

    /* <fa>.cat.mfi  (fmidx.c:246-275) */
    fp = open2(fa, ".cat.mfi");
    if (!fp) goto fail;
    ix->C = malloc(256 * 8);
    if (!ix->C) { rc = AS_ENOMEM; goto fail; }
    if (rd(fp, ix->C, 256 * 8) || rd(fp, &ix->view.o_ratio, sizeof(int)) || rd(fp, &ix->view.o_len, 8)) goto fail;
    if (ix->view.o_len > (1ULL << 40)) goto fail;
    ix->O = malloc(ix->view.o_len * 8 + 8);
    if (!ix->O) { rc = AS_ENOMEM; goto fail; }
    if (rd(fp, ix->O, ix->view.o_len * 8) || rd(fp, &n, 8)) goto fail;
    if (n < 2 || n >= (1ULL << 40)) goto fail;
    ix->bwt = malloc(n + 1);
    if (!ix->bwt) { rc = AS_ENOMEM; goto fail; }
    if (rd(fp, ix->bwt, n)) goto fail;
    ix->bwt[n] = 0;
    /* the sampled SA (csa, ratio 4) is in the file but unused by the mapper (fmidx.c:315-331): skipped */
    if (rd(fp, &ix->csa_ratio, sizeof(int)) || rd(fp, &ix->csa_len, 8)) { ix->csa_ratio = 0; ix->csa_len = 0; }
    fclose(fp);
    ix->view.length = n;
    ix->view.C = ix->C;
    ix->view.O = ix->O;
    ix->view.bwt = ix->bwt;

    /* <fa>.cat.lch  (lchash.c:114-127) */
    fp = open2(fa, ".cat.lch");
    if (!fp) goto fail;
    if (rd(fp, &ix->view.hlen, sizeof(int)) || rd(fp, &ix->view.lc_len, 8)) goto fail;
    if (ix->view.hlen < 1 || ix->view.hlen > 14 || ix->view.lc_len != (2ULL << (2 * ix->view.hlen))) goto fail;
    ix->lc = malloc(ix->view.lc_len * 8);
    if (!ix->lc) { rc = AS_ENOMEM; goto fail; }
    if (rd(fp, ix->lc, ix->view.lc_len * 8)) goto fail;
    fclose(fp);
    ix->view.lc = ix->lc;

    /* <fa>.cat  (alnmain.c:237-240) */
    fp = open2(fa, ".cat");
    if (!fp) goto fail;
    ix->content = malloc(n + 1);
    if (!ix->content) { rc = AS_ENOMEM; goto fail; }
    if (rd(fp, ix->content, n)) goto fail;
    ix->content[n] = 0;
    fclose(fp);
    ix->view.content = ix->content;
    ix->view.con_len = n;

    /* <fa>.cat.sa5  (alnmain.c:243-252): n entries of 5 bytes, kept packed */
    fp = open2(fa, ".cat.sa5");
    if (!fp) goto fail;
    ix->sa5 = malloc(n * 5);
    if (!ix->sa5) { rc = AS_ENOMEM; goto fail; }
    if (rd(fp, ix->sa5, n * 5)) goto fail;
    fclose(fp);
    fp = NULL;
    ix->view.sa = ix->sa5;
    ix->view.sa_entry_bytes = 5;

    *out = ix;
    return AS_OK;
fail:
    if (fp) fclose(fp);
    as_index_free(ix);
    return rc;
}

/* ---------------------------------------------------------------------------------------------- */
/* construction side: the .cat image and the writers of the reference's files                      */
/* ---------------------------------------------------------------------------------------------- */
static char comp_upper(char c) {                       /* asindex.c:119-123 revc_mapper (0 for other bytes) */
    switch (c) {
    case 'A': return 'T';
    case 'C': return 'G';
    case 'G': return 'C';
    case 'T': return 'A';
    default: return 'N';
    }
}

/* create_meta (asindex.c:78-117) on in-memory sequences: upper-case, N -> a base, forward then reverse
 * complement per sequence, one final '$'.  The reference draws the N replacement from lrand48() seeded with
 * time() (asindex.c:125); here it is a fixed splitmix64 stream, so the image is reproducible. */
char *as_cat_build(int nseq, const char *const *seqs, const uint64_t *lens, as_mta *mta, uint64_t *con_len) {
    if (nseq < 0 || (nseq > 0 && (!seqs || !lens)) || !con_len) return NULL;
    uint64_t total = 0;
    for (int i = 0; i < nseq; i++) total += lens[i];
    char *cat = malloc(2 * total + 2);
    if (!cat) return NULL;
    uint64_t off = 0, rs = 0x5DEECE66DULL;
    for (int i = 0; i < nseq; i++) {
        if (mta) { mta[i].offset = off; mta[i].seq_len = lens[i]; }
        char *fwd = cat + off, *rev = fwd + lens[i];
        for (uint64_t j = 0; j < lens[i]; j++) {
            char c = seqs[i][j];
            if (c == 'n' || c == 'N') {
                uint64_t z = (rs += 0x9E3779B97F4A7C15ULL);
                z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
                z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
                c = "ACGT"[(z ^ (z >> 31)) & 3];
            }
            if (c > 0x60) c -= 0x20;                           /* asindex.c:63-67 */
            fwd[j] = c;
        }
        for (uint64_t j = 0; j < lens[i]; j++) rev[j] = comp_upper(fwd[lens[i] - 1 - j]);
        off += 2 * lens[i];
    }
    cat[off] = '$';
    cat[off + 1] = 0;
    *con_len = off + 1;
    return cat;
}

as_host_index *as_index_alloc(uint64_t n, int hlen, int o_ratio, int csa_ratio, int mta_len) {
    if (n < 2 || hlen < 1 || hlen > 14 || o_ratio <= 0 || mta_len < 0) return NULL;
    as_host_index *ix = calloc(1, sizeof *ix);
    if (!ix) return NULL;
    ix->view.length = n;
    ix->view.o_ratio = o_ratio;
    ix->view.o_len = 4 * (n / (uint64_t)o_ratio + 1);
    ix->view.hlen = hlen;
    ix->view.lc_len = 2ULL << (2 * hlen);
    ix->view.con_len = n;
    ix->view.sa_entry_bytes = 5;
    ix->view.mta_len = mta_len;
    ix->csa_ratio = csa_ratio;
    ix->csa_len = csa_ratio > 0 ? n / (uint64_t)csa_ratio + 1 : 0;
    ix->C = calloc(256, 8);
    ix->O = calloc(ix->view.o_len, 8);
    ix->lc = calloc(ix->view.lc_len, 8);
    ix->csa = ix->csa_len ? calloc(ix->csa_len, 8) : NULL;
    ix->bwt = calloc(n + 1, 1);
    ix->content = calloc(n + 1, 1);
    ix->sa5 = calloc(n, 5);
    ix->mta = calloc((size_t)(mta_len ? mta_len : 1), sizeof(as_mta));
    ix->names = calloc((size_t)(mta_len ? mta_len : 1), sizeof(char *));
    if (!ix->C || !ix->O || !ix->lc || (ix->csa_len && !ix->csa) || !ix->bwt || !ix->content || !ix->sa5 || !ix->mta ||
        !ix->names) {
        as_index_free(ix);
        return NULL;
    }
    ix->view.C = ix->C; ix->view.O = ix->O; ix->view.lc = ix->lc; ix->view.bwt = ix->bwt;
    ix->view.content = ix->content; ix->view.sa = ix->sa5; ix->view.mta = ix->mta;
    return ix;
}

void as_index_out_of(as_host_index *ix, as_index_out *o) {
    memset(o, 0, sizeof *o);
    o->C = ix->C; o->O = ix->O; o->o_ratio = ix->view.o_ratio; o->bwt = ix->bwt; o->lc = ix->lc;
    o->sa = ix->sa5; o->sa_entry_bytes = 5; o->csa = ix->csa; o->csa_ratio = ix->csa_ratio;
}

static FILE *create2(const char *a, const char *b) {
    char *p = path2(a, b);
    if (!p) return NULL;
    FILE *fp = fopen(p, "wb");
    free(p);
    return fp;
}

static int wr(FILE *fp, const void *src, size_t bytes) { return fwrite(src, 1, bytes, fp) == bytes ? 0 : -1; }

/* <fa>.mta (asindex.c:89-93), <fa>.cat (:97-110), <fa>.cat.sa5 (sa_use.cc:8-18), <fa>.cat.mfi (fmidx.c:221-244),
 * <fa>.cat.lch (lchash.c:106-112) */
int as_index_write(const char *fa, const as_host_index *ix) {
    if (!fa || !ix) return AS_EINVAL;
    const uint64_t n = ix->view.length;
    FILE *fp = create2(fa, ".mta");
    if (!fp) return AS_EIO;
    int bad = 0;
    for (int i = 0; i < ix->view.mta_len; i++) {
        uint64_t nl = strlen(ix->names[i]);
        bad |= wr(fp, &nl, 8) | wr(fp, ix->names[i], nl) | wr(fp, &ix->mta[i].offset, 8) | wr(fp, &ix->mta[i].seq_len, 8);
    }
    bad |= fclose(fp);
    if (bad) return AS_EIO;
    fp = create2(fa, ".cat");
    if (!fp) return AS_EIO;
    bad = wr(fp, ix->content, n) | fclose(fp);
    if (bad) return AS_EIO;
    fp = create2(fa, ".cat.sa5");
    if (!fp) return AS_EIO;
    bad = wr(fp, ix->sa5, n * 5) | fclose(fp);
    if (bad) return AS_EIO;
    fp = create2(fa, ".cat.mfi");
    if (!fp) return AS_EIO;
    bad = wr(fp, ix->C, 256 * 8) | wr(fp, &ix->view.o_ratio, sizeof(int)) | wr(fp, &ix->view.o_len, 8) |
          wr(fp, ix->O, ix->view.o_len * 8) | wr(fp, &n, 8) | wr(fp, ix->bwt, n) | wr(fp, &ix->csa_ratio, sizeof(int)) |
          wr(fp, &ix->csa_len, 8);
    if (ix->csa) {
        bad |= wr(fp, ix->csa, ix->csa_len * 8);
    } else {
        /* an index loaded by as_index_load does not keep the sampled SA: csa[i] = SA[csa_ratio*i] (fmidx.c:153-163) */
        for (uint64_t i = 0; i < ix->csa_len && !bad; i++) {
            const uint64_t row = i * (uint64_t)ix->csa_ratio;
            uint64_t v = 0;
            if (row < n) {
                const uint8_t *e = ix->sa5 + 5 * row;
                v = (uint64_t)e[0] | (uint64_t)e[1] << 8 | (uint64_t)e[2] << 16 | (uint64_t)e[3] << 24 | (uint64_t)e[4] << 32;
            }
            bad |= wr(fp, &v, 8);
        }
    }
    bad |= fclose(fp);
    if (bad) return AS_EIO;
    fp = create2(fa, ".cat.lch");
    if (!fp) return AS_EIO;
    bad = wr(fp, &ix->view.hlen, sizeof(int)) | wr(fp, &ix->view.lc_len, 8) | wr(fp, ix->lc, ix->view.lc_len * 8);
    bad |= fclose(fp);
    return bad ? AS_EIO : AS_OK;
}
