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
