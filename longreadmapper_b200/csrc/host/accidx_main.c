/*
 * accidx_main.c -- `accidx_b200`: the asindex.c-shaped index builder over the C ABI (plain C, no CUDA headers).
 *
 *   accidx_b200 ref.fa                                   (argv of the reference's accidx: asindex.c:129-153)
 *
 * Writes the reference's five files beside ref.fa -- <fa>.mta, <fa>.cat, <fa>.cat.sa5, <fa>.cat.mfi, <fa>.cat.lch
 * (Notes.txt:6-25) -- byte-identical to what accidx writes for genomes without N (the reference replaces N by
 * lrand48() seeded with time(), asindex.c:125).  create_meta (asindex.c:78-117) runs on the host; the suffix sort
 * (pSAscan, psascan/sa_use.cc:8-18), BWT / C / O / CSA (fmidx.c:76-199) and the 4^12 interval table (lchash.c:52-73)
 * are built on the GPU by as_gpu_build and come back through as_gpu_export in the reference's layouts.
 * ACCSEQ_DEVICE selects the GPU (default 0).
 */
#include "accseq_b200.h"
#include "accseq_host.h"

#include <stdio.h>
#include <stdlib.h>
#include <string.h>

int main(int argc, const char **argv) {
    if (argc != 2) {
        fprintf(stderr, "usage: %s ref.fa\n", argv[0]);
        return 1;
    }
    int nseq = 0;
    char **names = NULL, **seqs = NULL;
    uint64_t *lens = NULL;
    int rc = as_fasta_load(argv[1], &nseq, &names, &seqs, &lens);
    if (rc != AS_OK || nseq == 0) {
        fprintf(stderr, "accidx_b200: cannot read sequences from %s (%d)\n", argv[1], rc);
        return 2;
    }
    as_mta *mta = calloc((size_t)nseq, sizeof *mta);
    uint64_t con_len = 0;
    char *cat = as_cat_build(nseq, (const char *const *)seqs, lens, mta, &con_len);
    if (!mta || !cat) {
        fprintf(stderr, "accidx_b200: out of memory\n");
        return 6;
    }
    for (int i = 0; i < nseq; i++) free(seqs[i]);
    free(seqs);
    const int device = getenv("ACCSEQ_DEVICE") ? atoi(getenv("ACCSEQ_DEVICE")) : 0;
    const int hlen = 12, o_ratio = 32, csa_ratio = 4;          /* asindex.c:138-148, fmidx.c:194 */
    as_gpu *h = NULL;
    rc = as_gpu_build(&h, device, cat, con_len, mta, nseq, hlen);
    if (rc != AS_OK) {
        fprintf(stderr, "accidx_b200: as_gpu_build failed (%d): %s\n", rc, as_last_error());
        return 3;
    }
    as_host_index *ix = as_index_alloc(con_len, hlen, o_ratio, csa_ratio, nseq);
    if (!ix) {
        fprintf(stderr, "accidx_b200: out of memory for the exported index\n");
        return 6;
    }
    memcpy(ix->content, cat, con_len);
    free(cat);
    for (int i = 0; i < nseq; i++) {
        ix->mta[i] = mta[i];
        ix->names[i] = names[i];                                /* freed by as_index_free */
    }
    free(names);
    free(mta);
    free(lens);
    as_index_out out;
    as_index_out_of(ix, &out);
    rc = as_gpu_export(h, &out);
    if (rc != AS_OK) {
        fprintf(stderr, "accidx_b200: as_gpu_export failed (%d): %s\n", rc, as_last_error());
        return 3;
    }
    as_gpu_close(h);
    rc = as_index_write(argv[1], ix);
    if (rc != AS_OK) {
        fprintf(stderr, "accidx_b200: cannot write the index files beside %s (%d)\n", argv[1], rc);
        return 5;
    }
    as_index_free(ix);
    return 0;
}
