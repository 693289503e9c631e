/*
 * oracle/ref_glue.c -- TEST INFRASTRUCTURE ONLY.
 * Link-time glue that lets the reference's own sources (compiled where they lie under /root/reference by
 * oracle/Makefile) produce runnable `accidx` / `accaln` binaries and a shared object:
 *   - sa_build(): the reference calls pSAscan (psascan/sa_use.cc:8-18), which needs libdivsufsort's
 *     cmake-generated headers and is therefore not built here.  The suffix array of a text is unique, so
 *     the oracle's SA-IS writes the byte-identical <text>.sa5 (5-byte little-endian entries).
 *   - simple_gact / parse_cigar / ms_borrow: the missing lisanhu/GACT submodule, provided by the oracle's
 *     documented stand-in (PARITY UNPINNED for CIGAR/ED).
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "gact/gact.h"
#include "oracle.h"

void sa_build(const char *fname, long ram_use) {
    (void)ram_use;
    FILE *fp = fopen(fname, "rb");
    if (!fp) { fprintf(stderr, "sa_build: cannot open %s\n", fname); exit(2); }
    fseek(fp, 0, SEEK_END);
    long n = ftell(fp);
    fseek(fp, 0, SEEK_SET);
    uint8_t *text = malloc((size_t)n + 1);
    if (fread(text, 1, (size_t)n, fp) != (size_t)n) { fprintf(stderr, "sa_build: short read\n"); exit(2); }
    fclose(fp);
    uint64_t *sa = malloc(sizeof(uint64_t) * (size_t)n);
    if (orc_suffix_array(text, (uint64_t)n, sa) != 0) { fprintf(stderr, "sa_build: text must end in a unique smallest byte\n"); exit(2); }
    char *out = malloc(strlen(fname) + 5);
    sprintf(out, "%s.sa5", fname);
    fp = fopen(out, "wb");
    for (long i = 0; i < n; i++) {
        uint32_t low = (uint32_t)sa[i];
        uint8_t high = (uint8_t)(sa[i] >> 32);
        fwrite(&low, 4, 1, fp);
        fwrite(&high, 1, 1, fp);
    }
    fclose(fp);
    free(out); free(sa); free(text);
}

mmstring ms_borrow(char *s, size_t l) {
    mmstring m;
    m.s = s;
    m.l = l;
    return m;
}

void simple_gact(mmstring q, mmstring t, cigar *result) {
    int n = 0;
    result->score = orc_gact(q.s, (int)q.l, t.s, (int)t.l, ORC_GACT_T, ORC_GACT_O, result->cigar, &n, NULL);
    result->n_cigar_op = n;
}

void parse_cigar(cigar *c, size_t len, char *buf) {
    if (len == 0) return;
    if (orc_parse_cigar(c->cigar, c->n_cigar_op, buf, 2 * len) < 0) {
        /* does not fit the reference's 2*len buffer: truncate safely */
        buf[0] = '*';
        if (2 * len > 1) buf[1] = 0;
    }
}
