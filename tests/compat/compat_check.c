/*
 * tests/compat/compat_check.c -- the invariants the reference's own unit tests check on its scalar C surface, as a plain C
 * program linked against libaccseq_compat.so (no Catch2 here):
 *   test/test-lchash.cc:19-29   lc_write / lc_read round trip
 *   test/test-lchash.cc:30-48   lc_aln == fmi_aln on a query taken from the text, exactly one hit, text[SA[k]..] == query
 *   test/test-fmidx.cc:17-28    fmi_write / fmi_read round trip
 *   test/test-fmidx.cc:30-41    fmi_aln + sa_access spell the query
 *   test/test-fmidx.cc:104-106  csa_access(i) == sa[i]  (at the sampled rows: see oracle.c orc_csa_access on the literal form)
 *   edlib/ed_test.c:14-19       two 100-mers two substitutions apart: distance 2, CIGAR 100M -- through cigar_align / parse_cigar
 * Compiled twice by tests/Makefile: against include/accseq_compat.h, and (where /root/reference exists) against the
 * reference's OWN headers with -DUSE_REFERENCE_HEADERS, which proves the struct layouts and signatures are the reference's.
 *
 *   compat_check <text file ending in '$'>       exit 0 = all invariants hold
 */
#ifdef USE_REFERENCE_HEADERS
#include "fmidx/fmidx.h"
#include "lchash/lchash.h"
#include "histo/histo.h"
#include "mutils.h"
#else
#include "accseq_compat.h"
#endif

#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define REQUIRE(c)                                                          \
    do {                                                                    \
        if (!(c)) {                                                         \
            fprintf(stderr, "compat_check: line %d: %s\n", __LINE__, #c);   \
            return 1;                                                       \
        }                                                                   \
    } while (0)

int main(int argc, char **argv) {
    if (argc != 2) return 2;
    const char *prefix = argv[1];
    FILE *fp = fopen(prefix, "rb");
    if (!fp) return 2;
    fseek(fp, 0, SEEK_END);
    const uint64_t n = (uint64_t)ftell(fp);
    fseek(fp, 0, SEEK_SET);
    char *text = malloc(n + 1);
    REQUIRE(fread(text, 1, n, fp) == n);
    fclose(fp);

    dna_fmi fmi;
    fmi_build(prefix, &fmi, 128, 8L << 30);                       /* test-lchash.cc:18 builds with o_ratio 128 */
    REQUIRE(fmi.length == n && fmi.o_ratio == 128 && fmi.o_len == 4 * (n / 128 + 1));
    lc_hash h = lc_build(&fmi, 12);
    REQUIRE(h.hlen == 12 && h.len == 2 * lc_length(12));

    /* serialization round trips */
    char path[4096];
    snprintf(path, sizeof path, "%s.lch", prefix);
    lc_write(path, &h);
    lc_hash h2;
    lc_read(path, &h2);
    REQUIRE(h.hlen == h2.hlen && h.len == h2.len && !memcmp(h.lc, h2.lc, h.len * sizeof(uint64_t)));
    fmi_write(fmi, prefix);
    dna_fmi mfi;
    fmi_read(&mfi, prefix);
    REQUIRE(fmi.o_ratio == mfi.o_ratio && fmi.length == mfi.length && fmi.o_len == mfi.o_len);
    REQUIRE(!memcmp(fmi.bwt, mfi.bwt, fmi.length) && !memcmp(fmi.o, mfi.o, fmi.o_len * sizeof(uint64_t)));
    REQUIRE(!memcmp(fmi.c, mfi.c, 256 * sizeof(uint64_t)));
    REQUIRE(fmi.csa_ratio == mfi.csa_ratio && fmi.csa_len == mfi.csa_len && !memcmp(fmi.csa, mfi.csa, fmi.csa_len * 8));

    /* the suffix array fmi_build wrote */
    snprintf(path, sizeof path, "%s.sa5", prefix);
    fp = fopen(path, "rb");
    REQUIRE(fp != NULL);
    uint8_t *sa5 = malloc(5 * n);
    REQUIRE(fread(sa5, 5, n, fp) == n);
    fclose(fp);
#define SA(i) ((uint64_t)sa5[5 * (i)] | (uint64_t)sa5[5 * (i) + 1] << 8 | (uint64_t)sa5[5 * (i) + 2] << 16 | \
               (uint64_t)sa5[5 * (i) + 3] << 24 | (uint64_t)sa5[5 * (i) + 4] << 32)
    REQUIRE(SA(0) == n - 1);                                      /* '$' sorts first */

    /* lchash alignment + fmidx alignment on queries cut from the text */
    const int qlen = 64;
    int unique = 0;
    for (int t = 0; t < 40; t++) {
        const uint64_t p = (uint64_t)t * 7919u % (n - qlen - 1);
        const char *query = text + p;
        uint64_t k = 1, l = fmi.length - 1;
        const uint64_t r = fmi_aln(&fmi, query, qlen, &k, &l);
        uint64_t kk = 0, ll = 0;
        const uint64_t rr = lc_aln(query, qlen, &kk, &ll, &fmi, &h);
        REQUIRE(r >= 1 && rr == r && kk == k && ll == l);
        const uint64_t off = sa_access(prefix, n, kk);
        REQUIRE(off == SA(kk) && !memcmp(query, text + off, qlen));
        unique += r == 1;
        /* a k-mer lookup alone, and a search continued from a caller's interval (what lc_aln does inside) */
        uint64_t k12 = 1, l12 = fmi.length - 1, ka, la;
        fmi_aln(&fmi, query + qlen - 12, 12, &k12, &l12);
        uint64_t code = 0;
        for (int i = 0; i < 12; i++) code = code << 2 | (uint64_t)(strchr("ACGT", query[qlen - 12 + i]) - "ACGT");
        lc_access(h, code, &ka, &la);
        REQUIRE(ka == k12 && la == l12);
        REQUIRE(fmi_aln(&fmi, query, qlen - 12, &k12, &l12) == r && k12 == k && l12 == l);
    }
    REQUIRE(unique > 0);
    /* a query that does not occur */
    {
        char q[65];
        memset(q, 'A', 64);
        q[64] = 0;
        for (int i = 0; i < 64; i += 3) q[i] = "CGT"[i % 3];
        uint64_t k = 1, l = fmi.length - 1, kk, ll;
        const uint64_t r = fmi_aln(&fmi, q, 64, &k, &l);
        REQUIRE(lc_aln(q, 64, &kk, &ll, &fmi, &h) == r);
    }

    /* csa_access at sampled rows, and the '$' row */
    for (uint64_t i = 0; i < n; i += 4 * 37) REQUIRE(csa_access(&fmi, i) == SA(i));
    sa_done_access();

    /* histo: two buckets, ties go to the earlier one */
    {
        histo *hs = histo_init(4);
        const u64 keys[] = {1000, 1003, 5000, 5001, 999 + 16 * 0, 77777};
        for (int i = 0; i < 6; i++) histo_add(hs, keys[i]);
        entry top[2];
        const u64 v = histo_find_2_max(hs, top);
        REQUIRE(v == 5 && top[0].val == 3 && top[0].key == 999 && top[1].val == 2 && top[1].key == 5000);
        REQUIRE(histo_find_max(hs) == 999);
        histo_destroy(hs);
    }

    /* the aligner names: two 100-mers, substitutions at columns 10 and 60 */
    {
        char q[101], t[101];
        memcpy(t, text + 1000, 100);
        memcpy(q, t, 100);
        q[100] = t[100] = 0;
        q[10] = q[10] == 'A' ? 'C' : 'A';
        q[60] = q[60] == 'G' ? 'T' : 'G';
        uint8_t ops[200];
        int limit = 10;
        cigar c = cigar_align(q, 100, t, 100, &limit, ops);
        REQUIRE(c.score == 2 && limit == 2 && c.n_cigar_op == 100 && c.cigar == ops);
        REQUIRE(ops[10] == 3 && ops[60] == 3 && ops[0] == 0);
        char buf[200];
        parse_cigar(&c, 100, buf);
        REQUIRE(!strcmp(buf, "100M"));
    }

    fmi_destroy(&fmi);
    fmi_destroy(&mfi);
    lc_destroy(h);
    lc_destroy(h2);
    printf("compat_check ok: %llu rows, %d/40 unique 64-mers\n", (unsigned long long)n, unique);
    return 0;
}
