/*
 * simgen.c -- seeded synthetic genomes and reads (host tooling, plain C).
 *
 * The reference's only generator is scripts/dna_txt_gen.py:12 -- one line of i.i.d. uniform `acgt`,
 * unseeded -- and it has no read simulator.  These functions draw from the same distribution with a
 * fixed-seed splitmix64 stream so every config of BASELINE.json is reproducible on any box, and add the
 * read sampler SURVEY.md 8(d) asks for (uniform position, both strands, sub/ins/del errors, truth kept).
 */
#include "accseq_host.h"

#include <stdlib.h>
#include <string.h>

static inline uint64_t splitmix64(uint64_t *s) {
    uint64_t z = (*s += 0x9E3779B97F4A7C15ULL);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
}

static inline double u01(uint64_t *s) { return (double)(splitmix64(s) >> 11) * (1.0 / 9007199254740992.0); }

void as_sim_genome(uint64_t seed, uint64_t n, char *out) {
    static const char alpha[4] = {'A', 'C', 'G', 'T'};
    uint64_t s = seed * 0xD1342543DE82EF95ULL + 1;
    uint64_t i = 0;
    while (i < n) {
        uint64_t r = splitmix64(&s);
        for (int b = 0; b < 32 && i < n; b++, i++) out[i] = alpha[(r >> (2 * b)) & 3];
    }
}

/* Plant repeats into a genome so that seeds hit many loci and vote ties occur (SURVEY Appendix A:
 * dispersed diverged copies, exact copies, tandem blocks, homopolymers). */
void as_sim_plant_repeats(uint64_t seed, char *g, uint64_t n, int n_families, uint64_t unit_len,
                          int copies, double divergence) {
    static const char alpha[4] = {'A', 'C', 'G', 'T'};
    uint64_t s = seed * 0x2545F4914F6CDD1DULL + 7;
    if (n < 4 * unit_len + 1024) return;
    for (int f = 0; f < n_families; f++) {
        uint64_t src = splitmix64(&s) % (n - unit_len);
        for (int c = 0; c < copies; c++) {
            uint64_t dst = splitmix64(&s) % (n - unit_len);
            if (dst + unit_len > src && src + unit_len > dst) continue;
            for (uint64_t i = 0; i < unit_len; i++) {
                char ch = g[src + i];
                if (u01(&s) < divergence) ch = alpha[splitmix64(&s) & 3];
                g[dst + i] = ch;
            }
        }
    }
    /* a tandem block and a homopolymer run */
    uint64_t p = splitmix64(&s) % (n - 1024);
    for (int i = 0; i < 480; i++) g[p + (uint64_t)i] = "ACGTTGCA"[i & 7];
    p = splitmix64(&s) % (n - 1024);
    for (int i = 0; i < 300; i++) g[p + (uint64_t)i] = 'A';
}

static char comp(char c) {
    switch (c) {
    case 'A': return 'T';
    case 'C': return 'G';
    case 'G': return 'C';
    default: return 'A';
    }
}

size_t as_sim_reads(uint64_t seed, const char *genome, uint64_t glen, size_t n_reads,
                    uint32_t len_min, uint32_t len_max, double err_min, double err_max,
                    double sub_frac, char *out, size_t stride, uint32_t *lens,
                    uint64_t *truth_pos, uint8_t *truth_strand) {
    static const char alpha[4] = {'A', 'C', 'G', 'T'};
    if (len_max < len_min || glen < (uint64_t)len_max * 2 + 64 || stride < (size_t)len_max + 1) return 0;
    char *tmp = malloc((size_t)len_max + 1);
    for (size_t r = 0; r < n_reads; r++) {
        uint64_t s = (seed + 0x632BE59BD9B4E019ULL) * (2 * (uint64_t)r + 1) + 0x9E3779B97F4A7C15ULL * (uint64_t)r;
        splitmix64(&s);
        uint32_t len = len_min + (uint32_t)(splitmix64(&s) % (uint64_t)(len_max - len_min + 1));
        double err = err_min + (err_max - err_min) * u01(&s);
        /* source window a little longer than the read so deletions cannot run off the end */
        uint64_t span = (uint64_t)len + (uint64_t)len / 4 + 16;
        uint64_t pos = splitmix64(&s) % (glen - span);
        int strand = (int)(splitmix64(&s) & 1);
        uint64_t gi = pos;
        uint32_t w = 0;
        while (w < len) {
            double x = u01(&s);
            if (x < err) {
                double y = u01(&s);
                if (y < sub_frac) {                       /* substitution: always a different base */
                    char c = genome[gi++];
                    char d;
                    do { d = alpha[splitmix64(&s) & 3]; } while (d == c);
                    tmp[w++] = d;
                } else if (y < sub_frac + (1.0 - sub_frac) / 2) {   /* insertion in the read */
                    tmp[w++] = alpha[splitmix64(&s) & 3];
                } else {                                  /* deletion from the read */
                    gi++;
                }
            } else {
                tmp[w++] = genome[gi++];
            }
            if (gi >= pos + span) break;
        }
        while (w < len) tmp[w++] = genome[gi < glen ? gi++ : glen - 1];
        char *dst = out + r * stride;
        if (strand) {
            for (uint32_t i = 0; i < len; i++) dst[i] = comp(tmp[len - 1 - i]);
        } else {
            memcpy(dst, tmp, len);
        }
        memset(dst + len, 0, stride - len);
        lens[r] = len;
        if (truth_pos) truth_pos[r] = pos;
        if (truth_strand) truth_strand[r] = (uint8_t)strand;
    }
    free(tmp);
    return n_reads;
}
