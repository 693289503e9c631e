/*
 * include/accseq_compat.h -- libaccseq_compat.so: the reference's SCALAR C surface on top of the batched C ABI.
 *
 * "the scalar signatures (lc_aln, fmi_aln, sa_access, cigar_align, histo_*) remain available as thin host
 * wrappers over the same index so existing callers/tests link unchanged".  Every name below has the signature and the
 * struct layout of the reference header cited beside it, so code compiled against the reference's OWN headers
 * (fmidx/fmidx.h, lchash/lchash.h, histo/histo.h, mutils.h, gact/gact.h) links against this library instead of
 * fmidx.c / lchash.c / histo.c / the missing gact submodule.  Include either the reference's headers or this one,
 * not both.
 *
 * What runs where:
 *   - lookups and alignment go to the GPU through libaccseq_b200.so, one element per call (a batch of 1): fmi_aln ->
 *     as_fmi_aln_from_batch, lc_aln -> as_lc_aln_batch, sa_access -> as_sa_access_batch, csa_access ->
 *     as_csa_access_batch(literal), cigar_align / simple_gact -> as_cigar_align_batch, fmi_build -> as_gpu_build +
 *     as_gpu_export, lc_build -> as_lc_build.  The device copy of an index is made the first time a dna_fmi is seen
 *     (keyed by its bwt pointer) and dropped by fmi_destroy.  No device -> the call prints as_last_error() and aborts:
 *     there is no CPU fallback behind these names.
 *   - containers and files are plain host code, as in the reference: lc_access / lc_length / lc_destroy / lc_write /
 *     lc_read (lchash.c:12-16,75-127), fmi_write / fmi_read / fmi_destroy (fmidx.c:201-275), sa_done_access, histo_*
 *     (histo.c:9-96: a list the caller owns; the GPU path votes in its own shared-memory table), ms_borrow, parse_cigar.
 */
#ifndef ACCSEQ_COMPAT_H
#define ACCSEQ_COMPAT_H

#include <stdint.h>
#include <stddef.h>
#include "gact/gact.h"

#ifdef __cplusplus
extern "C" {
#endif

/* ---- fmidx/fmidx.h:16-43, psascan/sa_use.h:17-20 ---- */
typedef struct {
    uint32_t low;
    uint8_t  high;
} ui40_t;

typedef struct _dna_fmi {
    uint64_t length, o_len, csa_len;
    uint64_t *c, *o, *csa;
    int o_ratio, csa_ratio;
    char *bwt;
} dna_fmi;

typedef struct {
    uint64_t start, len;
    ui40_t *mem;
} sa_mem;

extern sa_mem *sa_buf;

uint64_t sa_access(const char *prefix, uint64_t cache_sz, uint64_t loc);
uint64_t csa_access(dna_fmi *fmi, uint64_t loc);
void     sa_done_access(void);
void     fmi_build(const char *prefix, dna_fmi *idx, int o_ratio, long ram_use);
uint64_t fmi_aln(const dna_fmi *idx, const char *qry, int len, uint64_t *k, uint64_t *l);
void     fmi_write(dna_fmi idx, const char *prefix);
void     fmi_read(dna_fmi *idx, const char *prefix);
void     fmi_destroy(dna_fmi *idx);
void     sa_build(const char *fname, long ram_use);             /* psascan/sa_use.h:48 */

/* ---- lchash/lchash.h:16-30 ---- */
typedef struct {
    uint64_t *lc;
    uint64_t len;
    int hlen;
} lc_hash;

uint64_t lc_access(lc_hash h, uint64_t loc, uint64_t *k, uint64_t *l);
lc_hash  lc_build(const dna_fmi *fmi, int hlen);
uint64_t lc_length(int hlen);
uint64_t lc_aln(const char *qry, int qlen, uint64_t *k, uint64_t *l, const dna_fmi *fmi, const lc_hash *hash);
void     lc_destroy(lc_hash hash);
void     lc_write(const char *path, const lc_hash *h);
void     lc_read(const char *path, lc_hash *h);

/* ---- histo/histo.h:17-37 ---- */
typedef uint64_t u64;
typedef uint32_t u32;
typedef struct {
    u64 key, val, bucket;
} entry;
typedef struct {
    entry *entries;
    u32 cap, size, maxi, smaxi;
} histo;

histo *histo_init(u32 cap);
void   histo_add(histo *h, u64 key);
void   histo_destroy(histo *h);
u64    histo_get(histo *h, u64 key);
u64    histo_find_max(histo *h);
u64    histo_find_2_max(histo *h, entry *store);

/* ---- mutils.h:56-57 (the gact.h names -- ms_borrow, simple_gact, parse_cigar -- are declared by gact/gact.h) ---- */
cigar cigar_align(const char *qry, int qlen, const char *target, int tlen, int *limit, uint8_t *cigar_result);

/* ---- not in the reference: which GPU the wrappers use (default 0), and an explicit release of every device copy ---- */
void as_compat_set_device(int device);
void as_compat_release(void);

#ifdef __cplusplus
}
#endif
#endif
