/*
 * include/gact/gact.h -- the names the reference expects from its (un-vendored) `gact/` submodule.
 *
 * The reference includes "gact/gact.h" from accaln.h:19, mutils.h:23 and mutils.c:12 and uses exactly
 * these names: `cigar` (fields .cigar, .n_cigar_op, .score: mutils.c:99-104, alnmain.c:323-324),
 * `mmstring` + `ms_borrow` (mutils.c:97-98), `simple_gact` (mutils.c:102) and `parse_cigar`
 * (alnmain.c:498).  The submodule (lisanhu/GACT @ 5ee0690b) is absent from the reference tree, so this
 * header *defines* the layout; with `-I include` the reference's own sources compile unchanged against it
 * (oracle/Makefile does exactly that).
 */
#ifndef ACCSEQ_B200_GACT_H
#define ACCSEQ_B200_GACT_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct mmstring {
    uint64_t l;
    char *s;
} mmstring;

/* One op byte per alignment column, edlib codes (edlib.h EDLIB_EDOP_*):
 * 0 '=' match, 1 'I' (consumes query), 2 'D' (consumes target), 3 'X' mismatch. */
typedef struct cigar {
    uint8_t *cigar;      /* caller-owned storage, >= qlen + tlen bytes (alnmain.c:316-320) */
    int n_cigar_op;      /* number of columns written */
    int score;           /* number of non-match columns (reported as ED:I:), -1 = failure */
} cigar;

mmstring ms_borrow(char *s, size_t l);
void simple_gact(mmstring query, mmstring target, cigar *result);
/* run-length encode to SAM STANDARD text (M/I/D); `len` is the read length, buf holds 2*len bytes
 * (alnmain.c:497-498). Output is truncated to fit and always NUL-terminated. */
void parse_cigar(cigar *c, size_t len, char *buf);

#ifdef __cplusplus
}
#endif
#endif
