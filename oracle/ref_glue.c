/*
 * oracle/ref_glue.c -- TEST INFRASTRUCTURE ONLY.
 * Link-time glue that lets the reference's own sources (compiled where they lie under /root/reference by
 * oracle/Makefile) produce runnable `accidx` / `accaln` binaries and a shared object:
 *   - sa_build() is NOT here: oracle/Makefile compiles the reference's own pSAscan + libdivsufsort
 *     (psascan/sa_use.cc:8-18), so accidx_ref's <text>.sa5 comes from the reference's sorter.
 *   - simple_gact / parse_cigar / ms_borrow: the missing lisanhu/GACT submodule, provided by the oracle's
 *     documented stand-in (PARITY UNPINNED for CIGAR/ED).
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "gact/gact.h"
#include "oracle.h"

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
