/*
 * sam.c -- SAM text around the hot path (host tooling, plain C).
 *   header  gen_sam_header alnmain.c:62-75
 *   record  alnmain.c:488-526: QNAME FLAG RNAME off+1 MAPQ CIGAR * 0 0 SEQ QUAL ED:I:<ed>
 *   CIGAR   parse_cigar (un-vendored lisanhu/GACT; contract at alnmain.c:497-498): run-length text of the
 *           column ops, STANDARD alphabet M/I/D as edlibAlignmentToCigar writes it (edlib.cpp:262-311)
 */
#include "accseq_host.h"

#include <stdlib.h>
#include <string.h>

int as_sam_header(FILE *out, const as_host_index *ix, long rg_id) {
    if (!out || !ix) return AS_EINVAL;
    for (int i = 0; i < ix->view.mta_len; i++)
        fprintf(out, "@SQ\tSN:%s\tLN:%ld\n", ix->names[i], (long)ix->mta[i].seq_len);
    fprintf(out, "@RG\tID:%s%ld\tSM:SM_data\n", "accaln", rg_id);        /* reference: rg_id = time(NULL) */
    fprintf(out, "@PG\tID:%s\tPN:%s\n", "accaln", "accaln");
    return AS_OK;
}

int as_cigar_text(const uint8_t *ops, int n_ops, char *buf, size_t buflen) {
    static const char code[4] = {'M', 'I', 'D', 'M'};
    size_t w = 0;
    int i = 0;
    if (!buf || buflen == 0 || (n_ops > 0 && !ops)) return -1;
    while (i < n_ops) {
        if (ops[i] > 3) return -1;
        const char c = code[ops[i]];
        int run = 0;
        while (i < n_ops && ops[i] <= 3 && code[ops[i]] == c) { run++; i++; }
        char tmp[16];
        int d = 0;
        do { tmp[d++] = (char)('0' + run % 10); run /= 10; } while (run);
        if (w + (size_t)d + 2 > buflen) return -1;
        while (d) buf[w++] = tmp[--d];
        buf[w++] = c;
    }
    buf[w] = 0;
    return (int)w;
}

/* run bytes of AS_F_RLE_CIGAR ((len << 2) | class, class 0 M 1 I 2 D; a run longer than 63 continues in the next byte) */
int as_cigar_text_rle(const uint8_t *runs, int n_runs, char *buf, size_t buflen) {
    static const char code[4] = {'M', 'I', 'D', '?'};
    size_t w = 0;
    int i = 0;
    if (!buf || buflen == 0 || (n_runs > 0 && !runs)) return -1;
    while (i < n_runs) {
        const int cls = runs[i] & 3;
        long run = 0;
        if (cls == 3 || (runs[i] >> 2) == 0) return -1;
        while (i < n_runs && (runs[i] & 3) == cls) { run += runs[i] >> 2; i++; }
        char tmp[24];
        int d = 0;
        do { tmp[d++] = (char)('0' + run % 10); run /= 10; } while (run);
        if (w + (size_t)d + 2 > buflen) return -1;
        while (d) buf[w++] = tmp[--d];
        buf[w++] = code[cls];
    }
    buf[w] = 0;
    return (int)w;
}

int as_sam_record(FILE *out, const as_host_index *ix, const char *qname, size_t qname_len, const char *seq,
                  const char *qual, uint32_t len, const as_result *r, const uint8_t *ops) {
    return as_sam_record2(out, ix, qname, qname_len, seq, qual, len, r, ops, 0);
}

/* map_flags: the AS_F_* the batch was mapped with -- AS_F_RLE_CIGAR: `ops` holds run bytes; AS_F_KEEP_READS: `seq` is
 * still the read as sequenced, so strand-1 records print its reverse complement (alnmain.c:27-60,433-438) */
int as_sam_record2(FILE *out, const as_host_index *ix, const char *qname, size_t qname_len, const char *seq,
                   const char *qual, uint32_t len, const as_result *r, const uint8_t *ops, unsigned map_flags) {
    if (!out || !ix || !r) return AS_EINVAL;
    char stackbuf[4096], seqbuf[1024];
    char *cig = stackbuf, *rc = NULL;
    size_t need = 2 * (size_t)len + 16;                  /* the reference's buffer is 2*len (alnmain.c:497) */
    int heap = 0, rc_heap = 0;
    if (need > sizeof stackbuf) {
        cig = malloc(need);
        if (!cig) return AS_ENOMEM;
        heap = 1;
    }
    int bad = r->seq_id < 0 || !ops || r->n_cigar_op <= 0;
    if (!bad)
        bad = ((map_flags & AS_F_RLE_CIGAR) ? as_cigar_text_rle(ops, r->n_cigar_op, cig, need)
                                            : as_cigar_text(ops, r->n_cigar_op, cig, need)) < 0;
    if (bad) {
        cig[0] = '*';
        cig[1] = 0;
    }
    if ((map_flags & AS_F_KEEP_READS) && r->strand == 1 && r->seq_id >= 0 && len > 0) {
        rc = len <= sizeof seqbuf ? seqbuf : malloc(len);
        if (!rc) {
            if (heap) free(cig);
            return AS_ENOMEM;
        }
        rc_heap = rc != seqbuf;
        for (uint32_t i = 0; i < len; i++) {
            const char c = seq[len - 1 - i];
            /* _rev_comp_in_place (alnmain.c:31-47): either case complements to upper case, anything else becomes N */
            rc[i] = (c == 'A' || c == 'a') ? 'T' : (c == 'C' || c == 'c') ? 'G' : (c == 'G' || c == 'g') ? 'C' : (c == 'T' || c == 't') ? 'A' : 'N';
        }
        seq = rc;
    }
    /* seq_lookup failed: the reference prints uninitialised RNAME/POS (alnmain.c:409,460); defined here as * / 0 */
    const char *rname = r->seq_id >= 0 && r->seq_id < ix->view.mta_len ? ix->names[r->seq_id] : "*";
    const long pos = r->seq_id >= 0 ? (long)(r->off + 1) : 0L;
    fprintf(out, "%.*s\t%d\t%s\t%ld\t%d\t%s\t*\t%ld\t%d\t%.*s\t%.*s\tED:I:%d\n", (int)qname_len, qname, r->flag, rname, pos,
            r->mapq, cig, 0L, 0, (int)len, seq, (int)len, qual ? qual : "", r->ed);
    if (heap) free(cig);
    if (rc_heap) free(rc);
    return AS_OK;
}
