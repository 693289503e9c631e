/*
 * accaln_main.c -- `accaln_b200`: the alnmain.c-shaped host driver over the C ABI (plain C, no CUDA headers).
 *
 *   accaln_b200 ref.fa reads.fq [batch_size seed_len thres]          (argv of the reference: alnmain.c:560-571)
 *
 * Same inputs (the five index files accidx wrote beside ref.fa, a FASTQ), same stdout (SAM: header alnmain.c:62-75,
 * records alnmain.c:488-526) and the same closing `Sensitivity: valid/total` line on stderr (alnmain.c:541).  What
 * differs is the middle: where single_end() runs the loop nest alnmain.c:330-479 on the host, this driver hands the
 * refactor_reads_seq buffer of every batch to as_map_batch().  Defaults without argv[3..5] follow read_params
 * (alnmain.c:574-588): ./params if present, else 1000 / 20 / 300 -- except that the default batch is raised to 1M
 * reads, since a batch is the unit of GPU work and results do not depend on it.
 * ACCSEQ_DEVICE selects the GPU (default 0); ACCSEQ_RG_ID fixes the @RG id (the reference uses time(NULL)).
 */
#include "accseq_b200.h"
#include "accseq_host.h"

#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

static double now_s(void) {
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

static double t_start;
static void logline(const char *msg) { fprintf(stderr, "[ %-7s %8.2lf ] %s\n", "Verbose", now_s() - t_start, msg); }   /* mlog/logger.c:75 */

int main(int argc, const char **argv) {
    if (argc != 3 && argc != 6) {
        fprintf(stderr, "usage: %s ref.fa reads.fq [batch_size seed_len thres]\n", argv[0]);
        return 1;
    }
    t_start = now_s();
    as_params p;
    as_params_default(&p);
    /* compact D2H: run-length CIGAR bytes, reads not copied back (the emitter reverse-complements strand-1 records);
     * ACCALN_FULL_COPY=1 keeps the reference's layout (one op byte per column, reads reverse-complemented in place) */
    if (!getenv("ACCALN_FULL_COPY")) p.flags |= AS_F_RLE_CIGAR | AS_F_KEEP_READS;
    unsigned long batch = 1000000ul;
    if (argc == 6) {
        batch = strtoul(argv[3], NULL, 10);
        p.seed_len = (uint32_t)strtoul(argv[4], NULL, 10);
        p.thres = (uint32_t)strtoul(argv[5], NULL, 10);
    } else {
        FILE *fp = fopen("params", "r");                      /* alnmain.c:574-588 */
        if (fp) {
            unsigned long b;
            unsigned s, t;
            if (fscanf(fp, "%lu %u %u", &b, &s, &t) == 3) { batch = b; p.seed_len = s; p.thres = t; }
            fclose(fp);
        }
    }
    if (batch == 0) batch = 1;
    if (getenv("ACCSEQ_MYERS")) p.flags |= AS_F_MYERS;

    logline("Begin loading index");
    as_host_index *ix = NULL;
    int rc = as_index_load(argv[1], &ix);
    if (rc != AS_OK) {
        fprintf(stderr, "accaln_b200: cannot load the index beside %s (%d)\n", argv[1], rc);
        return 2;
    }
    const int device = getenv("ACCSEQ_DEVICE") ? atoi(getenv("ACCSEQ_DEVICE")) : 0;
    as_gpu *h = NULL;
    rc = as_gpu_open(&h, device, &ix->view);
    if (rc != AS_OK) {
        fprintf(stderr, "accaln_b200: as_gpu_open failed (%d): %s\n", rc, as_last_error());
        return 3;
    }
    const long rg = getenv("ACCSEQ_RG_ID") ? atol(getenv("ACCSEQ_RG_ID")) : (long)time(NULL);
    static char outbuf[4 << 20];
    setvbuf(stdout, outbuf, _IOFBF, sizeof outbuf);           /* alnmain.c:486 */
    as_sam_header(stdout, ix, rg);

    as_fastq *fq = as_fastq_open(argv[2]);
    if (!fq) {
        fprintf(stderr, "accaln_b200: cannot open %s\n", argv[2]);
        return 4;
    }
    logline("Begin processing queries");
    unsigned long long total = 0, valid = 0;
    as_result *res = NULL;
    uint8_t *store = NULL;
    size_t res_cap = 0, store_cap = 0;
    for (;;) {
        as_read_batch b;
        rc = as_fastq_next_batch(fq, batch, &b, as_host_alloc);      /* pinned: the batch buffer is the upload unit */
        if (rc != AS_OK) {
            fprintf(stderr, "accaln_b200: malformed FASTQ record after %llu reads (%d)\n", total, rc);
            return 5;
        }
        if (b.n == 0) break;
        const size_t store_stride = 2 * (size_t)b.max_len;             /* alnmain.c:316-320 */
        if (b.n > res_cap) {
            as_host_free(res);
            res = as_host_alloc(b.n * sizeof *res);
            res_cap = b.n;
        }
        if (b.n * store_stride > store_cap) {
            as_host_free(store);
            store = as_host_alloc(b.n * store_stride);
            store_cap = b.n * store_stride;
        }
        if (!res || !store) {
            fprintf(stderr, "accaln_b200: out of pinned memory\n");
            return 6;
        }
        rc = as_map_batch(h, &p, b.seqs, b.stride, b.lens, b.n, res, store, store_stride);   /* alnmain.c:330-479 */
        if (rc != AS_OK) {
            fprintf(stderr, "accaln_b200: as_map_batch failed (%d): %s\n", rc, as_last_error());
            return 7;
        }
        for (size_t i = 0; i < b.n; i++) {                              /* alnmain.c:488-526 */
            valid += res[i].valid != 0;
            as_sam_record2(stdout, ix, b.names[i], strlen(b.names[i]), b.seqs + i * b.stride, b.quals[i], b.lens[i],
                           &res[i], store + i * store_stride, p.flags);
        }
        fflush(stdout);
        total += b.n;
        as_read_batch_free(&b, as_host_free);
    }
    logline("Done aligning");
    fprintf(stderr, "[ %-7s %8.2lf ] Sensitivity: %llu/%llu\n", "Info", now_s() - t_start, valid, total);   /* alnmain.c:541 */
    as_stats st;
    if (as_gpu_stats(h, &st) == AS_OK)
        fprintf(stderr, "[ %-7s %8.2lf ] seeds %llu, SA hits %llu, DP cells %llu\n", "Info", now_s() - t_start,
                (unsigned long long)st.seeds, (unsigned long long)st.sa_hits, (unsigned long long)st.dp_cells);
    as_host_free(res);
    as_host_free(store);
    as_fastq_close(fq);
    as_gpu_close(h);
    as_index_free(ix);
    return 0;
}
