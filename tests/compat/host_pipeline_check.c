/*
 * tests/compat/host_pipeline_check.c -- the multi-threaded ingest (ingest.c) and emit (emit.c) of libaccseq_host.so against
 * the one-thread reader (fastq.c: kseq's grammar, accaln.c:45-58 / kseq.h:175-214) and the record-at-a-time SAM printer
 * (sam.c: alnmain.c:488-526) on the same FASTQ -- no GPU needed: the "results" are synthetic records derived from the reads.
 *
 *   host_pipeline_check <fastq (plain or .gz)> <threads> <block_bytes>        exit 0 = identical
 */
#define _GNU_SOURCE
#include "accseq_host.h"

#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define REQUIRE(c)                                                                \
    do {                                                                          \
        if (!(c)) {                                                               \
            fprintf(stderr, "host_pipeline_check: line %d: %s\n", __LINE__, #c);  \
            return 1;                                                             \
        }                                                                         \
    } while (0)

static uint64_t mix(uint64_t x) {
    x += 0x9E3779B97F4A7C15ULL;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ULL;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBULL;
    return x ^ (x >> 31);
}

int main(int argc, char **argv) {
    if (argc != 4) return 2;
    const int threads = atoi(argv[2]);
    const size_t block = (size_t)atol(argv[3]);
    /* the whole file through the serial reader */
    as_fastq *fq = as_fastq_open(argv[1]);
    REQUIRE(fq != NULL);
    as_read_batch all;
    REQUIRE(as_fastq_next_batch(fq, (size_t)1 << 30, &all, NULL) == AS_OK);
    as_fastq_close(fq);
    /* a two-sequence index stub for the printer */
    as_host_index ix;
    memset(&ix, 0, sizeof ix);
    as_mta mta[2] = {{0, 150000}, {300000, 70000}};
    char *names[2] = {"seqA", "seqB"};
    ix.mta = mta;
    ix.names = names;
    ix.view.mta_len = 2;

    as_ingest *in = as_ingest_open(argv[1], threads, block);
    REQUIRE(in != NULL);
    as_emitter *em = as_emitter_new(threads);
    size_t seen = 0;
    int batches = 0;
    char *want = NULL, *got = NULL;
    size_t want_len = 0, got_cap = 0;
    for (;;) {
        as_packed_batch b;
        REQUIRE(as_ingest_next(in, &b, NULL, NULL) == AS_OK);
        if (b.n == 0) break;
        batches++;
        REQUIRE(seen + b.n <= all.n);
        uint32_t *pk = calloc(b.wstride ? b.wstride : 1, 4);
        for (size_t i = 0; i < b.n; i++) {                       /* same records as the serial reader, same packing */
            const size_t k = seen + i;
            REQUIRE(b.lens[i] == all.lens[k]);
            REQUIRE(b.name_len[i] == strlen(all.names[k]) && !memcmp(b.name[i], all.names[k], b.name_len[i]));
            REQUIRE(!memcmp(b.seq[i], all.seqs + k * all.stride, b.lens[i]));
            REQUIRE(!memcmp(b.qual[i], all.quals[k], b.lens[i]));
            as_pack_read(all.seqs + k * all.stride, all.lens[k], pk, b.wstride);
            REQUIRE(!memcmp(pk, b.packed + i * b.wstride, b.wstride * 4));
        }
        free(pk);
        /* synthetic results in three "GPU shards": every kind of record the path produces */
        as_result_compact *res = calloc(b.n, sizeof *res);
        as_result *full = calloc(b.n, sizeof *full);
        size_t sb[4] = {0, b.n / 3, b.n / 3 + b.n / 5, b.n};
        uint8_t *runs[3];
        uint64_t runs_len[3] = {0, 0, 0};
        for (int s = 0; s < 3; s++) runs[s] = malloc((sb[s + 1] - sb[s]) * 16 + 16);
        uint8_t *rows = calloc(b.n, 16);
        for (int s = 0; s < 3; s++)
            for (size_t i = sb[s]; i < sb[s + 1]; i++) {
                const uint64_t h = mix(seen + i);
                const int kind = (int)(h % 11);
                as_result *r = &full[i];
                r->seq_id = kind == 0 ? -1 : (int)(h >> 8) % 2;
                r->strand = kind == 0 ? 0 : (int)(h >> 9) & 1;
                r->off = kind == 0 ? 0 : (h >> 16) % 60000;
                r->ed = kind == 1 ? -1 : (int)((h >> 40) % 9);
                r->flag = (kind == 0 || kind == 1) ? 4 : (r->strand ? 16 : 0);
                r->mapq = (r->flag & 4) ? 0 : 255;
                r->valid = !(r->flag & 4);
                int nr = (kind == 0 || kind == 1) ? 0 : 1 + (int)((h >> 50) % 9);
                if (kind == 2) nr = 16;
                uint8_t *row = rows + i * 16;
                uint32_t left = b.lens[i];
                for (int t = 0; t < nr; t++) {                     /* run bytes: M I M D M ..., a long M split at 63 */
                    const int cls = (t & 1) ? ((t & 2) ? 2 : 1) : 0;
                    uint32_t l = cls ? 1 + (uint32_t)((h >> (t + 3)) & 3) : (left > 63 ? 63 : (left ? left : 1));
                    if (kind == 3 && t == 1) { row[t] = (uint8_t)((63 << 2) | 0); continue; }     /* M continues: 63 + 63 */
                    row[t] = (uint8_t)((l << 2) | (unsigned)cls);
                    if (!cls) left -= l > left ? left : l;
                }
                r->n_cigar_op = nr;
                res[i].pos = (r->off & 0xFFFFFFFFFFULL) | ((uint64_t)(uint16_t)(r->seq_id + 1) << 40) | ((uint64_t)(uint8_t)r->flag << 56);
                res[i].ed = r->ed;
                res[i].n_runs = (uint32_t)nr | (r->strand ? 0x80000000u : 0u);
                memcpy(runs[s] + runs_len[s], row, (size_t)nr);
                runs_len[s] += (uint64_t)nr;
            }
        /* reference text: the record-at-a-time printer */
        size_t cap = 0;
        FILE *mem = open_memstream(&want, &cap);
        for (size_t i = 0; i < b.n; i++) {
            const size_t k = seen + i;
            REQUIRE(as_sam_record2(mem, &ix, all.names[k], strlen(all.names[k]), all.seqs + k * all.stride, all.quals[k], all.lens[k],
                                   &full[i], rows + i * 16, AS_F_RLE_CIGAR | AS_F_KEEP_READS) == AS_OK);
        }
        fclose(mem);
        want_len = cap;
        const char *seg[64];
        size_t seg_len[64];
        const uint8_t *rp[3] = {runs[0], runs[1], runs[2]};
        const int ns = as_emit_batch(em, &ix, &b, res, 3, sb, rp, runs_len, seg, seg_len);
        REQUIRE(ns == threads);
        size_t gl = 0;
        for (int s = 0; s < ns; s++) gl += seg_len[s];
        if (gl + 1 > got_cap) {
            got_cap = gl + 1;
            got = realloc(got, got_cap);
        }
        gl = 0;
        for (int s = 0; s < ns; s++) {
            memcpy(got + gl, seg[s], seg_len[s]);
            gl += seg_len[s];
        }
        REQUIRE(gl == want_len);
        REQUIRE(!memcmp(got, want, gl));
        /* and the record an as_result stands for comes back from the compact form */
        for (size_t i = 0; i < b.n; i += 7) {
            as_result back;
            as_result_from_compact(&res[i], b.lens[i], mta, 2, &back);
            REQUIRE(back.off == full[i].off && back.seq_id == full[i].seq_id && back.flag == full[i].flag && back.ed == full[i].ed &&
                    back.strand == full[i].strand && back.mapq == full[i].mapq && back.valid == full[i].valid &&
                    back.n_cigar_op == full[i].n_cigar_op);
        }
        free(want);
        want = NULL;
        free(res);
        free(full);
        free(rows);
        for (int s = 0; s < 3; s++) free(runs[s]);
        seen += b.n;
        as_packed_batch_free(&b, NULL);
    }
    REQUIRE(seen == all.n);
    printf("host_pipeline_check ok: %zu reads in %d batches, %d threads\n", seen, batches, threads);
    free(got);
    as_emitter_free(em);
    as_ingest_close(in);
    as_read_batch_free(&all, NULL);
    return 0;
}
