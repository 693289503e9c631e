/*
 * accaln_main.c -- `accaln_b200`: the alnmain.c-shaped host driver over the C ABI (plain C, no CUDA headers).
 *
 *   accaln_b200 ref.fa reads.fq [batch_size seed_len thres]          (argv of the reference: alnmain.c:560-571)
 *   accaln_b200 ref.fa r1.fq r2.fq [batch_size seed_len thres]       (paired end: a stub in the reference, alnmain.c:552-555)
 *
 * Same inputs (the five index files accidx wrote beside ref.fa, a FASTQ), same stdout (SAM: header alnmain.c:62-75,
 * records alnmain.c:488-526) and the same closing `Sensitivity: valid/total` line on stderr (alnmain.c:541).  What
 * differs is the middle: where single_end() runs the loop nest alnmain.c:330-479 on the host, batch after batch on one
 * thread, this driver runs three stages side by side:
 *
 *   ingest   N threads cut the FASTQ into records and pack them 2 bits per base into pinned memory (ingest.c);
 *   map      one host thread per GPU: the batch is split into contiguous read ranges, one per device, each mapped by
 *            as_map_batch_packed against that device's replica of the index -- no collective, the ranges are independent;
 *   emit     N threads format SAM text from the compact records (emit.c); the text is written in read order, so the
 *            output does not depend on the number of GPUs or threads.
 *
 * While batch k is mapped, batch k+1 is parsed and batch k-1 printed.  Defaults without argv[3..5] follow read_params
 * (alnmain.c:574-588): ./params if present, else 1000 / 20 / 300 -- except that the default batch is 1 M reads, since a
 * batch is the unit of GPU work and results do not depend on it (batch_size caps the reads per as_map_batch_packed call).
 *
 * Environment (host-program knobs; the libraries read none):
 *   ACCSEQ_DEVICES=0,1,..|all   GPUs to use (default: ACCSEQ_DEVICE, else 0)      ACCSEQ_THREADS=n  ingest/emit threads
 *   ACCSEQ_RG_ID=<id>           fixes the @RG id (the reference uses time(NULL))  ACCSEQ_MYERS=1    Myers pass (serial path)
 *   ACCALN_FULL_COPY=1 / ACCALN_SERIAL=1   the one-thread path over as_map_batch (reference buffer layout / run bytes):
 *                               taken by itself for input that is not 4-line FASTQ (kseq's full grammar: fastq.c)
 */
#define _GNU_SOURCE
#include "accseq_b200.h"
#include "accseq_host.h"

#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include <unistd.h>

static double now_s(void) {
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

static double t_start;
static void logline(const char *msg) { fprintf(stderr, "[ %-7s %8.2lf ] %s\n", "Verbose", now_s() - t_start, msg); }   /* mlog/logger.c:75 */

#define MAX_GPUS 16

/* ------------------------------------------------------------------------------------------------
 * serial path: fastq.c -> as_map_batch -> sam.c, one thread, one GPU (kseq's whole grammar; the reference's buffer layout)
 * ---------------------------------------------------------------------------------------------- */
static int run_serial(const char *fq_path, as_gpu *h, as_params p, const as_host_index *ix, unsigned long batch,
                      unsigned long long *total_out, unsigned long long *valid_out) {
    as_fastq *fq = as_fastq_open(fq_path);
    if (!fq) {
        fprintf(stderr, "accaln_b200: cannot open %s\n", fq_path);
        return 4;
    }
    unsigned long long total = 0, valid = 0;
    as_result *res = NULL;
    uint8_t *store = NULL;
    size_t res_cap = 0, store_cap = 0;
    for (;;) {
        as_read_batch b;
        int rc = as_fastq_next_batch(fq, batch, &b, as_host_alloc);      /* pinned: the batch buffer is the upload unit */
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
    as_host_free(res);
    as_host_free(store);
    as_fastq_close(fq);
    *total_out = total;
    *valid_out = valid;
    return 0;
}

/* ------------------------------------------------------------------------------------------------
 * paired-end mode (DESIGN.md "Paired-end mode"): the reference dispatches argc == 4 to pair_end(), a stub that returns -1
 * (alnmain.c:552-571), so there is no behaviour to match -- this is the mode its API reserves fields for (context.read2,
 * result.r_name / r_off, accaln.h:74,97-99), built on the single-end path: both mates go through as_map_batch_packed
 * exactly as single reads do, then the records are printed pair by pair with the mate fields filled in.
 * ---------------------------------------------------------------------------------------------- */
static int run_paired(const char *fq1, const char *fq2, as_gpu **gpus, int n_dev, const as_params *p, const as_host_index *ix,
                      unsigned long batch, unsigned long long *total_out, unsigned long long *valid_out) {
    as_fastq *f1 = as_fastq_open(fq1), *f2 = as_fastq_open(fq2);
    if (!f1 || !f2) {
        fprintf(stderr, "accaln_b200: cannot open %s\n", f1 ? fq2 : fq1);
        return 4;
    }
    const long max_insert = getenv("ACCSEQ_MAX_INSERT") ? atol(getenv("ACCSEQ_MAX_INSERT")) : 1000;
    as_emitter *em = as_emitter_new(1);
    unsigned long long total = 0, valid = 0;
    int status = 0;
    while (status == 0) {
        as_read_batch b[2];
        const int r1 = as_fastq_next_batch(f1, batch, &b[0], NULL), r2 = as_fastq_next_batch(f2, batch, &b[1], NULL);
        if (r1 != AS_OK || r2 != AS_OK) {
            fprintf(stderr, "accaln_b200: malformed FASTQ record after %llu pairs\n", total);
            status = 5;
            break;
        }
        if (b[0].n != b[1].n) {
            fprintf(stderr, "accaln_b200: %s and %s do not hold the same number of reads\n", fq1, fq2);
            status = 5;
            break;
        }
        const size_t n = b[0].n;
        if (n == 0) break;
        uint32_t *packed[2] = {0};
        as_result_compact *res[2] = {0};
        uint8_t *runs[2] = {0};
        uint32_t *nlen[2] = {0};
        for (int m = 0; m < 2 && status == 0; m++) {
            const size_t wst = ((size_t)b[m].max_len + 15) / 16 ? ((size_t)b[m].max_len + 15) / 16 : 1;
            uint64_t bases = 0, got = 0;
            for (size_t i = 0; i < n; i++) bases += b[m].lens[i];
            const size_t cap = (size_t)(2 * bases) + 4096;
            packed[m] = as_host_alloc(n * wst * sizeof(uint32_t));
            res[m] = as_host_alloc(n * sizeof(as_result_compact));
            runs[m] = as_host_alloc(cap);
            nlen[m] = malloc(n * sizeof(uint32_t));
            if (!packed[m] || !res[m] || !runs[m] || !nlen[m]) {
                status = 6;
                break;
            }
            for (size_t i = 0; i < n; i++) nlen[m][i] = (uint32_t)strlen(b[m].names[i]);
            as_pack_reads(b[m].seqs, b[m].stride, b[m].lens, n, packed[m], wst);
            /* mate 1 on the first device, mate 2 on the second when there is one: the two files are independent batches */
            const int rc = as_map_batch_packed(gpus[n_dev > 1 ? m : 0], p, packed[m], wst, b[m].lens, n, res[m], runs[m], cap, &got);
            if (rc != AS_OK) {
                fprintf(stderr, "accaln_b200: as_map_batch_packed failed (%d): %s\n", rc, as_last_error());
                status = 7;
            }
        }
        if (status == 0) {
            const char **seq[2], **qual[2];
            for (int m = 0; m < 2; m++) {
                seq[m] = malloc(n * sizeof(char *));
                qual[m] = malloc(n * sizeof(char *));
                for (size_t i = 0; i < n; i++) {
                    seq[m][i] = b[m].seqs + i * b[m].stride;
                    qual[m][i] = b[m].quals[i];
                }
            }
            const char *text;
            size_t text_len;
            if (as_emit_pairs(em, ix, n, (const char *const *)b[0].names, nlen[0], seq[0], qual[0], b[0].lens, res[0], runs[0],
                              (const char *const *)b[1].names, nlen[1], seq[1], qual[1], b[1].lens, res[1], runs[1], max_insert,
                              &text, &text_len) != AS_OK)
                status = 6;
            else if (fwrite(text, 1, text_len, stdout) != text_len)
                status = 8;
            fflush(stdout);
            for (size_t i = 0; i < n; i++) valid += !(AS_RC_FLAG(res[0][i]) & 4u) + !(AS_RC_FLAG(res[1][i]) & 4u);
            total += 2 * n;
            for (int m = 0; m < 2; m++) {
                free(seq[m]);
                free(qual[m]);
            }
        }
        for (int m = 0; m < 2; m++) {
            as_host_free(packed[m]);
            as_host_free(res[m]);
            as_host_free(runs[m]);
            free(nlen[m]);
            as_read_batch_free(&b[m], NULL);
        }
    }
    as_emitter_free(em);
    as_fastq_close(f1);
    as_fastq_close(f2);
    *total_out = total;
    *valid_out = valid;
    return status;
}

/* ------------------------------------------------------------------------------------------------
 * pipelined path
 * ---------------------------------------------------------------------------------------------- */
/* Pinned buffers are recycled: cudaHostAlloc costs about a millisecond per 4 MB, more than mapping the reads they carry.
 * A block keeps its capacity in a 64-byte header; pin_get hands out the smallest cached block that fits. */
#define POOL_SLOTS 64
static struct {
    pthread_mutex_t mu;
    void *blk[POOL_SLOTS];
} g_pool = {PTHREAD_MUTEX_INITIALIZER, {0}};

static void *pin_get(size_t bytes) {
    pthread_mutex_lock(&g_pool.mu);
    int best = -1;
    for (int i = 0; i < POOL_SLOTS; i++) {
        if (!g_pool.blk[i]) continue;
        const size_t cap = *(size_t *)g_pool.blk[i];
        if (cap >= bytes && cap <= 2 * bytes + (1 << 20) && (best < 0 || cap < *(size_t *)g_pool.blk[best])) best = i;
    }
    void *b = NULL;
    if (best >= 0) {
        b = g_pool.blk[best];
        g_pool.blk[best] = NULL;
    }
    pthread_mutex_unlock(&g_pool.mu);
    if (!b) {
        const size_t cap = bytes + bytes / 8 + 4096;
        b = as_host_alloc(cap + 64);
        if (!b) return NULL;
        *(size_t *)b = cap;
    }
    return (char *)b + 64;
}

static void pin_put(void *p) {
    if (!p) return;
    void *b = (char *)p - 64;
    pthread_mutex_lock(&g_pool.mu);
    for (int i = 0; i < POOL_SLOTS; i++)
        if (!g_pool.blk[i]) {
            g_pool.blk[i] = b;
            b = NULL;
            break;
        }
    pthread_mutex_unlock(&g_pool.mu);
    if (b) as_host_free(b);
}

static void pin_drain(void) {
    for (int i = 0; i < POOL_SLOTS; i++) {
        if (g_pool.blk[i]) as_host_free(g_pool.blk[i]);
        g_pool.blk[i] = NULL;
    }
}

static double g_stage_s[4];        /* busy seconds of ingest / map / emit / write, for the closing Info line */
typedef struct {
    as_packed_batch b;
    as_result_compact *res;        /* pinned, b.n records */
    uint8_t *runs[MAX_GPUS];       /* pinned, per shard */
    size_t runs_cap[MAX_GPUS];
    uint64_t runs_len[MAX_GPUS];
    size_t shard_begin[MAX_GPUS + 1];
    int n_shards;
    int last;                      /* end-of-input marker (b.n == 0) */
    int rc;
} job;

/* a one-slot hand-over between two stages */
typedef struct {
    pthread_mutex_t mu;
    pthread_cond_t cv;
    job *slot;
} chan;

static void chan_init(chan *c) {
    pthread_mutex_init(&c->mu, NULL);
    pthread_cond_init(&c->cv, NULL);
    c->slot = NULL;
}
static void chan_put(chan *c, job *j) {
    pthread_mutex_lock(&c->mu);
    while (c->slot) pthread_cond_wait(&c->cv, &c->mu);
    c->slot = j;
    pthread_cond_broadcast(&c->cv);
    pthread_mutex_unlock(&c->mu);
}
static job *chan_get(chan *c) {
    pthread_mutex_lock(&c->mu);
    while (!c->slot) pthread_cond_wait(&c->cv, &c->mu);
    job *j = c->slot;
    c->slot = NULL;
    pthread_cond_broadcast(&c->cv);
    pthread_mutex_unlock(&c->mu);
    return j;
}

typedef struct {
    as_ingest *in;                 /* closed by main once the last batch is printed: record text points into its mapping */
    chan *to_map;
    int rc;
} ingest_arg;

static void *ingest_main(void *v) {
    ingest_arg *a = v;
    as_ingest *in = a->in;
    if (!in) a->rc = 4;
    for (;;) {
        job *j = calloc(1, sizeof *j);
        if (in && a->rc == 0) {
            const double t0 = now_s();
            const int rc = as_ingest_next(in, &j->b, pin_get, pin_put);
            g_stage_s[0] += now_s() - t0;
            if (rc != AS_OK) a->rc = rc == AS_EIO ? 50 : 6;
        }
        if (!in || a->rc != 0 || j->b.n == 0) {
            j->last = 1;
            j->rc = a->rc;
            chan_put(a->to_map, j);
            break;
        }
        chan_put(a->to_map, j);
    }
    return NULL;
}

typedef struct {
    as_gpu *h;
    const as_params *p;
    job *j;
    int shard;
    unsigned long batch;
    int rc;
} shard_arg;

static void *shard_main(void *v) {
    shard_arg *a = v;
    job *j = a->j;
    const size_t r0 = j->shard_begin[a->shard], r1 = j->shard_begin[a->shard + 1];
    uint64_t used = 0;
    a->rc = AS_OK;
    /* batch_size caps the reads of one call (alnmain.c:300: the reference's batch is its unit of work too) */
    for (size_t s = r0; s < r1 && a->rc == AS_OK; s += a->batch) {
        const size_t cn = r1 - s < a->batch ? r1 - s : a->batch;
        uint64_t got = 0;
        const double t0 = now_s();
        a->rc = as_map_batch_packed(a->h, a->p, j->b.packed + s * j->b.wstride, j->b.wstride, j->b.lens + s, cn, j->res + s,
                                    j->runs[a->shard] + used, j->runs_cap[a->shard] - used, &got);
        used += got;
        if (getenv("ACCSEQ_VERBOSE")) {
            float ms[8];
            double us[3];
            as_gpu_last_timing(a->h, ms);
            as_gpu_last_host_timing(a->h, us);
            fprintf(stderr, "  shard %d: %zu reads, call %.2f ms (device total %.2f: part1 %.2f part2 %.2f gact %.2f h2d %.2f d2h %.2f; host enqueue %.0f us wait %.0f us)\n",
                    a->shard, cn, 1e3 * (now_s() - t0), ms[7], ms[0], ms[1], ms[2], ms[5], ms[6], us[0], us[1]);
        }
        if (a->rc != AS_OK && a->rc != AS_ERANGE)              /* as_last_error() is per thread: say it here */
            fprintf(stderr, "accaln_b200: as_map_batch_packed failed on GPU shard %d (%d): %s\n", a->shard, a->rc, as_last_error());
    }
    j->runs_len[a->shard] = used;
    return NULL;
}

typedef struct {
    chan *from_ingest, *to_emit;
    as_gpu **gpus;
    int n_gpus;
    const as_params *p;
    unsigned long batch;
    int rc;
} map_arg;

static void *map_main(void *v) {
    map_arg *a = v;
    for (;;) {
        job *j = chan_get(a->from_ingest);
        if (!j->last && a->rc == 0) {
            const size_t n = j->b.n;
            const int G = (size_t)a->n_gpus < n ? a->n_gpus : (n ? (int)n : 1);
            j->n_shards = G;
            for (int s = 0; s <= G; s++) j->shard_begin[s] = (size_t)(((unsigned __int128)n * (unsigned)s) / (unsigned)G);
            const double t0 = now_s();
            if (getenv("ACCSEQ_VERBOSE")) fprintf(stderr, "map: batch of %zu reads at %.3f\n", n, now_s() - t_start);
            j->res = pin_get(n * sizeof *j->res);
            int ok = j->res != NULL;
            for (int s = 0; s < G && ok; s++) {
                /* run bytes: 2 per column would hold any CIGAR; reads at a few percent error need a twentieth of that, so start
                 * at 16 per read plus an eighth of a byte per base and let AS_ERANGE ask for more */
                uint64_t bases = 0;
                for (size_t i = j->shard_begin[s]; i < j->shard_begin[s + 1]; i++) bases += j->b.lens[i];
                j->runs_cap[s] = (j->shard_begin[s + 1] - j->shard_begin[s]) * 16 + (size_t)(bases / 8) + 4096;
                j->runs[s] = pin_get(j->runs_cap[s]);
                ok = j->runs[s] != NULL;
            }
            if (!ok) {
                fprintf(stderr, "accaln_b200: out of pinned memory\n");
                a->rc = 6;
            } else {
                pthread_t th[MAX_GPUS];
                shard_arg sa[MAX_GPUS];
                for (int s = 0; s < G; s++) {
                    sa[s] = (shard_arg){a->gpus[s], a->p, j, s, a->batch, 0};
                    pthread_create(&th[s], NULL, shard_main, &sa[s]);
                }
                for (int s = 0; s < G; s++) pthread_join(th[s], NULL);
                for (int s = 0; s < G; s++) {
                    if (sa[s].rc == AS_ERANGE) {                         /* a shard needed more room for its run bytes: redo it */
                        uint64_t bases = 0;
                        for (size_t i = j->shard_begin[s]; i < j->shard_begin[s + 1]; i++) bases += j->b.lens[i];
                        pin_put(j->runs[s]);
                        j->runs_cap[s] = (size_t)(2 * bases) + 4096;
                        j->runs[s] = pin_get(j->runs_cap[s]);
                        if (!j->runs[s]) {
                            a->rc = 6;
                            break;
                        }
                        shard_main(&sa[s]);
                    }
                    if (sa[s].rc != AS_OK) {
                        a->rc = 7;
                        break;
                    }
                }
            }
            j->rc = a->rc;
            g_stage_s[1] += now_s() - t0;
        }
        const int last = j->last;
        chan_put(a->to_emit, j);
        if (last) break;
    }
    return NULL;
}

static void job_free(job *j) {
    as_packed_batch_free(&j->b, pin_put);
    pin_put(j->res);
    for (int s = 0; s < MAX_GPUS; s++) pin_put(j->runs[s]);
    free(j);
}

static int parse_devices(const char *spec, int *dev, int have) {
    int n = 0;
    if (!spec || !*spec) return 0;
    if (!strcmp(spec, "all")) {
        for (int i = 0; i < have && i < MAX_GPUS; i++) dev[n++] = i;
        return n;
    }
    const char *p = spec;
    while (*p && n < MAX_GPUS) {
        char *end;
        long a = strtol(p, &end, 10), b = a;
        if (end == p) return -1;
        if (*end == '-') {
            p = end + 1;
            b = strtol(p, &end, 10);
            if (end == p) return -1;
        }
        for (long d = a; d <= b && n < MAX_GPUS; d++) dev[n++] = (int)d;
        p = *end == ',' ? end + 1 : end;
        if (*end && *end != ',') return -1;
    }
    return n;
}

int main(int argc, const char **argv) {
    /* argc 3 / 6: single end (alnmain.c:560-571).  argc 4 / 7: the reference's paired-end invocation `ref.fa r1.fq r2.fq`, which
     * it dispatches to a stub (alnmain.c:552-555); here it runs run_paired() above */
    if (argc != 3 && argc != 6 && argc != 4 && argc != 7) {
        fprintf(stderr, "usage: %s ref.fa reads.fq [batch_size seed_len thres]\n       %s ref.fa reads_1.fq reads_2.fq [batch_size seed_len thres]\n",
                argv[0], argv[0]);
        return 1;
    }
    const int paired = argc == 4 || argc == 7;
    const int a0 = paired ? 4 : 3;                            /* index of batch_size in argv */
    t_start = now_s();
    as_params p;
    as_params_default(&p);
    unsigned long batch = 1000000ul;
    if (argc == a0 + 3) {
        batch = strtoul(argv[a0], NULL, 10);
        p.seed_len = (uint32_t)strtoul(argv[a0 + 1], NULL, 10);
        p.thres = (uint32_t)strtoul(argv[a0 + 2], NULL, 10);
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

    logline("Begin loading index");
    as_host_index *ix = NULL;
    int rc = as_index_load(argv[1], &ix);
    if (rc != AS_OK) {
        fprintf(stderr, "accaln_b200: cannot load the index beside %s (%d)\n", argv[1], rc);
        return 2;
    }
    int dev[MAX_GPUS], n_dev = parse_devices(getenv("ACCSEQ_DEVICES"), dev, as_device_count());
    if (n_dev < 0) {
        fprintf(stderr, "accaln_b200: cannot parse ACCSEQ_DEVICES=%s\n", getenv("ACCSEQ_DEVICES"));
        return 1;
    }
    if (n_dev == 0) {
        dev[0] = getenv("ACCSEQ_DEVICE") ? atoi(getenv("ACCSEQ_DEVICE")) : 0;
        n_dev = 1;
    }
    const int serial = !paired && (getenv("ACCALN_FULL_COPY") || getenv("ACCALN_SERIAL") || getenv("ACCSEQ_MYERS"));
    if (serial) n_dev = 1;
    if (paired && n_dev > 2) n_dev = 2;
    as_gpu *gpus[MAX_GPUS] = {0};
    for (int g = 0; g < n_dev; g++) {                         /* the index is replicated: one copy per device */
        rc = as_gpu_open(&gpus[g], dev[g], &ix->view);
        if (rc != AS_OK) {
            fprintf(stderr, "accaln_b200: as_gpu_open failed on device %d (%d): %s\n", dev[g], rc, as_last_error());
            return 3;
        }
    }
    const long rg = getenv("ACCSEQ_RG_ID") ? atol(getenv("ACCSEQ_RG_ID")) : (long)time(NULL);
    static char outbuf[4 << 20];
    setvbuf(stdout, outbuf, _IOFBF, sizeof outbuf);           /* alnmain.c:486 */
    as_sam_header(stdout, ix, rg);
    fflush(stdout);

    logline("Begin processing queries");
    unsigned long long total = 0, valid = 0;
    int status = 0;
    if (paired) {
        status = run_paired(argv[2], argv[3], gpus, n_dev, &p, ix, batch, &total, &valid);
    } else if (!serial) {
        long ncpu = sysconf(_SC_NPROCESSORS_ONLN);
        /* parsing + packing a read costs about twice what printing it does: 5/8 of the cores parse, 3/8 print */
        int threads = getenv("ACCSEQ_THREADS") ? atoi(getenv("ACCSEQ_THREADS")) : (int)(ncpu > 2 ? (5 * ncpu + 7) / 8 : 1);
        int ethreads = getenv("ACCSEQ_THREADS") ? threads : (int)(ncpu > 2 ? (3 * ncpu + 7) / 8 : 1);
        if (threads < 1) threads = 1;
        if (threads > 64) threads = 64;
        if (ethreads < 1) ethreads = 1;
        if (ethreads > 64) ethreads = 64;
        chan c1, c2;
        chan_init(&c1);
        chan_init(&c2);
        ingest_arg ia = {as_ingest_open(argv[2], threads, 0), &c1, 0};
        map_arg ma = {&c1, &c2, gpus, n_dev, &p, batch, 0};
        pthread_t t_in, t_map;
        pthread_create(&t_in, NULL, ingest_main, &ia);
        pthread_create(&t_map, NULL, map_main, &ma);
        as_emitter *em = as_emitter_new(ethreads);
        const char *seg[64];
        size_t seg_len[64];
        int fallback = 0;
        for (;;) {
            job *j = chan_get(&c2);
            if (j->rc == 50 && total == 0) fallback = 1;      /* not 4-line FASTQ, nothing printed yet: the serial reader */
            else if (j->rc != 0 && status == 0) status = j->rc == 50 ? 5 : j->rc;
            if (!j->last && j->rc == 0 && status == 0) {
                const uint8_t *rp[MAX_GPUS];
                for (int s = 0; s < j->n_shards; s++) rp[s] = j->runs[s];
                const double t0 = now_s();
                const int ns = as_emit_batch(em, ix, &j->b, j->res, j->n_shards, j->shard_begin, rp, j->runs_len, seg, seg_len);
                const double t1 = now_s();
                if (ns < 0) status = 6;
                /* the segments go straight to the descriptor: stdio would copy 350 bytes per read once more */
                fflush(stdout);
                for (int s = 0; s < ns && status == 0; s++) {
                    size_t done = 0;
                    while (done < seg_len[s]) {
                        const ssize_t w = write(STDOUT_FILENO, seg[s] + done, seg_len[s] - done);
                        if (w <= 0) {
                            status = 8;
                            break;
                        }
                        done += (size_t)w;
                    }
                }
                g_stage_s[2] += t1 - t0;
                g_stage_s[3] += now_s() - t1;
                for (size_t i = 0; i < j->b.n; i++) valid += !(AS_RC_FLAG(j->res[i]) & 4u);
                total += j->b.n;
            }
            const int last = j->last;
            job_free(j);
            if (last) break;
        }
        pthread_join(t_in, NULL);
        pthread_join(t_map, NULL);
        as_emitter_free(em);
        as_ingest_close(ia.in);
        pin_drain();
        fprintf(stderr, "[ %-7s %8.2lf ] stage busy seconds: ingest %.3f, map %.3f, emit %.3f, write %.3f (%d ingest threads, %d emit threads)\n",
                "Info", now_s() - t_start, g_stage_s[0], g_stage_s[1], g_stage_s[2], g_stage_s[3], threads, ethreads);
        if (fallback) {
            logline("input is not 4-line FASTQ: one-thread reader");
            p.flags |= AS_F_RLE_CIGAR | AS_F_KEEP_READS;
            status = run_serial(argv[2], gpus[0], p, ix, batch, &total, &valid);
        } else if (status == 5) {
            fprintf(stderr, "accaln_b200: malformed FASTQ record after %llu reads\n", total);
        } else if (status == 4) {
            fprintf(stderr, "accaln_b200: cannot open %s\n", argv[2]);
        }
    } else {
        /* compact D2H: run-length CIGAR bytes, reads not copied back (the emitter reverse-complements strand-1 records);
         * ACCALN_FULL_COPY=1 keeps the reference's layout (one op byte per column, reads reverse-complemented in place) */
        if (!getenv("ACCALN_FULL_COPY")) p.flags |= AS_F_RLE_CIGAR | AS_F_KEEP_READS;
        if (getenv("ACCSEQ_MYERS")) p.flags |= AS_F_MYERS;
        status = run_serial(argv[2], gpus[0], p, ix, batch, &total, &valid);
    }
    if (status != 0) return status;
    logline("Done aligning");
    fprintf(stderr, "[ %-7s %8.2lf ] Sensitivity: %llu/%llu\n", "Info", now_s() - t_start, valid, total);   /* alnmain.c:541 */
    /* the only cross-GPU step: the shard counters, summed (what one ncclAllReduce would do across processes) */
    as_stats sum;
    memset(&sum, 0, sizeof sum);
    for (int g = 0; g < n_dev; g++) {
        as_stats st;
        if (as_gpu_stats(gpus[g], &st) != AS_OK) continue;
        sum.seeds += st.seeds;
        sum.sa_hits += st.sa_hits;
        sum.dp_cells += st.dp_cells;
        sum.total += st.total;
        sum.valid += st.valid;
    }
    fprintf(stderr, "[ %-7s %8.2lf ] %d GPU(s): mapped %llu/%llu, seeds %llu, SA hits %llu, DP cells %llu\n", "Info", now_s() - t_start,
            n_dev, (unsigned long long)sum.valid, (unsigned long long)sum.total, (unsigned long long)sum.seeds,
            (unsigned long long)sum.sa_hits, (unsigned long long)sum.dp_cells);
    for (int g = 0; g < n_dev; g++) as_gpu_close(gpus[g]);
    as_index_free(ix);
    return 0;
}
