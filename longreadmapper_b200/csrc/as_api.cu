/*
 * as_api.cu -- the batched entry points of include/accseq_b200.h: as_map_batch replaces the loop nest
 * alnmain.c:330-479 for one batch; the *_batch primitives expose single stages for parity tests.
 */
#include "as_common.cuh"
#include <stdlib.h>
#include <time.h>

static double now_us(void) {
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return 1e6 * (double)ts.tv_sec + 1e-3 * (double)ts.tv_nsec;
}

extern "C" void as_params_default(as_params *p) {
    if (!p) return;
    p->seed_len = 20;      /* alnmain.c:577-580 */
    p->thres = 300;
    p->gact_T = 320;       /* Darwin GACT operating point (README.md:5; DESIGN.md "GACT spec") */
    p->gact_O = 120;
    p->flags = 0;
    p->gact_band = 0;      /* adaptive banded first pass */
    p->vote_ectx = -1;
    p->gact_order = -1;
    p->gact_blocks_per_sm = 0;
    p->chunk_reads = 0;
    p->chunk_tail = -1;
    p->vote_thread = -1;
    p->gact_runs = -1;
    p->vote_blocks_per_sm = 0;
}

__global__ void k_rle_ops(uint8_t *__restrict__ ops, size_t stride, as_result *__restrict__ res, size_t n);
__global__ void k_rle_ops_warp(uint8_t *__restrict__ ops, size_t stride, as_result *__restrict__ res, size_t n);

static size_t pack_wstride(uint32_t max_len) { return ((size_t)max_len + 15) / 16 + 3; }

/* all device-side stages of n reads, asynchronous on h->stream; ev[0..4] bracket the stages.  `slot0` is the index
 * of the first read inside the handle's per-read scratch (qpack / vote / toff), so chunks of one batch do not overlap. */
__global__ void k_restride(const uint32_t *__restrict__ in, size_t win, uint32_t *__restrict__ lens, size_t n,
                           uint32_t *__restrict__ out, size_t ws, unsigned long long *__restrict__ bad);

static int run_stages(as_gpu *h, const as_params *p, char *d_reads, size_t stride, const uint32_t *d_lens, size_t n,
                      uint32_t max_len, as_result *d_res, uint8_t *d_store, size_t store_stride, bool stop_after_vote,
                      VoteOut **vote_out, cudaEvent_t *ev, size_t slot0, size_t slots,
                      const uint32_t *d_packed_in = nullptr, size_t wstride_in = 0) {
    const size_t ws = pack_wstride(max_len);
    AS_TRY(as_scratch(h, &h->qpack, slots * ws * sizeof(uint32_t)));
    AS_TRY(as_scratch(h, &h->vote, slots * sizeof(VoteOut)));
    AS_TRY(as_scratch(h, &h->toff, slots * sizeof(uint64_t)));
    uint32_t *d_pack = (uint32_t *)h->qpack.p + slot0 * ws;
    VoteOut *d_vote = (VoteOut *)h->vote.p + slot0;
    uint64_t *d_toff = (uint64_t *)h->toff.p + slot0;
    if (vote_out) *vote_out = d_vote;
    AS_CUDA(cudaEventRecord(ev[0], h->stream));
    if (d_packed_in) {                       /* reads arrived 2-bit packed (as_map_batch_packed): only the row stride changes */
        const uint64_t total = (uint64_t)n * ws, blocks = (total + 255) / 256, cap = (uint64_t)h->sm_count * 32;
        k_restride<<<(unsigned)(blocks < cap ? blocks : cap), 256, 0, h->stream>>>(d_packed_in, wstride_in, const_cast<uint32_t *>(d_lens), n,
                                                                                   d_pack, ws, h->d_stats + 15);
        h->launches++;
        AS_CUDA(cudaGetLastError());
    } else {
        AS_TRY(as_launch_pack(h, d_reads, stride, d_lens, n, d_pack, ws));
    }
    AS_TRY(as_launch_vote(h, p, d_pack, ws, d_lens, n, max_len, d_vote));
    AS_CUDA(cudaEventRecord(ev[1], h->stream));
    if (stop_after_vote) return AS_OK;
    AS_TRY(as_launch_locus(h, d_reads, stride, d_lens, n, d_pack, ws, d_vote, d_res, d_toff));
    AS_CUDA(cudaEventRecord(ev[2], h->stream));
    /* columns past n_cigar_op stay zero in DEVICE storage, so equal inputs give equal bytes whatever ran before
     * (the reference leaves them uninitialised, alnmain.c:316-320) */
    /* packed path (the caller reads nothing but the run bytes): one-tile reads leave run bytes straight from GACT's traceback
     * events, every other path writes all the columns it reports -- nothing reads a byte it did not write, no clearing */
    const bool emit_runs = d_packed_in && (p->flags & AS_F_RLE_CIGAR) && d_store && !(p->flags & AS_F_NO_ALIGN) && p->gact_runs != 0;
    if (!(p->flags & AS_F_NO_ALIGN) && d_store && !emit_runs) AS_CUDA(cudaMemsetAsync(d_store, 0, n * store_stride, h->stream));
    h->gact_emit_runs = emit_runs;
    if (!(p->flags & AS_F_NO_ALIGN))
        AS_TRY(as_launch_gact(h, p, d_pack, ws, d_lens, h->ix.text2, d_toff, nullptr, n, max_len, max_len, d_store,
                              store_stride, nullptr, nullptr, d_res));
    if ((p->flags & AS_F_RLE_CIGAR) && !(p->flags & AS_F_NO_ALIGN) && d_store) {
        if (store_stride > 1024) {
            uint64_t blocks = (n + 7) / 8, cap = (uint64_t)h->sm_count * 8;
            k_rle_ops_warp<<<(unsigned)(blocks < cap ? blocks : cap), 256, 0, h->stream>>>(d_store, store_stride, d_res, n);
        } else {
            uint64_t blocks = (n + 255) / 256, cap = (uint64_t)h->sm_count * 8;
            k_rle_ops<<<(unsigned)(blocks < cap ? blocks : cap), 256, 0, h->stream>>>(d_store, store_stride, d_res, n);
        }
        h->launches++;
        AS_CUDA(cudaGetLastError());
    }
    AS_CUDA(cudaEventRecord(ev[3], h->stream));
    if ((p->flags & AS_F_MYERS) && !(p->flags & AS_F_NO_ALIGN))
        AS_TRY(as_launch_myers(h, d_pack, ws, d_lens, h->ix.text2, d_toff, nullptr, n, max_len, nullptr, nullptr, d_res));
    AS_CUDA(cudaEventRecord(ev[4], h->stream));
    return AS_OK;
}

/*
 * AS_F_RLE_CIGAR: the column ops of every read become run bytes in place -- (len << 2) | class, class 0 M ('=' and 'X',
 * the STANDARD alphabet of edlibAlignmentToCigar, edlib.cpp:262-311), 1 I, 2 D, len 1..63 (longer runs continue in the
 * next byte) -- and n_cigar_op becomes the number of run bytes.  It is what parse_cigar (alnmain.c:497-498) prints, one
 * pass earlier: a 150-bp read at 2 % error needs about ten bytes instead of 150, so the D2H copy shrinks by as much.
 */
__global__ void __launch_bounds__(256) k_rle_ops(uint8_t *__restrict__ ops, size_t stride, as_result *__restrict__ res, size_t n) {
    /* one thread per read: a sequential scan of its columns, four per 32-bit load; the run bytes go back into the same
     * row, always behind the read position (runs never outnumber columns) */
    const bool words = (stride & 3) == 0 && (reinterpret_cast<uintptr_t>(ops) & 3) == 0;
    for (size_t r = (size_t)blockIdx.x * blockDim.x + threadIdx.x; r < n; r += (size_t)gridDim.x * blockDim.x) {
        const int nops = res[r].n_cigar_op;
        if (nops <= 0) continue;
        if (nops & AS_NOPS_RUNS) {                 /* GACT wrote the run bytes itself (GactArgs.emit_runs) */
            res[r].n_cigar_op = nops & ~AS_NOPS_RUNS;
            continue;
        }
        uint8_t *row = ops + r * stride;
        int out = 0, open_cls = -1, open_len = 0;
        for (int base = 0; base < nops; base += 4) {
            uint32_t x;
            if (words) x = *reinterpret_cast<const uint32_t *>(row + base);
            else {
                x = 0;
                for (int t = 0; t < 4 && base + t < nops; t++) x |= (uint32_t)row[base + t] << (8 * t);
            }
            const int lim = nops - base < 4 ? nops - base : 4;
#pragma unroll
            for (int t = 0; t < 4; t++) {
                if (t < lim) {
                    const uint32_t c = (x >> (8 * t)) & 0xFFu;
                    const int cls = c == 1u ? 1 : (c == 2u ? 2 : 0);       /* '=' and 'X' are M */
                    if (cls != open_cls) {
                        while (open_len > 0) {
                            const int l = open_len > 63 ? 63 : open_len;
                            row[out++] = (uint8_t)((l << 2) | open_cls);
                            open_len -= l;
                        }
                        open_cls = cls;
                    }
                    open_len++;
                }
            }
        }
        while (open_len > 0) {
            const int l = open_len > 63 ? 63 : open_len;
            row[out++] = (uint8_t)((l << 2) | open_cls);
            open_len -= l;
        }
        res[r].n_cigar_op = out;
    }
}

/* the same with one warp per read (ballot of class changes per 32 columns): long reads, where a thread per read would
 * leave most of the GPU idle */
__global__ void __launch_bounds__(256) k_rle_ops_warp(uint8_t *__restrict__ ops, size_t stride, as_result *__restrict__ res, size_t n) {
    const uint32_t lane = threadIdx.x & 31u;
    const size_t warp0 = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = ((size_t)gridDim.x * blockDim.x) >> 5;
    for (size_t r = warp0; r < n; r += nwarps) {
        const int nops = res[r].n_cigar_op;
        if (nops <= 0) continue;
        if (nops & AS_NOPS_RUNS) {
            if (lane == 0) res[r].n_cigar_op = nops & ~AS_NOPS_RUNS;
            continue;
        }
        uint8_t *row = ops + r * stride;
        int out = 0, open_cls = -1, open_len = 0;
        for (int base = 0; base < nops; base += 32) {
            const bool valid = base + (int)lane < nops;
            const uint32_t c = valid ? row[base + lane] : 0u;
            const int cls = !valid ? 3 : (c == 1u ? 1 : (c == 2u ? 2 : 0));
            int prev = __shfl_up_sync(0xFFFFFFFFu, cls, 1);
            if (lane == 0) prev = open_cls;
            uint32_t heads = __ballot_sync(0xFFFFFFFFu, valid && cls != prev);
            const int nvalid = nops - base < 32 ? nops - base : 32;
            __syncwarp();
            int pos = 0;                                   /* columns of the strip already attributed to a run */
            while (pos < nvalid) {
                /* the run that covers column `pos` reaches to the next head after pos (or the strip's end) */
                const uint32_t later = heads & ~((2u << pos) - 1u);          /* heads at columns > pos */
                const int end = later ? __ffs((int)later) - 1 : nvalid;
                const int c0 = __shfl_sync(0xFFFFFFFFu, cls, pos);
                if ((heads >> pos) & 1u) {                 /* a new run opens here: close the one that was open */
                    while (open_len > 0) {
                        const int l = open_len > 63 ? 63 : open_len;
                        if (lane == 0) row[out] = (uint8_t)((l << 2) | open_cls);
                        out++;
                        open_len -= l;
                    }
                    open_cls = c0;
                }
                open_len += end - pos;
                pos = end;
            }
        }
        while (open_len > 0) {
            const int l = open_len > 63 ? 63 : open_len;
            if (lane == 0) row[out] = (uint8_t)((l << 2) | open_cls);
            out++;
            open_len -= l;
        }
        if (lane == 0) res[r].n_cigar_op = out;
    }
}

/* per-batch counters that are cheap to take from the results on the device */
__global__ void __launch_bounds__(256) k_count_results(const as_result *__restrict__ res, size_t n,
                                                       unsigned long long *__restrict__ stats, int *__restrict__ max_ops) {
    unsigned long long valid = 0, f4 = 0, f16 = 0, sed = 0;
    int mx = 0;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const as_result r = res[i];
        valid += r.valid != 0;
        f4 += (r.flag & 4) != 0;
        f16 += (r.flag & 16) != 0;
        if (r.valid && r.ed > 0) sed += (unsigned long long)r.ed;
        mx = max(mx, r.n_cigar_op);
    }
#pragma unroll
    for (int d = 16; d >= 1; d >>= 1) {
        valid += __shfl_xor_sync(0xFFFFFFFFu, valid, d);
        f4 += __shfl_xor_sync(0xFFFFFFFFu, f4, d);
        f16 += __shfl_xor_sync(0xFFFFFFFFu, f16, d);
        sed += __shfl_xor_sync(0xFFFFFFFFu, sed, d);
        mx = max(mx, __shfl_xor_sync(0xFFFFFFFFu, mx, d));
    }
    if ((threadIdx.x & 31) == 0) {
        if (max_ops && mx) atomicMax(max_ops, mx);
        if (valid) atomicAdd(stats + 1, valid);
        if (f4) atomicAdd(stats + 2, f4);
        if (f16) atomicAdd(stats + 3, f16);
        if (sed) atomicAdd(stats + 4, sed);
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(stats + 0, (unsigned long long)n);
}

static int count_results(as_gpu *h, const as_result *d_res, size_t n, int *d_max_ops) {
    uint64_t blocks = (n + 255) / 256, cap = (uint64_t)h->sm_count * 4;
    k_count_results<<<(unsigned)(blocks < cap ? blocks : cap), 256, 0, h->stream>>>(d_res, n, h->d_stats, d_max_ops);
    h->launches++;
    AS_CUDA(cudaGetLastError());
    return AS_OK;
}

static int check_map_args(as_gpu *h, const as_params *p, const void *reads, size_t stride, const uint32_t *lens,
                          const as_result *results) {
    if (!h || !p || !reads || !lens || !results || stride == 0) {
        as_set_error("as_map_batch: NULL argument");
        return AS_EINVAL;
    }
    if (!h->ix.lc || !h->ix.text2 || !(h->ix.sa_lo || h->d_csa_lo)) {
        as_set_error("as_map_batch: lookup-only handle (opened without lc table, suffix array or text)");
        return AS_EINVAL;
    }
    return AS_OK;
}

static float ev_ms(cudaEvent_t a, cudaEvent_t b) {
    float v = 0.f;
    if (cudaEventElapsedTime(&v, a, b) != cudaSuccess) {
        cudaGetLastError();
        return 0.f;
    }
    return v;
}

/* ms: [0] PART 1, [1] PART 2, [2] GACT, [3] Myers, [4] unused -- kernel stage times summed over the chunks of the call;
 * [5] span of the H2D copies, [6] span of the D2H copies (both overlap the kernels when the call was chunked), [7] whole call */
static int finish_timing(as_gpu *h) {
    for (int i = 0; i < 8; i++) h->last_ms[i] = 0.f;
    const int nc = h->n_chunks_last;
    for (int c = 0; c < nc; c++) {
        cudaEvent_t *e = h->cev[c];
        h->last_ms[0] += ev_ms(e[0], e[1]);
        h->last_ms[1] += ev_ms(e[1], e[2]);
        h->last_ms[2] += ev_ms(e[2], e[3]);
        h->last_ms[3] += ev_ms(e[3], e[4]);
    }
    h->last_ms[7] = ev_ms(h->ev[0], h->ev[6]);
    h->timing_pending = false;
    return AS_OK;
}

static int ensure_pipeline(as_gpu *h) {
    if (h->cev_ready) return AS_OK;
    AS_CUDA(cudaStreamCreateWithFlags(&h->s_h2d, cudaStreamNonBlocking));
    AS_CUDA(cudaStreamCreateWithFlags(&h->s_d2h, cudaStreamNonBlocking));
    for (auto &row : h->cev)
        for (cudaEvent_t &e : row) AS_CUDA(cudaEventCreate(&e));
    AS_CUDA(cudaEventCreate(&h->ev_d2h0));
    AS_CUDA(cudaMalloc((void **)&h->d_chunk_max, AS_MAX_CHUNKS * sizeof(int)));
    AS_CUDA(cudaMallocHost((void **)&h->h_chunk_max, AS_MAX_CHUNKS * sizeof(int)));
    h->cev_ready = true;
    return AS_OK;
}

/* Chunk schedule of one as_map_batch call.  The D2H stream is the longest of the three (results + CIGAR columns + the
 * reverse-complemented reads), so the first chunk is small -- it reaches the D2H stream early -- and the following
 * ones grow to the size at which the kernels run at full efficiency; as_params.chunk_reads forces a fixed size. */
static int plan_chunks(const as_params *p, size_t n, size_t stride, size_t *begin /* AS_MAX_CHUNKS + 1 */, bool light_d2h) {
    size_t fixed = p->chunk_reads;
    int nc = 0;
    begin[0] = 0;
    if (fixed) {
        if ((n + fixed - 1) / fixed > AS_MAX_CHUNKS) fixed = (n + AS_MAX_CHUNKS - 1) / AS_MAX_CHUNKS;
        for (size_t r = 0; r < n; r += fixed) begin[++nc] = (r + fixed < n) ? r + fixed : n;
        return nc;
    }
    const size_t min_reads = 32768;
    /* long reads: cap a chunk at ~256 MB of sequence */
    /* with run-byte CIGARs and no read copy-back the D2H stream is short: fewer, larger chunks keep the kernels efficient */
    size_t big = light_d2h ? n / 3 : n / 4, by_bytes = ((size_t)256 << 20) / (stride ? stride : 1);
    if (big > by_bytes) big = by_bytes;
    if (big < min_reads) big = min_reads;
    size_t c = light_d2h ? big / 3 : big / 4;                  /* ramp: big/4, big/2, big, big, ... */
    if (c < min_reads) c = min_reads;
    size_t r = 0;
    while (r < n) {
        size_t take = c;
        if (nc == AS_MAX_CHUNKS - 1 || n - r < take + min_reads / 2) take = n - r;
        r += take;
        begin[++nc] = r;
        if (c < big) c = (2 * c < big) ? 2 * c : big;
    }
    /* ramp down as well: nothing overlaps the last chunk's D2H, so the last chunk is the smaller part of a split */
    const int tail = p->chunk_tail >= 0 ? p->chunk_tail : 1;
    for (int t = 0; t < tail && nc < AS_MAX_CHUNKS; t++) {
        const size_t b0 = begin[nc - 1], sz = begin[nc] - b0;
        if (sz < 4 * min_reads) break;
        begin[nc] = b0 + (sz * 2 / 3 + 31) / 32 * 32;
        begin[++nc] = n;
    }
    return nc;
}

static int map_batch_impl(as_gpu *h, const as_params *p, char *reads, size_t stride, const uint32_t *lens, size_t n,
                          as_result *results, uint8_t *store_mem, size_t store_stride);

extern "C" int as_map_batch(as_gpu *h, const as_params *p, char *reads, size_t stride, const uint32_t *lens, size_t n,
                            as_result *results, uint8_t *store_mem, size_t store_stride) {
    if (h) h->reads_done_last = h->reads_touched_last = 0;
    AS_TRY(check_map_args(h, p, reads, stride, lens, results));
    if (n == 0) return AS_OK;
    AS_CUDA(cudaSetDevice(h->device));
    AS_TRY(ensure_pipeline(h));
    const int rc = map_batch_impl(h, p, reads, stride, lens, n, results, store_mem, store_stride);
    if (rc != AS_OK) {           /* copies may be in flight on the caller's buffers: let them end before the caller sees the error */
        cudaStreamSynchronize(h->s_h2d);
        cudaStreamSynchronize(h->stream);
        cudaStreamSynchronize(h->s_d2h);
        cudaGetLastError();
    }
    return rc;
}

static int map_batch_impl(as_gpu *h, const as_params *p, char *reads, size_t stride, const uint32_t *lens, size_t n,
                          as_result *results, uint8_t *store_mem, size_t store_stride) {
    AS_TRY(as_scratch(h, &h->reads, n * stride));
    AS_TRY(as_scratch(h, &h->lens, n * sizeof(uint32_t)));
    size_t cb[AS_MAX_CHUNKS + 1];
    const int nc = plan_chunks(p, n, stride, cb, (p->flags & AS_F_KEEP_READS) && (p->flags & AS_F_RLE_CIGAR));
    h->n_chunks_last = nc;
    AS_CUDA(cudaEventRecord(h->ev[0], h->stream));
    AS_CUDA(cudaStreamWaitEvent(h->s_h2d, h->ev[0], 0));
    AS_CUDA(cudaStreamWaitEvent(h->s_d2h, h->ev[0], 0));
    /* every H2D copy is queued up front: they stream in order while the kernels of earlier chunks run -- and while the
     * host looks at the lengths below */
    for (int c = 0; c < nc; c++) {
        const size_t r0 = cb[c], cn = cb[c + 1] - cb[c];
        AS_CUDA(cudaMemcpyAsync((char *)h->reads.p + r0 * stride, reads + r0 * stride, cn * stride, cudaMemcpyHostToDevice, h->s_h2d));
        AS_CUDA(cudaMemcpyAsync((uint32_t *)h->lens.p + r0, lens + r0, cn * sizeof(uint32_t), cudaMemcpyHostToDevice, h->s_h2d));
        AS_CUDA(cudaEventRecord(h->cev[c][5], h->s_h2d));
    }
    uint32_t max_len = 0;
    {
        uint32_t m0 = 0, m1 = 0, m2 = 0, m3 = 0;
        size_t i = 0;
        for (; i + 4 <= n; i += 4) {
            m0 = lens[i] > m0 ? lens[i] : m0;
            m1 = lens[i + 1] > m1 ? lens[i + 1] : m1;
            m2 = lens[i + 2] > m2 ? lens[i + 2] : m2;
            m3 = lens[i + 3] > m3 ? lens[i + 3] : m3;
        }
        for (; i < n; i++) m0 = lens[i] > m0 ? lens[i] : m0;
        m0 = m0 > m1 ? m0 : m1;
        m2 = m2 > m3 ? m2 : m3;
        max_len = m0 > m2 ? m0 : m2;
    }
    const bool want_ops = store_mem != nullptr && !(p->flags & AS_F_NO_ALIGN);
    if ((size_t)max_len + 1 > stride || (want_ops && store_stride < 2 * (size_t)max_len)) {
        if ((size_t)max_len + 1 > stride)
            as_set_error("as_map_batch: a read of %u bases does not fit stride %zu", max_len, stride);
        else
            as_set_error("as_map_batch: store_stride %zu < 2*max_read_len (alnmain.c:316-320)", store_stride);
        return AS_EINVAL;
    }
    AS_TRY(as_scratch(h, &h->results, n * sizeof(as_result)));
    if (want_ops) AS_TRY(as_scratch(h, &h->store, n * store_stride));
    char *d_reads = (char *)h->reads.p;
    uint32_t *d_lens = (uint32_t *)h->lens.p;
    as_result *d_res = (as_result *)h->results.p;
    uint8_t *d_store = want_ops ? (uint8_t *)h->store.p : nullptr;

    /* per-read scratch of every chunk is carved out of one allocation made before anything is enqueued */
    {
        VoteOut *unused;
        (void)unused;
        const size_t ws = pack_wstride(max_len);
        AS_TRY(as_scratch(h, &h->qpack, n * ws * sizeof(uint32_t)));
        AS_TRY(as_scratch(h, &h->vote, n * sizeof(VoteOut)));
        AS_TRY(as_scratch(h, &h->toff, n * sizeof(uint64_t)));
        if (!(p->flags & AS_F_NO_ALIGN) && max_len > (uint32_t)p->gact_T) AS_TRY(as_gact_order_reserve(h, n));
    }
    h->last_h2d_bytes = n * stride + n * sizeof(uint32_t);
    const bool copy_reads_back = !(p->flags & AS_F_KEEP_READS);
    h->last_d2h_bytes = n * sizeof(as_result) + (copy_reads_back ? n * stride : 0) + (size_t)nc * sizeof(int);
    AS_CUDA(cudaMemsetAsync(h->d_chunk_max, 0, AS_MAX_CHUNKS * sizeof(int), h->stream));
    /* the CIGAR columns of chunk c are copied once the chunk's widest row is known: only that many columns of
     * every row cross PCIe (the reference leaves columns >= n_cigar_op uninitialised, alnmain.c:316-320) */
    auto finish_chunk = [&](int c) -> int {
        const size_t r0 = cb[c], cn = cb[c + 1] - cb[c];
        AS_CUDA(cudaEventSynchronize(h->cev[c][7]));
        size_t w = (size_t)(h->h_chunk_max[c] > 0 ? h->h_chunk_max[c] : 0);
        w = (w + 15) & ~(size_t)15;
        if (w > store_stride) w = store_stride;
        h->last_d2h_bytes += w * cn;
        if (w)
            AS_CUDA(cudaMemcpy2DAsync(store_mem + r0 * store_stride, store_stride, d_store + r0 * store_stride, store_stride,
                                      w, cn, cudaMemcpyDeviceToHost, h->s_d2h));
        return AS_OK;
    };
    for (int c = 0; c < nc; c++) {
        const size_t r0 = cb[c], cn = cb[c + 1] - cb[c];
        AS_CUDA(cudaStreamWaitEvent(h->stream, h->cev[c][5], 0));
        AS_TRY(run_stages(h, p, d_reads + r0 * stride, stride, d_lens + r0, cn, max_len, d_res + r0,
                          d_store ? d_store + r0 * store_stride : nullptr, store_stride, false, nullptr, h->cev[c], r0, n));
        AS_TRY(count_results(h, d_res + r0, cn, h->d_chunk_max + c));
        AS_CUDA(cudaEventRecord(h->cev[c][6], h->stream));
        if (want_ops && c > 0) {
            AS_TRY(finish_chunk(c - 1));
            h->reads_done_last = cb[c];                      /* chunk c-1: results, reads and columns are all on their way */
        }
        AS_CUDA(cudaStreamWaitEvent(h->s_d2h, h->cev[c][6], 0));
        if (c == 0) AS_CUDA(cudaEventRecord(h->ev_d2h0, h->s_d2h));
        AS_CUDA(cudaMemcpyAsync(h->h_chunk_max + c, h->d_chunk_max + c, sizeof(int), cudaMemcpyDeviceToHost, h->s_d2h));
        AS_CUDA(cudaEventRecord(h->cev[c][7], h->s_d2h));
        h->reads_touched_last = cb[c + 1];
        AS_CUDA(cudaMemcpyAsync(results + r0, d_res + r0, cn * sizeof(as_result), cudaMemcpyDeviceToHost, h->s_d2h));
        /* strand-1 reads were reverse-complemented in place (alnmain.c:433-438; the OpenACC copy(reads_mem) :421) */
        if (copy_reads_back)
            AS_CUDA(cudaMemcpyAsync(reads + r0 * stride, d_reads + r0 * stride, cn * stride, cudaMemcpyDeviceToHost, h->s_d2h));
        if (!want_ops) h->reads_done_last = cb[c + 1];
    }
    if (want_ops) AS_TRY(finish_chunk(nc - 1));
    h->reads_done_last = n;
    AS_CUDA(cudaEventRecord(h->ev[7], h->s_d2h));
    AS_CUDA(cudaStreamWaitEvent(h->stream, h->ev[7], 0));
    AS_CUDA(cudaEventRecord(h->ev[6], h->stream));
    AS_CUDA(cudaStreamSynchronize(h->stream));
    finish_timing(h);
    h->last_ms[5] = ev_ms(h->ev[0], h->cev[nc - 1][5]);
    h->last_ms[6] = ev_ms(h->ev_d2h0, h->ev[7]);
    h->last_ms[7] = ev_ms(h->ev[0], h->ev[6]);
    return AS_OK;
}

/* ---------------------------------------------------------------------------------------------- */
/* packed reads in, compact records + dense run bytes out                                          */
/* ---------------------------------------------------------------------------------------------- */
/* host rows of `win` words -> device rows of `ws` words (3 words of slack for the funnel-shift readers); bits past the
 * read's length are cleared whatever the host left there */
__global__ void __launch_bounds__(256) k_restride(const uint32_t *__restrict__ in, size_t win, uint32_t *__restrict__ lens,
                                                  size_t n, uint32_t *__restrict__ out, size_t ws,
                                                  unsigned long long *__restrict__ bad) {
    const uint64_t total = (uint64_t)n * ws;
    for (uint64_t g = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; g < total; g += (uint64_t)gridDim.x * blockDim.x) {
        const uint64_t r = g / ws;
        const uint32_t w = (uint32_t)(g - r * ws), b0 = w << 4;
        uint32_t len = lens[r];
        if ((uint64_t)len > 16 * (uint64_t)win) {             /* does not fit its row: mapped as the row's bases, call fails */
            len = (uint32_t)(16 * win);
            if (w == ws - 1) {                                 /* one thread per row stores the clamped length; the others clamp
                                                                * whichever of the two values they happen to read the same way */
                lens[r] = len;
                atomicAdd(bad, 1ULL);
            }
        }
        uint32_t v = 0;
        if (b0 < len && w < win) {
            v = in[r * win + w];
            if (len - b0 < 16) v &= (1u << (2 * (len - b0))) - 1u;
        }
        out[g] = v;
    }
}

#define CR_READS 256         /* reads per block of the compaction kernels */
/* pass 1: the compact record of every read, and the run bytes of each block of CR_READS reads */
__global__ void __launch_bounds__(CR_READS) k_compact_records(const as_result *__restrict__ res, size_t n,
                                                              as_result_compact *__restrict__ out,
                                                              unsigned long long *__restrict__ block_sum) {
    __shared__ unsigned long long s_w[CR_READS / 32];
    const size_t i = (size_t)blockIdx.x * CR_READS + threadIdx.x;
    unsigned long long mine = 0;
    if (i < n) {
        const as_result r = res[i];
        as_result_compact c;
        const uint32_t nr = r.n_cigar_op > 0 && r.seq_id >= 0 ? (uint32_t)r.n_cigar_op : 0u;
        c.pos = (r.off & 0xFFFFFFFFFFULL) | ((uint64_t)(uint16_t)(r.seq_id + 1) << 40) | ((uint64_t)(uint8_t)r.flag << 56);
        c.ed = r.ed;
        c.n_runs = nr | ((((uint32_t)r.flag >> 8) & 0xFu) << 27) | (r.strand ? 0x80000000u : 0u);
        out[i] = c;
        mine = nr;
    }
#pragma unroll
    for (int d = 16; d >= 1; d >>= 1) mine += __shfl_xor_sync(0xFFFFFFFFu, mine, d);
    if ((threadIdx.x & 31) == 0) s_w[threadIdx.x >> 5] = mine;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long t = 0;
        for (int w = 0; w < CR_READS / 32; w++) t += s_w[w];
        block_sum[blockIdx.x] = t;
    }
}

/* pass 2: exclusive scan of the block sums by one block, offset by the run bytes of the chunks before (*base, advanced
 * here); the chunk's total lands in *total */
__global__ void __launch_bounds__(1024) k_scan_block_sums(unsigned long long *__restrict__ block_sum, size_t nb,
                                                          unsigned long long *__restrict__ base, unsigned long long *__restrict__ total) {
    __shared__ unsigned long long s_w[32];
    __shared__ unsigned long long s_carry;
    const unsigned long long base0 = *base;                   /* run bytes of the chunks before this one */
    __syncthreads();
    if (threadIdx.x == 0) s_carry = base0;
    __syncthreads();
    for (size_t b0 = 0; b0 < nb; b0 += 1024) {
        const size_t i = b0 + threadIdx.x;
        const unsigned long long v = i < nb ? block_sum[i] : 0;
        unsigned long long inc = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const unsigned long long o = __shfl_up_sync(0xFFFFFFFFu, inc, d);
            if ((threadIdx.x & 31) >= d) inc += o;
        }
        if ((threadIdx.x & 31) == 31) s_w[threadIdx.x >> 5] = inc;
        __syncthreads();
        if (threadIdx.x < 32) {
            unsigned long long w = s_w[threadIdx.x], wi = w;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const unsigned long long o = __shfl_up_sync(0xFFFFFFFFu, wi, d);
                if (threadIdx.x >= d) wi += o;
            }
            s_w[threadIdx.x] = wi - w;                       /* exclusive over the warps */
        }
        __syncthreads();
        const unsigned long long excl = s_carry + s_w[threadIdx.x >> 5] + inc - v;
        if (i < nb) block_sum[i] = excl;
        __syncthreads();
        if (threadIdx.x == 1023) s_carry = excl + v;
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        *total = s_carry - base0;
        *base = s_carry;
    }
}

/* pass 3: every warp places the run bytes of its 32 reads behind one another in the dense buffer */
__global__ void __launch_bounds__(CR_READS) k_gather_runs(const as_result_compact *__restrict__ rec, size_t n,
                                                          const uint8_t *__restrict__ rows, size_t stride,
                                                          const unsigned long long *__restrict__ block_off,
                                                          uint8_t *__restrict__ dense, unsigned long long cap) {
    __shared__ unsigned long long s_w[CR_READS / 32];
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const size_t i = (size_t)blockIdx.x * CR_READS + threadIdx.x;
    const uint32_t nr = i < n ? AS_RC_NRUNS(rec[i]) : 0u;
    unsigned long long inc = nr;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const unsigned long long o = __shfl_up_sync(0xFFFFFFFFu, inc, d);
        if (lane >= (uint32_t)d) inc += o;
    }
    if (lane == 31) s_w[warp] = inc;
    __syncthreads();
    unsigned long long base = block_off[blockIdx.x];
    for (uint32_t w = 0; w < warp; w++) base += s_w[w];
    const unsigned long long my_off = base + inc - nr;
    for (int src = 0; src < 32; src++) {
        const uint32_t cnt = __shfl_sync(0xFFFFFFFFu, nr, src);
        if (cnt == 0) continue;
        const unsigned long long off = __shfl_sync(0xFFFFFFFFu, my_off, src);
        if (off + cnt > cap) continue;                       /* the caller's buffer is too small: reported through the total */
        const uint8_t *s = rows + ((size_t)blockIdx.x * CR_READS + warp * 32 + (size_t)src) * stride;
        for (uint32_t b = lane; b < cnt; b += 32) dense[off + b] = s[b];
    }
}

static int map_batch_packed_impl(as_gpu *h, const as_params *p, const uint32_t *packed, size_t wstride, const uint32_t *lens,
                                 size_t n, as_result_compact *results, uint8_t *runs, size_t runs_cap, uint64_t *runs_len);

extern "C" int as_map_batch_packed(as_gpu *h, const as_params *p, const uint32_t *packed, size_t wstride, const uint32_t *lens,
                                   size_t n, as_result_compact *results, uint8_t *runs, size_t runs_cap, uint64_t *runs_len) {
    if (!h || !p || !packed || !lens || !results || wstride == 0 || (!runs && runs_cap) || !runs_len) {
        as_set_error("as_map_batch_packed: NULL argument");
        return AS_EINVAL;
    }
    if (!h->ix.lc || !h->ix.text2 || !(h->ix.sa_lo || h->d_csa_lo)) {
        as_set_error("as_map_batch_packed: lookup-only handle (opened without lc table, suffix array or text)");
        return AS_EINVAL;
    }
    if (p->flags & (AS_F_MYERS | AS_F_NO_ALIGN)) {
        as_set_error("as_map_batch_packed: AS_F_MYERS / AS_F_NO_ALIGN have no field in the compact record; use as_map_batch");
        return AS_EINVAL;
    }
    *runs_len = 0;
    if (n == 0) return AS_OK;
    AS_CUDA(cudaSetDevice(h->device));
    AS_TRY(ensure_pipeline(h));
    const int rc = map_batch_packed_impl(h, p, packed, wstride, lens, n, results, runs, runs_cap, runs_len);
    if (rc != AS_OK && rc != AS_ERANGE) {
        cudaStreamSynchronize(h->s_h2d);
        cudaStreamSynchronize(h->stream);
        cudaStreamSynchronize(h->s_d2h);
        cudaGetLastError();
    }
    return rc;
}

/* chunk plan of the packed path: copies are a small fraction of the kernel time, so a small first chunk (its H2D is the
 * only exposed one), two large ones, and a small last chunk (its D2H is the only exposed one) */
static int plan_chunks_packed(const as_params *p, size_t n, size_t row_bytes, size_t *begin) {
    int nc = 0;
    begin[0] = 0;
    if (p->chunk_reads) {
        size_t fixed = p->chunk_reads;
        if ((n + fixed - 1) / fixed > AS_MAX_CHUNKS) fixed = (n + AS_MAX_CHUNKS - 1) / AS_MAX_CHUNKS;
        for (size_t r = 0; r < n; r += fixed) begin[++nc] = (r + fixed < n) ? r + fixed : n;
        return nc;
    }
    const size_t min_reads = 32768;
    /* long reads: the copies are nothing next to the kernels, and GACT hands a launch's reads out longest first so that the
     * 32 lanes of a warp finish together -- every extra chunk costs a tail of half-empty warps (c5: 480 ms in 256 MB chunks
     * against 174 ms whole).  So as few chunks as 16 GB of op rows each allow, all the same size. */
    size_t by_bytes = ((size_t)16 << 30) / (row_bytes ? row_bytes : 1);
    if (by_bytes < 1024) by_bytes = 1024;
    if (row_bytes >= 2048 || (n < 4 * min_reads && n <= by_bytes)) {
        const size_t parts = (n + by_bytes - 1) / by_bytes;
        for (size_t k = 1; k <= parts && nc < AS_MAX_CHUNKS; k++) begin[++nc] = (size_t)(((unsigned __int128)n * k) / parts);
        begin[nc] = n;
        return nc;
    }
    /* Chunks alternate between two lanes (streams) and share the GPU, so what matters is that both lanes stay busy: one
     * eighth first -- its H2D is the only exposed one --, one eighth last -- its D2H is the only exposed one --, and three
     * quarters in between as two chunks (more when the op rows of long reads would not fit); as_params.chunk_tail > 0
     * sets the number of chunks in between (A/B runs).  Measured on c2 (1 M reads, kernels alone 3.9 ms): 4.58 ms with
     * two chunks in between or four equal chunks, 4.6 with three, 4.7 with one or four, 5.0 with six. */
    /* (A/B runs: chunk_tail = 100 * d + m asks for edges of n / d reads and m chunks in between) */
    const size_t edge_div = p->chunk_tail >= 100 ? (size_t)(p->chunk_tail / 100) : 8;
    size_t edge = n / (edge_div < 3 ? 3 : edge_div);
    if (edge < min_reads) edge = min_reads;
    if (edge > by_bytes) edge = by_bytes;
    const size_t inner = n - 2 * edge;
    int mids = p->chunk_tail > 0 ? p->chunk_tail % 100 : 2;
    if (mids == 0) mids = 2;
    if ((inner + (size_t)mids - 1) / (size_t)mids > by_bytes) mids = (int)((inner + by_bytes - 1) / by_bytes);
    if (mids < 1) mids = 1;
    if (mids > AS_MAX_CHUNKS - 2) mids = AS_MAX_CHUNKS - 2;
    size_t r = edge;
    begin[++nc] = r;
    for (int k = 0; k < mids; k++) {
        r = edge + (size_t)(((unsigned __int128)inner * (unsigned)(k + 1)) / (unsigned)mids);
        begin[++nc] = r;
    }
    begin[++nc] = n;
    return nc;
}

static int ensure_twin(as_gpu *h) {
    if (h->twin) return AS_OK;
    as_gpu *t = as_handle_new(h->device);
    if (!t) return AS_ENODEV;
    t->ix = h->ix;
    t->opts = h->opts;
    t->is_twin = true;
    h->twin = t;
    return AS_OK;
}

static int map_batch_packed_impl(as_gpu *h, const as_params *p, const uint32_t *packed, size_t wstride, const uint32_t *lens,
                                 size_t n, as_result_compact *results, uint8_t *runs, size_t runs_cap, uint64_t *runs_len) {
    const double t_in = now_us();
    AS_TRY(ensure_twin(h));
    h->twin->ix = h->ix;                                  /* tables attached since (as_gpu_attach_*) */
    size_t cb[AS_MAX_CHUNKS + 1];
    const int nc = plan_chunks_packed(p, n, 32 * wstride, cb);          /* 32 * wstride >= 2 * max_len: the op row of a read */
    h->n_chunks_last = nc;
    size_t max_chunk = 0;
    for (int c = 0; c < nc; c++) max_chunk = cb[c + 1] - cb[c] > max_chunk ? cb[c + 1] - cb[c] : max_chunk;
    /* the reads start crossing PCIe before the host has looked at a single length */
    AS_TRY(as_scratch(h, &h->pack_in, n * wstride * sizeof(uint32_t)));
    AS_TRY(as_scratch(h, &h->lens, n * sizeof(uint32_t)));
    uint32_t *d_in = (uint32_t *)h->pack_in.p;
    uint32_t *d_lens = (uint32_t *)h->lens.p;
    AS_CUDA(cudaEventRecord(h->ev[0], h->stream));
    AS_CUDA(cudaStreamWaitEvent(h->s_h2d, h->ev[0], 0));
    for (int c = 0; c < nc; c++) {
        const size_t r0 = cb[c], cn = cb[c + 1] - cb[c];
        AS_CUDA(cudaMemcpyAsync(d_in + r0 * wstride, packed + r0 * wstride, cn * wstride * sizeof(uint32_t), cudaMemcpyHostToDevice, h->s_h2d));
        AS_CUDA(cudaMemcpyAsync(d_lens + r0, lens + r0, cn * sizeof(uint32_t), cudaMemcpyHostToDevice, h->s_h2d));
        AS_CUDA(cudaEventRecord(h->cev[c][5], h->s_h2d));
    }
    /* The host never reads the lengths (a scan of 1 M of them costs more than the kernels' launch): every buffer and
     * launch shape is sized from the row width, max_len <= 16 * wstride; k_restride clamps a length that does not fit its
     * row and raises a flag, which turns the call into AS_EINVAL. */
    const uint32_t max_len = (uint32_t)(16 * wstride < 0xFFFFFFF0ull ? 16 * wstride : 0xFFFFFFF0ull);
    const size_t ws = pack_wstride(max_len), store_stride = (2 * (size_t)max_len + 15) & ~(size_t)15;
    /* every buffer is sized before anything is enqueued (growing one waits for the stream).  The op rows and the
     * as_result records are chunk-local scratch: the kernels of all chunks share one stream, so a chunk has read them
     * before the next one writes.  The dense run bytes of the whole batch collect in one buffer (worst case: every
     * column its own run), so no chunk ever waits for another one's copy to leave. */
    as_gpu *lane[2] = {h, nc > 1 ? h->twin : h};          /* chunk c runs on lane[c & 1] */
    const size_t nblk_max = (max_chunk + CR_READS - 1) / CR_READS;
    for (int k = 0; k < (nc > 1 ? 2 : 1); k++) {
        as_gpu *g = lane[k];
        AS_TRY(as_scratch(g, &g->qpack, n * ws * sizeof(uint32_t)));
        AS_TRY(as_scratch(g, &g->vote, n * sizeof(VoteOut)));
        AS_TRY(as_scratch(g, &g->toff, n * sizeof(uint64_t)));
        AS_TRY(as_scratch(g, &g->results, max_chunk * sizeof(as_result)));
        AS_TRY(as_scratch(g, &g->store, max_chunk * store_stride));
        AS_TRY(as_scratch(g, &g->run_tot, (AS_MAX_CHUNKS + nblk_max + 8) * sizeof(unsigned long long)));
        if (max_len > (uint32_t)p->gact_T) AS_TRY(as_gact_order_reserve(g, max_chunk));
    }
    AS_TRY(as_scratch(h, &h->runs, n * store_stride));
    AS_TRY(as_scratch(h, &h->cres, n * sizeof(as_result_compact)));
    if (!h->h_run_tot) {
        AS_CUDA(cudaMallocHost((void **)&h->h_run_tot, (AS_MAX_CHUNKS + 2) * sizeof(unsigned long long)));
        AS_CUDA(cudaStreamCreateWithFlags(&h->s_d2h2, cudaStreamNonBlocking));
    }
    uint8_t *d_dense = (uint8_t *)h->runs.p;
    as_result_compact *d_cres = (as_result_compact *)h->cres.p;
    /* run_tot of the first lane: [0] running base of the dense buffer, [1 .. AS_MAX_CHUNKS] total per chunk; the per-block
     * sums behind them are per lane */
    unsigned long long *d_base = (unsigned long long *)h->run_tot.p, *d_tot = d_base + 1;

    AS_CUDA(cudaStreamWaitEvent(h->s_d2h, h->ev[0], 0));
    AS_CUDA(cudaStreamWaitEvent(h->s_d2h2, h->ev[0], 0));
    if (lane[1] != h) AS_CUDA(cudaStreamWaitEvent(lane[1]->stream, h->ev[0], 0));
    h->last_h2d_bytes = n * wstride * sizeof(uint32_t) + n * sizeof(uint32_t);
    h->last_d2h_bytes = n * sizeof(as_result_compact) + (size_t)nc * sizeof(unsigned long long);
    as_params q = *p;
    q.flags |= AS_F_RLE_CIGAR | AS_F_KEEP_READS;
    AS_CUDA(cudaMemsetAsync(d_base, 0, sizeof(unsigned long long), h->stream));
    AS_CUDA(cudaMemsetAsync(h->d_stats + 15, 0, sizeof(unsigned long long), h->stream));       /* reads too long for their row */
    if (lane[1] != h) AS_CUDA(cudaMemsetAsync(lane[1]->d_stats + 15, 0, sizeof(unsigned long long), lane[1]->stream));
    h->dbg_us[2] = now_us() - t_in;
    /* the kernels of every chunk are enqueued first, chunks alternating between the two lanes: the kernels size their grids
     * to fill the GPU, so a chunk's blocks start as the other lane's retire -- the tails overlap, the bulk does not */
    for (int c = 0; c < nc; c++) {
        const size_t r0 = cb[c], cn = cb[c + 1] - cb[c];
        if (c == 1) h->dbg_us[3] = now_us() - t_in;
        as_gpu *g = lane[c & 1];
        as_result *d_res = (as_result *)g->results.p;
        uint8_t *d_rows = (uint8_t *)g->store.p;
        unsigned long long *d_bsum = (unsigned long long *)g->run_tot.p + 1 + AS_MAX_CHUNKS;
        AS_CUDA(cudaStreamWaitEvent(g->stream, h->cev[c][5], 0));
        AS_TRY(run_stages(g, &q, nullptr, 0, d_lens + r0, cn, max_len, d_res, d_rows, store_stride, false, nullptr, h->cev[c], r0, n,
                          d_in + r0 * wstride, wstride));
        AS_TRY(count_results(g, d_res, cn, nullptr));
        const unsigned nblk = (unsigned)((cn + CR_READS - 1) / CR_READS);
        k_compact_records<<<nblk, CR_READS, 0, g->stream>>>(d_res, cn, d_cres + r0, d_bsum);
        if (c > 0) AS_CUDA(cudaStreamWaitEvent(g->stream, h->cev[c - 1][6], 0));     /* the dense offsets continue chunk c-1's */
        k_scan_block_sums<<<1, 1024, 0, g->stream>>>(d_bsum, nblk, d_base, d_tot + c);
        k_gather_runs<<<nblk, CR_READS, 0, g->stream>>>(d_cres + r0, cn, d_rows, store_stride, d_bsum, d_dense,
                                                       (unsigned long long)(n * store_stride));
        g->launches += 3;
        AS_CUDA(cudaGetLastError());
        AS_CUDA(cudaEventRecord(h->cev[c][6], g->stream));
        AS_CUDA(cudaStreamWaitEvent(h->s_d2h, h->cev[c][6], 0));
        if (c == 0) AS_CUDA(cudaEventRecord(h->ev_d2h0, h->s_d2h));
        AS_CUDA(cudaMemcpyAsync(h->h_run_tot + c, d_tot + c, sizeof(unsigned long long), cudaMemcpyDeviceToHost, h->s_d2h));
        AS_CUDA(cudaEventRecord(h->cev[c][7], h->s_d2h));
        AS_CUDA(cudaMemcpyAsync(results + r0, d_cres + r0, cn * sizeof(as_result_compact), cudaMemcpyDeviceToHost, h->s_d2h));
    }
    AS_CUDA(cudaMemcpyAsync(h->h_run_tot + AS_MAX_CHUNKS, h->d_stats + 15, sizeof(unsigned long long), cudaMemcpyDeviceToHost, h->s_d2h));
    AS_CUDA(cudaMemcpyAsync(h->h_run_tot + AS_MAX_CHUNKS + 1, lane[1]->d_stats + 15, sizeof(unsigned long long), cudaMemcpyDeviceToHost, h->s_d2h));
    AS_CUDA(cudaEventRecord(h->ev[7], h->s_d2h));
    const double t_enq = now_us();
    /* ... then the dense run bytes of each chunk leave, on their own stream, as soon as the chunk's total is known here */
    uint64_t host_off = 0;
    int rc_out = AS_OK;
    for (int c = 0; c < nc; c++) {
        AS_CUDA(cudaEventSynchronize(h->cev[c][7]));
        const uint64_t tot = h->h_run_tot[c];
        if (host_off + tot > runs_cap) rc_out = AS_ERANGE;
        else if (tot) AS_CUDA(cudaMemcpyAsync(runs + host_off, d_dense + host_off, tot, cudaMemcpyDeviceToHost, h->s_d2h2));
        host_off += tot;
    }
    h->last_d2h_bytes += host_off;
    AS_CUDA(cudaEventRecord(h->ev[8], h->s_d2h2));
    AS_CUDA(cudaStreamWaitEvent(h->stream, h->ev[7], 0));
    AS_CUDA(cudaStreamWaitEvent(h->stream, h->ev[8], 0));
    AS_CUDA(cudaEventRecord(h->ev[6], h->stream));
    AS_CUDA(cudaStreamSynchronize(h->stream));
    const double t_done = now_us();
    finish_timing(h);
    h->last_host_us[0] = t_enq - t_in;          /* host time until every kernel of the call was enqueued */
    h->last_host_us[1] = t_done - t_enq;        /* waiting for the device */
    h->last_host_us[2] = now_us() - t_done;     /* reading the event times */
    h->last_ms[5] = ev_ms(h->ev[0], h->cev[nc - 1][5]);
    h->last_ms[6] = ev_ms(h->ev_d2h0, h->ev[6]);
    h->last_ms[7] = ev_ms(h->ev[0], h->ev[6]);
    *runs_len = host_off;
    if (h->h_run_tot[AS_MAX_CHUNKS] + h->h_run_tot[AS_MAX_CHUNKS + 1]) {
        as_set_error("as_map_batch_packed: %llu reads are longer than their row of %zu words holds",
                     h->h_run_tot[AS_MAX_CHUNKS] + (lane[1] != h ? h->h_run_tot[AS_MAX_CHUNKS + 1] : 0), wstride);
        return AS_EINVAL;
    }
    if (rc_out == AS_ERANGE) as_set_error("as_map_batch_packed: %llu run bytes do not fit runs_cap %zu", (unsigned long long)host_off, runs_cap);
    return rc_out;
}

extern "C" int as_map_batch_device(as_gpu *h, const as_params *p, char *d_reads, size_t stride, const uint32_t *d_lens,
                                   size_t n, uint32_t max_len, as_result *d_results, uint8_t *d_store, size_t store_stride) {
    AS_TRY(check_map_args(h, p, d_reads, stride, d_lens, d_results));
    if (n == 0) return AS_OK;
    AS_CUDA(cudaSetDevice(h->device));
    AS_TRY(ensure_pipeline(h));
    if ((size_t)max_len + 1 > stride) return AS_EINVAL;
    if (d_store && !(p->flags & AS_F_NO_ALIGN) && store_stride < 2 * (size_t)max_len) return AS_EINVAL;
    h->n_chunks_last = 1;
    AS_CUDA(cudaEventRecord(h->ev[0], h->stream));
    AS_TRY(run_stages(h, p, d_reads, stride, d_lens, n, max_len, d_results, d_store, store_stride, false, nullptr, h->cev[0], 0, n));
    AS_TRY(count_results(h, d_results, n, nullptr));
    AS_CUDA(cudaEventRecord(h->ev[6], h->stream));
    h->timing_pending = true;
    return AS_OK;                     /* asynchronous: as_gpu_sync() or the caller's stream sync completes it */
}

extern "C" int as_gpu_sync(as_gpu *h) {
    if (!h) return AS_EINVAL;
    AS_CUDA(cudaSetDevice(h->device));
    AS_CUDA(cudaStreamSynchronize(h->stream));
    return AS_OK;
}

extern "C" int as_gpu_last_timing(as_gpu *h, float ms[8]) {
    if (!h || !ms) return AS_EINVAL;
    AS_CUDA(cudaSetDevice(h->device));
    AS_CUDA(cudaStreamSynchronize(h->stream));
    if (h->timing_pending) finish_timing(h);
    for (int i = 0; i < 8; i++) ms[i] = h->last_ms[i];
    return AS_OK;
}

/* when the events of chunk `chunk` of the last as_map_batch* call fired, in ms after the call's start: [0] kernels begin,
 * [1] PART 1 done, [2] PART 2 done, [3] GACT done, [4] Myers done, [5] the chunk's H2D done, [6] all kernels done,
 * [7] the chunk's fixed-size D2H done.  Returns the number of chunks of the call. */
extern "C" int as_gpu_last_chunk_timeline(as_gpu *h, int chunk, float ms[8]) {
    if (!h || !ms) return AS_EINVAL;
    if (chunk < 0 || chunk >= h->n_chunks_last) return h->n_chunks_last;
    AS_CUDA(cudaSetDevice(h->device));
    for (int i = 0; i < 8; i++) ms[i] = ev_ms(h->ev[0], h->cev[chunk][i]);
    return h->n_chunks_last;
}

extern "C" int as_gpu_last_host_timing(as_gpu *h, double us[3]) {
    if (!h || !us) return AS_EINVAL;
    for (int i = 0; i < 3; i++) us[i] = h->last_host_us[i];
    if (getenv("AS_DEBUG_HOST_US"))
        fprintf(stderr, "host us: h2d enqueued %.0f, lens scanned %.0f, buffers sized %.0f, chunk 0 enqueued %.0f\n", h->dbg_us[0],
                h->dbg_us[1], h->dbg_us[2], h->dbg_us[3]);
    return AS_OK;
}

extern "C" int as_gpu_last_transfer(as_gpu *h, uint64_t *h2d_bytes, uint64_t *d2h_bytes) {
    if (!h) return AS_EINVAL;
    if (h2d_bytes) *h2d_bytes = h->last_h2d_bytes;
    if (d2h_bytes) *d2h_bytes = h->last_d2h_bytes;
    return AS_OK;
}

extern "C" size_t as_gpu_last_reads_done(as_gpu *h, size_t *touched) {
    if (touched) *touched = h ? h->reads_touched_last : 0;
    return h ? h->reads_done_last : 0;
}

extern "C" int as_gpu_stats(as_gpu *h, as_stats *out) {
    if (!h || !out) return AS_EINVAL;
    AS_CUDA(cudaSetDevice(h->device));
    AS_CUDA(cudaStreamSynchronize(h->stream));
    unsigned long long v[18];
    AS_CUDA(cudaMemcpy(v, h->d_stats, sizeof v, cudaMemcpyDeviceToHost));
    if (h->twin) {
        unsigned long long w[18];
        AS_CUDA(cudaStreamSynchronize(h->twin->stream));
        AS_CUDA(cudaMemcpy(w, h->twin->d_stats, sizeof w, cudaMemcpyDeviceToHost));
        for (int i = 0; i < 18; i++) v[i] += w[i];
    }
    out->total = v[0]; out->valid = v[1]; out->n_flag4 = v[2]; out->n_flag16 = v[3]; out->sum_ed = v[4];
    out->dp_cells = v[5]; out->seeds = v[6]; out->sa_hits = v[7]; out->bs_steps = v[8]; out->overflow_reads = v[9];
    out->dp_cells_eval = v[10]; out->gact_redone = v[11]; out->fm_steps_exec = v[12]; out->text_verified = v[13];
    out->overflow_queued = v[14];
    out->vote_deferred = v[16];         /* v[15]: reads too long for their row (as_map_batch_packed) */
    out->gact_queued = v[17];
    return AS_OK;
}

extern "C" int as_gpu_stats_reset(as_gpu *h) {
    if (!h) return AS_EINVAL;
    AS_CUDA(cudaSetDevice(h->device));
    AS_CUDA(cudaMemsetAsync(h->d_stats, 0, 24 * sizeof(unsigned long long), h->stream));
    if (h->twin) AS_CUDA(cudaMemsetAsync(h->twin->d_stats, 0, 24 * sizeof(unsigned long long), h->twin->stream));
    return AS_OK;
}

extern "C" uint64_t as_gpu_launch_count(as_gpu *h) { return h ? h->launches + (h->twin ? h->twin->launches : 0) : 0; }
extern "C" uint64_t as_gpu_index_length(as_gpu *h) { return h ? h->ix.N : 0; }

/* ---------------------------------------------------------------------------------------------- */
/* PART 1 only                                                                                     */
/* ---------------------------------------------------------------------------------------------- */
extern "C" int as_vote_batch(as_gpu *h, const as_params *p, const char *reads, size_t stride, const uint32_t *lens,
                             size_t n, uint64_t *best_key, uint64_t *best_val, int32_t *phases, int32_t *from_outer) {
    if (!h || !p || !reads || !lens || stride == 0) return AS_EINVAL;
    if (n == 0) return AS_OK;
    if (!h->ix.lc || !h->ix.text2 || !(h->ix.sa_lo || h->d_csa_lo)) {
        as_set_error("as_vote_batch: lookup-only handle (opened without lc table, suffix array or text)");
        return AS_EINVAL;
    }
    AS_CUDA(cudaSetDevice(h->device));
    uint32_t max_len = 0;
    for (size_t i = 0; i < n; i++)
        if (lens[i] > max_len) max_len = lens[i];
    if ((size_t)max_len + 1 > stride) return AS_EINVAL;
    AS_TRY(as_scratch(h, &h->reads, n * stride));
    AS_TRY(as_scratch(h, &h->lens, n * sizeof(uint32_t)));
    AS_CUDA(cudaMemcpyAsync(h->reads.p, reads, n * stride, cudaMemcpyHostToDevice, h->stream));
    AS_CUDA(cudaMemcpyAsync(h->lens.p, lens, n * sizeof(uint32_t), cudaMemcpyHostToDevice, h->stream));
    VoteOut *d_vote = nullptr;
    AS_TRY(ensure_pipeline(h));
    AS_TRY(run_stages(h, p, (char *)h->reads.p, stride, (const uint32_t *)h->lens.p, n, max_len, nullptr, nullptr, 0, true, &d_vote,
                      h->cev[0], 0, n));
    AS_TRY(as_pinned(h, n * sizeof(VoteOut)));
    AS_CUDA(cudaMemcpyAsync(h->h_pin, d_vote, n * sizeof(VoteOut), cudaMemcpyDeviceToHost, h->stream));
    AS_CUDA(cudaStreamSynchronize(h->stream));
    const VoteOut *v = (const VoteOut *)h->h_pin;
    for (size_t i = 0; i < n; i++) {
        if (best_key) best_key[i] = v[i].best_key;
        if (best_val) best_val[i] = v[i].best_val;
        if (phases) phases[i] = v[i].phases;
        if (from_outer) from_outer[i] = v[i].from_outer;
    }
    return AS_OK;
}

/* ---------------------------------------------------------------------------------------------- */
/* aligner on explicit (query, target) pairs                                                       */
/* ---------------------------------------------------------------------------------------------- */
static int upload_pairs(as_gpu *h, const char *q, size_t q_stride, const uint32_t *qlens, const char *t, size_t t_stride,
                        const uint32_t *tlens, size_t n, uint32_t *max_q, uint32_t *max_t, size_t *qws, size_t *tws) {
    uint32_t mq = 0, mt = 0;
    for (size_t i = 0; i < n; i++) {
        if (qlens[i] > mq) mq = qlens[i];
        if (tlens[i] > mt) mt = tlens[i];
    }
    if (mq > q_stride || mt > t_stride) {
        as_set_error("pair batch: a sequence is longer than its stride");
        return AS_EINVAL;
    }
    *max_q = mq;
    *max_t = mt;
    *qws = pack_wstride(mq);
    *tws = pack_wstride(mt);
    AS_TRY(as_scratch(h, &h->misc[0], n * q_stride + 16));
    AS_TRY(as_scratch(h, &h->misc[1], n * t_stride + 16));
    AS_TRY(as_scratch(h, &h->misc[2], 2 * n * sizeof(uint32_t)));
    AS_TRY(as_scratch(h, &h->qpack, n * *qws * sizeof(uint32_t)));
    AS_TRY(as_scratch(h, &h->tpack, n * *tws * sizeof(uint32_t)));
    AS_CUDA(cudaMemcpyAsync(h->misc[0].p, q, n * q_stride, cudaMemcpyHostToDevice, h->stream));
    AS_CUDA(cudaMemcpyAsync(h->misc[1].p, t, n * t_stride, cudaMemcpyHostToDevice, h->stream));
    uint32_t *d_l = (uint32_t *)h->misc[2].p;
    AS_CUDA(cudaMemcpyAsync(d_l, qlens, n * sizeof(uint32_t), cudaMemcpyHostToDevice, h->stream));
    AS_CUDA(cudaMemcpyAsync(d_l + n, tlens, n * sizeof(uint32_t), cudaMemcpyHostToDevice, h->stream));
    AS_TRY(as_launch_pack(h, (const char *)h->misc[0].p, q_stride, d_l, n, (uint32_t *)h->qpack.p, *qws));
    AS_TRY(as_launch_pack(h, (const char *)h->misc[1].p, t_stride, d_l + n, n, (uint32_t *)h->tpack.p, *tws));
    return AS_OK;
}

__global__ void k_pair_offsets(uint64_t *toff, size_t n, size_t bases_per_pair) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        toff[i] = (uint64_t)i * bases_per_pair;
}

static int pair_offsets(as_gpu *h, size_t n, size_t tws) {
    AS_TRY(as_scratch(h, &h->toff, n * sizeof(uint64_t)));
    uint64_t blocks = (n + 255) / 256;
    k_pair_offsets<<<(unsigned)(blocks < 1024 ? blocks : 1024), 256, 0, h->stream>>>((uint64_t *)h->toff.p, n, tws * 16);
    h->launches++;
    AS_CUDA(cudaGetLastError());
    return AS_OK;
}

extern "C" int as_cigar_align_batch(as_gpu *h, const as_params *p, const char *q, size_t q_stride, const uint32_t *qlens,
                                    const char *t, size_t t_stride, const uint32_t *tlens, size_t n, uint8_t *ops,
                                    size_t ops_stride, int32_t *n_ops, int32_t *score) {
    if (!h || !p || !q || !t || !qlens || !tlens || !n_ops || !score) return AS_EINVAL;
    if (n == 0) return AS_OK;
    AS_CUDA(cudaSetDevice(h->device));
    uint32_t mq, mt;
    size_t qws, tws;
    AS_TRY(upload_pairs(h, q, q_stride, qlens, t, t_stride, tlens, n, &mq, &mt, &qws, &tws));
    if (ops && ops_stride < (size_t)mq + mt) {
        as_set_error("as_cigar_align_batch: ops_stride %zu < max qlen + max tlen", ops_stride);
        return AS_EINVAL;
    }
    AS_TRY(pair_offsets(h, n, tws));
    AS_TRY(as_scratch(h, &h->misc[3], 2 * n * sizeof(int32_t)));
    if (ops) AS_TRY(as_scratch(h, &h->store, n * ops_stride));
    if (ops) AS_CUDA(cudaMemsetAsync(h->store.p, 0, n * ops_stride, h->stream));     /* the aligner writes only non-'=' ops of one-tile pairs */
    int32_t *d_no = (int32_t *)h->misc[3].p;
    const uint32_t *d_l = (const uint32_t *)h->misc[2].p;
    AS_TRY(as_launch_gact(h, p, (const uint32_t *)h->qpack.p, qws, d_l, (const uint32_t *)h->tpack.p, (const uint64_t *)h->toff.p,
                          d_l + n, n, mq, mt, ops ? (uint8_t *)h->store.p : nullptr, ops_stride, d_no, d_no + n, nullptr));
    AS_CUDA(cudaMemcpyAsync(n_ops, d_no, n * sizeof(int32_t), cudaMemcpyDeviceToHost, h->stream));
    AS_CUDA(cudaMemcpyAsync(score, d_no + n, n * sizeof(int32_t), cudaMemcpyDeviceToHost, h->stream));
    if (ops) AS_CUDA(cudaMemcpyAsync(ops, h->store.p, n * ops_stride, cudaMemcpyDeviceToHost, h->stream));
    AS_CUDA(cudaStreamSynchronize(h->stream));
    return AS_OK;
}

extern "C" int as_edit_alignment_batch(as_gpu *h, const char *q, size_t q_stride, const uint32_t *qlens, const char *t,
                                       size_t t_stride, const uint32_t *tlens, size_t n, const int32_t *ks, int32_t *dist,
                                       uint8_t *ops, size_t ops_stride, int32_t *n_ops) {
    if (!h || !q || !t || !qlens || !tlens || !dist || !ops || !n_ops) return AS_EINVAL;
    if (n == 0) return AS_OK;
    AS_CUDA(cudaSetDevice(h->device));
    uint32_t mq, mt;
    size_t qws, tws;
    /* limits first (host scan): the query's words stay in registers, and past 1 MiB of alignment data edlib itself leaves
     * the traceback for Hirschberg's divide and conquer (edlib.cpp:1131-1143), whose tie-breaks are not this function's */
    for (size_t i = 0; i < n; i++) {
        const uint64_t nb = ((uint64_t)qlens[i] + 63) / 64;
        if (qlens[i] > AS_PATH_MAX_QLEN || 20ull * nb * tlens[i] + 8ull * tlens[i] >= (1ull << 20)) {
            as_set_error("as_edit_alignment_batch: pair %zu (%u x %u bases) is beyond the traceback's range: query <= %d bases "
                         "and 20 * ceil(qlen / 64) * tlen + 8 * tlen < 2^20", i, qlens[i], tlens[i], AS_PATH_MAX_QLEN);
            return AS_ERANGE;
        }
        if ((size_t)qlens[i] + tlens[i] > ops_stride) {
            as_set_error("as_edit_alignment_batch: ops_stride %zu < qlen + tlen of pair %zu", ops_stride, i);
            return AS_EINVAL;
        }
    }
    AS_TRY(upload_pairs(h, q, q_stride, qlens, t, t_stride, tlens, n, &mq, &mt, &qws, &tws));
    AS_TRY(pair_offsets(h, n, tws));
    AS_TRY(as_scratch(h, &h->misc[3], 3 * n * sizeof(int32_t)));
    AS_TRY(as_scratch(h, &h->store, n * ops_stride));
    int32_t *d_d = (int32_t *)h->misc[3].p, *d_n = d_d + n;
    int32_t *d_k = nullptr;
    if (ks) {
        d_k = d_d + 2 * n;
        AS_CUDA(cudaMemcpyAsync(d_k, ks, n * sizeof(int32_t), cudaMemcpyHostToDevice, h->stream));
    }
    const uint32_t *d_l = (const uint32_t *)h->misc[2].p;
    AS_TRY(as_launch_myers_path(h, (const uint32_t *)h->qpack.p, qws, d_l, (const uint32_t *)h->tpack.p, (const uint64_t *)h->toff.p,
                                d_l + n, n, mq, mt, d_k, d_d, (uint8_t *)h->store.p, ops_stride, d_n));
    AS_CUDA(cudaMemcpyAsync(dist, d_d, n * sizeof(int32_t), cudaMemcpyDeviceToHost, h->stream));
    AS_CUDA(cudaMemcpyAsync(n_ops, d_n, n * sizeof(int32_t), cudaMemcpyDeviceToHost, h->stream));
    AS_CUDA(cudaMemcpyAsync(ops, h->store.p, n * ops_stride, cudaMemcpyDeviceToHost, h->stream));
    AS_CUDA(cudaStreamSynchronize(h->stream));
    return AS_OK;
}

extern "C" int as_edit_distance_batch(as_gpu *h, const char *q, size_t q_stride, const uint32_t *qlens, const char *t,
                                      size_t t_stride, const uint32_t *tlens, size_t n, const int32_t *ks, int32_t *dist) {
    if (!h || !q || !t || !qlens || !tlens || !dist) return AS_EINVAL;
    if (n == 0) return AS_OK;
    AS_CUDA(cudaSetDevice(h->device));
    uint32_t mq, mt;
    size_t qws, tws;
    AS_TRY(upload_pairs(h, q, q_stride, qlens, t, t_stride, tlens, n, &mq, &mt, &qws, &tws));
    AS_TRY(pair_offsets(h, n, tws));
    AS_TRY(as_scratch(h, &h->misc[3], 2 * n * sizeof(int32_t)));
    int32_t *d_d = (int32_t *)h->misc[3].p;
    int32_t *d_k = nullptr;
    if (ks) {
        d_k = d_d + n;
        AS_CUDA(cudaMemcpyAsync(d_k, ks, n * sizeof(int32_t), cudaMemcpyHostToDevice, h->stream));
    }
    const uint32_t *d_l = (const uint32_t *)h->misc[2].p;
    AS_TRY(as_launch_myers(h, (const uint32_t *)h->qpack.p, qws, d_l, (const uint32_t *)h->tpack.p, (const uint64_t *)h->toff.p,
                           d_l + n, n, mq, d_k, d_d, nullptr));
    AS_CUDA(cudaMemcpyAsync(dist, d_d, n * sizeof(int32_t), cudaMemcpyDeviceToHost, h->stream));
    AS_CUDA(cudaStreamSynchronize(h->stream));
    return AS_OK;
}
