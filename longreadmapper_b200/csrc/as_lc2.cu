/*
 * as_lc2.cu -- the deep k-mer table: the SA interval of every h2-mer, h2 chosen so that 4^h2 is about 4x the rows, one
 * level more while the table stays within 4 GB (14 for the 4.6 Mbp genome, 16 for 3.1 Gbp), 8 bytes per entry.
 *
 * The reference's lchash (lchash.c:52-104) answers the last hlen = 12 bases of a seed from a 256 MiB table and walks
 * the other 8 through the FM index -- 16 dependent sector gathers on a genome where a 12-mer still has hundreds of
 * rows.  With 180 GB of HBM the same question can be asked for the last h2 bases at once: most h2-mers of a genome
 * have one row, and a one-row interval is finished against the text (as_vote.cu), so a seed costs one table sector,
 * one SA sector and one text sector.  The 12-mer table stays (it is what `.lch` holds and what as_lc_aln_batch
 * answers from); PART 1 uses this one when seed_len >= h2.  Results are identical: the interval of the seed's last h2
 * bases is what lc_access + (h2 - hlen) steps of fmi_aln (fmidx.c:295-313) would have reached, including the
 * reference's k = 1 start, which drops the occurrence that ends on the text's last base (see k_lc_codes, as_build.cu).
 *
 * Entry: bits 40..63 = info, 0 = absent.  info >> 22 is a tag:
 *   0: bits 0..39 = first row k, info = rows in the interval, 0x3FFFFF = "2^22-1 or more" (PART 1 then takes the 12-mer
 *      table and walks);
 *   1: ONE row: bits 0..39 = SA[k] ITSELF (the text position, not the row) and info & 0x3FFFFF = the 11 text bases in
 *      front of it (base SA-11 lowest);
 *   2: TWO rows: bits 0..39 = k, info bits 0..9 / 10..19 = the 5 text bases in front of SA[k] / SA[k+1].
 * Tags 1 and 2 let PART 1 decide the rest of the walk from the entry itself (a row survives iff the seed's remaining
 * bases equal the text in front of it).  A one-row entry finishes its seed with NO further gather: the vote needs the
 * position, never the row (on the 4.6 Mbp genome 87 % of the h2-mers a read presents have one row, and PART 1 runs at
 * the DRAM's random-transaction rate -- tools/gather_probe.cu -- so a gather saved is time saved); a seed that dies at
 * a two-row entry -- most seeds of a noisy long read, whose h2-mer exists in a human-size genome by chance -- costs ONE
 * gather and only survivors fetch their position.  Rows closer to the text start than the longest possible rest of a
 * seed (32 bases) keep tag 0.  Built from the device index alone (SA + 2-bit text), so handles
 * opened from the reference's files get it too.
 */
#include "as_common.cuh"
#include <stdlib.h>

#define LC2_NONE 0xFFFFFFFFFFFFFFFFULL
#define LC2_SAT 0x3FFFFFULL

/* LSB-first code of the h-mer starting suffix SA[r], or LC2_NONE (row 0, or the h-mer would touch the last base / '$') */
__device__ __forceinline__ uint64_t lc2_code(const DevIndex &ix, int h, uint64_t r) {
    const uint64_t p = as_sa(ix, r);
    if (r < 1 || p + (uint64_t)h >= ix.N - 1) return LC2_NONE;
    const uint64_t x = h <= 16 ? (uint64_t)as_bases16(ix.text2, p) : as_bases32(ix.text2, p);
    return h >= 32 ? x : (x & ((1ULL << (2 * h)) - 1ULL));
}

__global__ void __launch_bounds__(256) k_lc2_fill(DevIndex ix, int h, unsigned long long *__restrict__ lc2) {
    const uint64_t RUN = 16, N = ix.N;
    const uint64_t nruns = (N + RUN - 1) / RUN;
    for (uint64_t q = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; q < nruns; q += (uint64_t)gridDim.x * blockDim.x) {
        const uint64_t r0 = q * RUN, r1 = r0 + RUN < N ? r0 + RUN : N;
        uint64_t cur = lc2_code(ix, h, r0);
        uint64_t start = r0;                 /* first row of cur's run, as far as this thread has seen */
        bool open_left = r0 > 0 && cur != LC2_NONE && lc2_code(ix, h, r0 - 1) == cur;
        for (uint64_t r = r0; r < r1; r++) {
            const uint64_t next = r + 1 < N ? lc2_code(ix, h, r + 1) : LC2_NONE;
            if (next != cur) {
                if (cur != LC2_NONE) {       /* r closes cur's run: the thread that sees the end writes the entry */
                    uint64_t k = start;
                    if (open_left) {         /* the run began before r0: gallop back, then bisect */
                        uint64_t lo = 0, hi = r0 - 1, step = 1;      /* code(hi) == cur; find the smallest such row */
                        while (hi >= step && lc2_code(ix, h, hi - step) == cur) {
                            hi -= step;
                            step <<= 1;
                        }
                        lo = hi >= step ? hi - step + 1 : 0;          /* code(lo-1) != cur (or lo == 0), code(hi) == cur */
                        while (lo < hi) {
                            const uint64_t mid = lo + ((hi - lo) >> 1);
                            if (lc2_code(ix, h, mid) == cur) hi = mid; else lo = mid + 1;
                        }
                        k = lo;
                    }
                    const uint64_t cnt = r - k + 1;
                    uint64_t info = cnt >= LC2_SAT ? LC2_SAT : cnt;
                    if (cnt == 1) {
                        const uint64_t p = as_sa(ix, k);
                        if (p >= 32) {
                            info = (1ULL << 22) | (as_bases16(ix.text2, p - 11) & 0x3FFFFFu);
                            k = p;                    /* tag 1 carries the position */
                        }
                    } else if (cnt == 2) {
                        const uint64_t p0 = as_sa(ix, k), p1 = as_sa(ix, k + 1);
                        if (p0 >= 5 && p1 >= 5)
                            info = (2ULL << 22) | (as_bases16(ix.text2, p0 - 5) & 0x3FFu) | ((uint64_t)(as_bases16(ix.text2, p1 - 5) & 0x3FFu) << 10);
                    }
                    lc2[cur] = k | (info << 40);
                }
                start = r + 1;
                open_left = false;
            }
            cur = next;
        }
    }
}

/*
 * sax[row] = SA[row] (40 bits) | the 12 text bases in front of that position (24 bits, base SA-12 lowest).  A one-row
 * interval is finished by comparing the seed's remaining bases with the text in front of SA[row] (as_vote.cu); with the
 * bases stored next to the SA value that is ONE gather instead of two (SA, then text) -- on the long-read and human-scale
 * workloads PART 1 moves 5 TB/s of DRAM sectors, so a third fewer gathers is a third less time there.  8 bytes per row on
 * top of the split SA (which stays: export and the sa_access primitive read it); skipped when memory is short.
 */
__global__ void __launch_bounds__(256) k_sax_fill(DevIndex ix, unsigned long long *__restrict__ sax) {
    for (uint64_t r = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; r < ix.N; r += (uint64_t)gridDim.x * blockDim.x) {
        const uint64_t p = as_sa(ix, r);
        uint32_t prev;
        if (p >= 12) prev = as_bases16(ix.text2, p - 12) & 0xFFFFFFu;
        else prev = (as_bases16(ix.text2, 0) << (2u * (12u - (uint32_t)p))) & 0xFFFFFFu;   /* bases [0,p) at the top */
        sax[r] = p | ((unsigned long long)prev << 40);
    }
}

static int build_sax(as_gpu *h) {
    DevIndex &ix = h->ix;
    ix.sax = nullptr;
    if (h->opts.no_sax || h->opts.sampled_sa > 0) return AS_OK;
    size_t free_b = 0, total_b = 0;
    AS_CUDA(cudaMemGetInfo(&free_b, &total_b));
    if (ix.N * 8ULL > free_b / 2) return AS_OK;                   /* leave room for the batches */
    AS_CUDA(cudaMalloc(&h->d_sax, ix.N * sizeof(unsigned long long)));
    uint64_t blocks = (ix.N + 255) / 256, cap = (uint64_t)h->sm_count * 16;
    k_sax_fill<<<(unsigned)(blocks < cap ? blocks : cap), 256, 0, h->stream>>>(ix, (unsigned long long *)h->d_sax);
    h->launches++;
    AS_CUDA(cudaGetLastError());
    AS_CUDA(cudaStreamSynchronize(h->stream));
    ix.sax = (const unsigned long long *)h->d_sax;
    return AS_OK;
}

static int build_lc2_table(as_gpu *h) {
    DevIndex &ix = h->ix;
    int h2 = 1;
    while (h2 < 16 && (1ULL << (2 * h2)) < ix.N) h2++;           /* ceil(log4 N) */
    h2 += 1;
    /* one level deeper while the table stays small (<= 4 GB): fewer entries of several rows, i.e. fewer seeds that need a
     * second gather or a backward-search step (4.6 Mbp genome: h2 13 -> 14, PART 1 1.51 -> 1.27 ms per 1 M reads) */
    if (h2 < 16 && (8ULL << (2 * (h2 + 1))) <= (4ULL << 30)) h2++;
    if (h2 > 16) h2 = 16;
    if (h->opts.lc2_hlen > 0) h2 = h->opts.lc2_hlen;
    if (h2 <= ix.hlen || h2 > 16) return AS_OK;                   /* nothing deeper than the reference's table to offer */
    size_t free_b = 0, total_b = 0;
    AS_CUDA(cudaMemGetInfo(&free_b, &total_b));
    while (h2 > ix.hlen && (8ULL << (2 * h2)) > free_b / 3) h2--;  /* leave room for the batches */
    if (h2 <= ix.hlen) return AS_OK;
    const uint64_t entries = 1ULL << (2 * h2);
    AS_CUDA(cudaMalloc(&h->d_lc2, entries * sizeof(unsigned long long)));
    AS_CUDA(cudaMemsetAsync(h->d_lc2, 0, entries * sizeof(unsigned long long), h->stream));
    const uint64_t nruns = (ix.N + 15) / 16;
    uint64_t blocks = (nruns + 255) / 256, cap = (uint64_t)h->sm_count * 16;
    k_lc2_fill<<<(unsigned)(blocks < cap ? blocks : cap), 256, 0, h->stream>>>(ix, h2, (unsigned long long *)h->d_lc2);
    h->launches++;
    AS_CUDA(cudaGetLastError());
    AS_CUDA(cudaStreamSynchronize(h->stream));
    ix.lc2 = (const unsigned long long *)h->d_lc2;
    ix.hlen2 = h2;
    return AS_OK;
}

/* the derived tables of a device index: the deep k-mer table first (it buys more per byte), then SA+context */
int as_build_lc2(as_gpu *h) {
    DevIndex &ix = h->ix;
    ix.lc2 = nullptr;
    ix.hlen2 = 0;
    ix.sax = nullptr;
    if (h->opts.l2_fetch_gran > 0) AS_CUDA(cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, (size_t)h->opts.l2_fetch_gran));
    if (!ix.text2 || !ix.sa_lo || ix.N < 2) return AS_OK;
    AS_TRY(build_lc2_table(h));
    AS_TRY(build_sax(h));
    if (h->opts.sampled_sa > 0) {
        /* as_open_opts.sampled_sa: keep SA[r * i] only (_build_csa, fmidx.c:152-163).  The deep table above was built from
         * the full array; from here on positions come from csa_access' LF walk (as_locate). */
        AS_TRY(as_csa_sample(h, h->opts.sampled_sa));
        AS_CUDA(cudaStreamSynchronize(h->stream));
        if (h->d_sa_lo) cudaFree(h->d_sa_lo);
        if (h->d_sa_hi) cudaFree(h->d_sa_hi);
        h->d_sa_lo = h->d_sa_hi = nullptr;
        ix.sa_lo = nullptr;
        ix.sa_hi = nullptr;
    }
    return AS_OK;
}
