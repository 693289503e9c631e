/*
 * pack.c -- host side of as_map_batch_packed: 2-bit packing of reads (the layout include/accseq_b200.h describes) and the
 * as_result a compact record stands for.  Plain C, no CUDA.
 */
#include "accseq_host.h"

#include <string.h>

/* four ASCII bases -> eight bits: per byte ((c >> 1) & 3) ^ ((c >> 2) & 1) = A0 C1 G2 T3 (upper or lower case; the code the
 * device's as_base_code gives every other byte too), then the four 2-bit codes gathered into the top byte by one 32-bit
 * multiply: the partial products that matter land on bits 24-25, 26-27, 28-29, 30-31, the others below bit 24 or past bit 31 */
static inline uint32_t pack4(uint32_t x) {
    const uint32_t c = ((x >> 1) & 0x03030303u) ^ ((x >> 2) & 0x01010101u);
    return (uint32_t)(c * 0x01041040u) >> 24;
}

static inline uint32_t pack8(uint64_t x) { return pack4((uint32_t)x) | pack4((uint32_t)(x >> 32)) << 8; }

#if defined(__x86_64__) || defined(__i386__)
#include <immintrin.h>
/* 16 bases per step: the same codes per byte, then b0 + 4 b1 per byte pair (PMADDUBSW), pairs of pairs + 16 x (PMADDWD), and
 * the four bytes that result gathered into one word (PSHUFB) */
__attribute__((target("ssse3"))) static size_t pack_words_ssse3(const char *seq, uint32_t len, uint32_t *out, size_t nwords) {
    const __m128i m3 = _mm_set1_epi8(3), m1 = _mm_set1_epi8(1);
    const __m128i w14 = _mm_set1_epi16(0x0401), w116 = _mm_set1_epi32(0x00100001);
    const __m128i pick = _mm_setr_epi8(0, 4, 8, 12, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1);
    size_t w = 0;
    for (uint32_t i = 0; i + 16 <= len && w < nwords; i += 16, w++) {
        const __m128i x = _mm_loadu_si128((const __m128i *)(seq + i));
        const __m128i c = _mm_xor_si128(_mm_and_si128(_mm_srli_epi16(x, 1), m3), _mm_and_si128(_mm_srli_epi16(x, 2), m1));
        const __m128i q = _mm_madd_epi16(_mm_maddubs_epi16(c, w14), w116);
        out[w] = (uint32_t)_mm_cvtsi128_si32(_mm_shuffle_epi8(q, pick));
    }
    return w;
}
#endif

void as_pack_read(const char *seq, uint32_t len, uint32_t *out, size_t nwords) {
    size_t w = 0;
    uint32_t i = 0;
#if defined(__x86_64__) || defined(__i386__)
    static int have = -1;
    if (have < 0) have = __builtin_cpu_supports("ssse3") ? 1 : 0;
    if (have) {
        w = pack_words_ssse3(seq, len, out, nwords);
        i = (uint32_t)(16 * w);
    }
#endif
    for (; i + 16 <= len && w < nwords; i += 16, w++) {
        uint64_t a, b;
        memcpy(&a, seq + i, 8);
        memcpy(&b, seq + i + 8, 8);
        out[w] = pack8(a) | pack8(b) << 16;
    }
    if (i < len && w < nwords) {
        uint32_t v = 0;
        for (uint32_t t = 0; i + t < len; t++) {
            const uint32_t c = (uint8_t)seq[i + t];
            v |= ((((c >> 1) & 3u) ^ ((c >> 2) & 1u))) << (2 * t);
        }
        out[w++] = v;
    }
    for (; w < nwords; w++) out[w] = 0;
}

int as_pack_reads(const char *ascii, size_t stride, const uint32_t *lens, size_t n, uint32_t *packed, size_t wstride) {
    if (!ascii || !lens || !packed) return AS_EINVAL;
    for (size_t r = 0; r < n; r++) {
        if (((size_t)lens[r] + 15) / 16 > wstride || lens[r] > stride) return AS_EINVAL;
        as_pack_read(ascii + r * stride, lens[r], packed + r * wstride, wstride);
    }
    return AS_OK;
}

/* the fields of as_result a compact record determines (alnmain.c:458-477; loc per seq_lookup, alnmain.c:151-176) */
void as_result_from_compact(const as_result_compact *c, uint32_t len, const as_mta *mta, int mta_len, as_result *r) {
    memset(r, 0, sizeof *r);
    const int seq_id = AS_RC_SEQ_ID(*c);
    r->off = AS_RC_OFF(*c);
    r->seq_id = seq_id;
    r->flag = (int)AS_RC_FLAG(*c);
    r->strand = (int)AS_RC_STRAND(*c);
    r->mapq = (r->flag & 4) ? 0 : 255;
    r->ed = c->ed;
    r->valid = !(r->flag & 4);
    r->n_cigar_op = (int32_t)AS_RC_NRUNS(*c);
    r->myers = -2;
    if (seq_id >= 0 && seq_id < mta_len && mta) {
        const uint64_t start = mta[seq_id].offset, sl = mta[seq_id].seq_len;
        r->loc = r->strand ? start + 2 * sl - r->off - len : start + r->off;
    }
}
