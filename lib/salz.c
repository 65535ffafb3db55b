/*
 * salz.c - host side of the B200-native salz library.
 *
 *   salz_encode_safe : argument checks, then one block through the CUDA
 *                      pipeline (include/salz_cuda.h).  Replaces reference
 *                      lib/salz.c:777-823.  NO CPU fallback.
 *   salz_decode_safe : host C decoder for the unchanged stream format
 *                      (reference lib/salz.c:829-1228).
 *   encode_vnibble_le, vnibble_size : kept as exported symbols
 *                      (reference lib/salz.c:352,565).
 */
#include "salz.h"
#include "salz_cuda.h"

#include <pthread.h>
#include <stdlib.h>
#include <string.h>

/* ---- encoder: process-wide lazily created GPU context ------------------- */

static pthread_mutex_t g_lock = PTHREAD_MUTEX_INITIALIZER;
static salz_cuda_ctx *g_ctx;
static size_t g_ctx_block;

static void drop_ctx(void)
{
    if (g_ctx) {
        salz_cuda_ctx_destroy(g_ctx);
        g_ctx = NULL;
        g_ctx_block = 0;
    }
}

/* Releases the process-wide encoder context (device workspace, pinned buffers).
 * Not part of the reference API: a host application that wants the memory back
 * before exit calls it; the next salz_encode_safe creates a fresh context.
 * Nothing is torn down at exit: by then the CUDA runtime's own exit handler
 * may already have run, and the driver reclaims everything anyway. */
void salz_shutdown(void)
{
    pthread_mutex_lock(&g_lock);
    drop_ctx();
    pthread_mutex_unlock(&g_lock);
}

int salz_encode_safe(const uint8_t *src, size_t src_len, uint8_t *dst, size_t *dst_len)
{
    if (src == NULL || dst == NULL || dst_len == NULL)
        return -1;
    /* the reference fails for < 8 bytes and crashes for exactly 8 */
    if (src_len <= 8 || *dst_len < 4)
        return -1;

    int rc = -1;
    pthread_mutex_lock(&g_lock);
    if (g_ctx == NULL || g_ctx_block < src_len) {
        size_t cap = 1u << 20;
        while (cap < src_len)
            cap <<= 1;
        const char *dev = getenv("SALZ_CUDA_DEVICE");
        drop_ctx();
        g_ctx = salz_cuda_ctx_create(dev ? atoi(dev) : 0, cap, 1);
        g_ctx_block = g_ctx ? cap : 0;
    }
    if (g_ctx != NULL) {
        const uint8_t *srcs[1] = { src };
        uint8_t *dsts[1] = { dst };
        size_t lens[1] = { src_len };
        size_t caps[1] = { *dst_len };
        int rcs[1] = { -1 };
        if (salz_cuda_encode_blocks(g_ctx, srcs, lens, dsts, caps, rcs, 1) == 0 && rcs[0] == 0) {
            *dst_len = caps[0];
            rc = 0;
        }
    }
    pthread_mutex_unlock(&g_lock);
    return rc;
}

/* ---- vnibble helpers -------------------------------------------------------
 * Bijective base-8 code, 4 bits per nibble: 3 payload bits plus a stop flag
 * (0x8) on the last nibble.  k nibbles cover 8^k values starting at
 * 8 + 8^2 + ... + 8^(k-1). */

size_t vnibble_size(uint32_t val)
{
    uint64_t first_beyond = 8, span = 8;
    size_t k = 1;
    while (k < 11 && val >= first_beyond) {
        span <<= 3;
        first_beyond += span;
        k++;
    }
    return k;
}

/* Nibbles packed little-endian into *res so that writing the low 4*k bits
 * MSB-first emits them in stream order; returns k. */
size_t encode_vnibble_le(uint32_t val, uint64_t *res)
{
    uint64_t packed = (val & 7u) | 8u;
    int64_t rest = ((int64_t)val >> 3) - 1;
    size_t k = 1;
    while (rest >= 0) {
        packed |= (uint64_t)(rest & 7) << (4 * k);
        rest = (rest >> 3) - 1;
        k++;
    }
    /* the reference stores bytes, i.e. the value modulo 2^64 in LE order */
    uint8_t *p = (uint8_t *)res;
    for (size_t i = 0; i < (k + 1) / 2; i++)
        p[i] = (uint8_t)(packed >> (8 * i));
    return k;
}

/* ---- decoder ----------------------------------------------------------------
 * Payload = tokens.  Bit words (u64, little-endian, consumed MSB first) are
 * pulled from the byte stream when a bit is needed and none is buffered;
 * literal bytes and the low offset byte are read directly in between. */

struct bitsrc {
    const uint8_t *p, *end;
    uint64_t word;
    unsigned left;
};

static inline int refill(struct bitsrc *s)
{
    if (s->end - s->p < 8)
        return 0;
    memcpy(&s->word, s->p, 8);
    s->p += 8;
    s->left = 64;
    return 1;
}

static inline int take_bits(struct bitsrc *s, unsigned cnt, uint32_t *out)
{
    uint32_t v = 0;
    while (cnt) {
        if (s->left == 0 && !refill(s))
            return 0;
        unsigned t = cnt < s->left ? cnt : s->left;
        v = (v << t) | (uint32_t)(s->word >> (64 - t)); /* t <= 4 here */
        s->word <<= t;
        s->left -= t;
        cnt -= t;
    }
    *out = v;
    return 1;
}

static inline int take_unary(struct bitsrc *s, uint32_t *out)
{
    uint32_t zeros = 0;
    for (;;) {
        if (s->left == 0 && !refill(s))
            return 0;
        if (s->word == 0) {
            zeros += s->left;
            s->left = 0;
            continue;
        }
        unsigned lead = (unsigned)__builtin_clzll(s->word);
        zeros += lead;
        s->word = (lead == 63) ? 0 : (s->word << (lead + 1));
        s->left -= lead + 1;
        *out = zeros;
        return 1;
    }
}

int salz_decode_safe(const uint8_t *src, size_t src_len, uint8_t *dst, size_t *dst_len)
{
    if (src == NULL || dst == NULL || dst_len == NULL || src_len < 4)
        return -1;

    uint32_t hdr;
    memcpy(&hdr, src, 4);
    const uint32_t type = hdr >> 24;
    const size_t payload = hdr & 0xffffffu;
    if (type > 1 || payload > src_len - 4)
        return -1;

    const size_t cap = *dst_len;
    if (type == 0) {
        if (payload > cap)
            return -1;
        memcpy(dst, src + 4, payload);
        *dst_len = payload;
        return 0;
    }

    struct bitsrc s = { src + 4, src + 4 + payload, 0, 0 };
    size_t out = 0;
    while (s.p < s.end) {
        uint32_t tok;
        if (!take_bits(&s, 1, &tok))
            return -1;
        if (tok == 0) {
            if (s.p >= s.end || out >= cap)
                return -1;
            dst[out++] = *s.p++;
            continue;
        }
        uint32_t hi = 0, nib;
        for (int k = 0;; k++) {
            if (!take_bits(&s, 4, &nib))
                return -1;
            hi = k ? (((hi + 1) << 3) | (nib & 7u)) : (nib & 7u);
            if ((nib & 8u) || k == 10)
                break;
        }
        if (s.p >= s.end)
            return -1;
        /* 32-bit arithmetic as in the reference (lib/salz.c:1101-1124): an over-long vnibble wraps */
        const uint32_t offs = ((hi << 8) | *s.p++) + 1u;
        uint32_t q, low;
        if (!take_unary(&s, &q) || !take_bits(&s, 3, &low))
            return -1;
        const uint32_t len = ((q << 3) | low) + 3u;
        /* the reference does not check the offset against the bytes produced so far and reads
         * before its output buffer (undefined); such streams are rejected here */
        if (len > cap - out || offs > out || offs == 0)
            return -1;
        if (offs >= len) {
            memcpy(dst + out, dst + out - offs, len);
            out += len;
        } else {
            for (size_t i = 0; i < len; i++, out++)
                dst[out] = dst[out - offs];
        }
    }
    *dst_len = out;
    return 0;
}
