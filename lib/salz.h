/*
 * salz.h - public codec API of the B200-native salz library.
 *
 * Drop-in for the reference's lib/salz.h:25-58: the same three entry points
 * with the same argument meaning and 0 / -1 error convention.
 * salz_encode_safe runs on the GPU through include/salz_cuda.h (no CPU
 * fallback: it returns -1 when no CUDA device can be used);
 * salz_decode_safe is host C.
 */
#ifndef SALZ_H
#define SALZ_H

#include <stddef.h>
#include <stdint.h>

#include "common.h"

#ifdef __cplusplus
extern "C" {
#endif

/* Upper bound of an encoded block: 4-byte header, one byte per input byte
 * and one token bit per input byte rounded up to whole 64-bit words
 * (reference lib/salz.h:25-28; used by callers to size both buffers). */
static inline int salz_encoded_len_max(size_t plain_len)
{
    return 4 + plain_len + roundup(plain_len, 64) / 8;
}

/* Encode one block.  *dst_len: capacity in, encoded length out (success only).
 * Returns 0, or -1 (NULL buffers, src_len <= 8, insufficient capacity, no GPU). */
extern int salz_encode_safe(const uint8_t *src, size_t src_len, uint8_t *dst,
                            size_t *dst_len);

/* Decode one block.  *dst_len: capacity in, decoded length out.
 * Returns 0, or -1 (NULL buffers, malformed / truncated stream, overflow). */
extern int salz_decode_safe(const uint8_t *src, size_t src_len, uint8_t *dst,
                            size_t *dst_len);

/* Exported by the reference library as well (lib/salz.c:352,565). */
extern size_t encode_vnibble_le(uint32_t val, uint64_t *res);
extern size_t vnibble_size(uint32_t val);

/* Not in the reference: releases the process-wide GPU context that
 * salz_encode_safe creates lazily (device workspace and pinned buffers). */
extern void salz_shutdown(void);

#ifdef __cplusplus
}
#endif
#endif /* SALZ_H */
