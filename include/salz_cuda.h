/*
 * salz_cuda.h - C ABI of the B200-native salz compression hot path.
 *
 * Plain C: pointers and sizes only.  This is the boundary the host code
 * (lib/salz.c, programs/salzcli.c, or any FFI) binds instead of the
 * reference's per-block CPU encoder:
 *
 *   salz_cuda_encode_blocks  == a loop of salz_encode_safe   (ref lib/salz.c:777-823,
 *                                                              API ref lib/salz.h:42-43)
 *   salz_cuda_encode_stream  == the CLI block loop + framing  (ref programs/salzcli.c:143-179)
 *   salz_cuda_debug_stages   == the intermediates of          (ref lib/salz.c:463-662)
 *
 * Every entry point runs hand-written sm_100a kernels; there is no CPU
 * fallback.  Without a CUDA device ctx_create returns NULL and every other
 * call fails with -1.
 */
#ifndef SALZ_CUDA_H
#define SALZ_CUDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct salz_cuda_ctx salz_cuda_ctx;

/* Per-context counters, accumulated since ctx_create / stats_reset. */
typedef struct salz_cuda_stats {
    uint64_t blocks;          /* blocks encoded */
    uint64_t input_bytes;     /* sum of block lengths */
    uint64_t output_bytes;    /* sum of encoded lengths (block headers included) */
    uint64_t launches;        /* kernels launched */
    uint64_t sort_launches;   /* of which radix scatter passes */
    uint64_t sort_elems;      /* elements moved by the scatter passes (sum over passes) */
    uint64_t sort_bytes;      /* algorithmic bytes of those passes (16 or 24 B per element) */
    uint64_t doubling_rounds; /* prefix-doubling rounds executed (sum over batches) */
    uint64_t alg_bytes_sa;    /* algorithmic HBM bytes, suffix sorting */
    uint64_t alg_bytes_rest;  /* algorithmic HBM bytes, LCP/ANSV/parse/emit */
    double ms_sa;             /* CUDA-event time per stage, milliseconds */
    double ms_sort;           /* of which inside the radix scatter passes */
    double ms_lcp;
    double ms_ansv;
    double ms_parse;
    double ms_emit;
    double ms_total;          /* whole device pipeline (excludes H2D/D2H) */
    double ms_h2d;
    double ms_d2h;
} salz_cuda_stats;

/* Number of CUDA devices visible (0 when none / no driver). */
int salz_cuda_device_count(void);

/*
 * Create a context on `device` with workspace for batches of up to
 * `max_batch_blocks` blocks of up to `max_block_len` bytes each.
 * (replaces encode_ctx_create's per-call calloc, ref lib/salz.c:175-247)
 * Returns NULL on failure (no device, out of memory, bad arguments).
 */
salz_cuda_ctx *salz_cuda_ctx_create(int device, size_t max_block_len,
                                    size_t max_batch_blocks);
void salz_cuda_ctx_destroy(salz_cuda_ctx *ctx);

/* Text of the last error on this context ("" if none). */
const char *salz_cuda_last_error(const salz_cuda_ctx *ctx);

/*
 * Encode `nblocks` independent blocks held in host memory.
 * For block i: srcs[i]/src_lens[i] in, dsts[i] out with capacity dst_lens[i]
 * in and encoded length out, rcs[i] = 0 / -1 with salz_encode_safe's
 * conventions (src_len <= 8, NULL buffers or insufficient capacity => -1,
 * dst_lens[i] untouched on failure).  Output bytes are identical to the
 * reference's.  Returns 0 if the batch ran (inspect rcs), -1 on a CUDA error.
 */
int salz_cuda_encode_blocks(salz_cuda_ctx *ctx, const uint8_t *const *srcs,
                            const size_t *src_lens, uint8_t *const *dsts,
                            size_t *dst_lens, int *rcs, size_t nblocks);

/*
 * Encode a contiguous host stream cut at `block_len` boundaries and write
 * the CLI framing for every block: u32 encoded_len (native endian) followed
 * by the encoded block (ref programs/salzcli.c:162-169).  The 8-byte file
 * header is the caller's.  `block_sizes` (optional) receives the per-block
 * encoded_len; `*nblocks` (optional) the block count.  Fails (-1) when any
 * block is <= 8 bytes (the reference CLI fails on such files too) or when
 * dst_cap is too small.
 */
int salz_cuda_encode_stream(salz_cuda_ctx *ctx, const uint8_t *src,
                            size_t src_len, size_t block_len, uint8_t *dst,
                            size_t dst_cap, size_t *dst_len,
                            uint32_t *block_sizes, size_t *nblocks);

/* Same, but `d_src` and `d_dst` are device pointers on the context's device
 * (inputs already resident in HBM; used for kernel-only timing).
 * block_sizes is a HOST array. */
int salz_cuda_encode_stream_device(salz_cuda_ctx *ctx, const uint8_t *d_src,
                                   size_t src_len, size_t block_len,
                                   uint8_t *d_dst, size_t dst_cap,
                                   size_t *dst_len, uint32_t *block_sizes,
                                   size_t *nblocks);

/*
 * Test hook: run one block (N bytes, n = N - 8) and copy the intermediates
 * to host arrays (any may be NULL):
 *   sa[n]      suffix array of src[0..n)                 (ref lib/salz.c:463-469)
 *   plcp[n], lcp[n]                                      (ref libsais.h:319,329)
 *   psvnsv[2n] (PSV, NSV) text positions, -1 = none      (ref lib/salz.c:471-490)
 *   cand[4n]   (p-psv, psv_len, p-nsv, nsv_len)          (ref lib/salz.c:540-560)
 *   parse[3n]  (offs, len, cost) after the optimal parse (ref lib/salz.c:610-662)
 *   dst/dst_len encoded block
 */
int salz_cuda_debug_stages(salz_cuda_ctx *ctx, const uint8_t *src, size_t N,
                           int32_t *sa, int32_t *plcp, int32_t *lcp,
                           int32_t *psvnsv, int32_t *cand, int32_t *parse,
                           uint8_t *dst, size_t *dst_len);

/* Suffix arrays only, for a batch of texts laid out at `stride` bytes in one
 * host buffer (SA-build GB/s measurements).  sa_out[i*stride + r]. */
int salz_cuda_suffix_arrays(salz_cuda_ctx *ctx, const uint8_t *texts,
                            const uint32_t *lens, size_t stride, size_t ntexts,
                            int32_t *sa_out);

/* Page-locked host buffers for callers that stream large inputs (the CLI):
 * copies from/to them run at full PCIe speed.  NULL on failure. */
void *salz_cuda_host_alloc(size_t bytes);
void salz_cuda_host_free(void *p);

int salz_cuda_get_stats(const salz_cuda_ctx *ctx, salz_cuda_stats *out);
void salz_cuda_reset_stats(salz_cuda_ctx *ctx);

/* Tunables (tests, A/B timing): ranking inside the register-staged radix scatter pass
 * (pairs of blocks above 2 MiB, and SALZ_SCATTER=0).
 * mode 1 (default) = warp peer masks by shared-memory atomicOr,
 * mode 0           = warp peer masks by match_any.
 * The TMA-staged pass of packed lists is chosen by the environment (SALZ_SCATTER). */
void salz_cuda_set_sort_mode(salz_cuda_ctx *ctx, int mode);

/* Segments per block of the optimal parse (replaces nothing in the reference:
 * optimize_factorization, lib/salz.c:610-662, is one serial loop per block).
 * 0 (default) = automatic: enough chains to fill the GPU when the batch has few
 * blocks, one chain per block otherwise; k >= 1 = at most k segments (rounded
 * down to a power of two, segments never shorter than 4096 positions); k <= -1
 * = |k| segments with the shortened first pass forced (automatic mode decides
 * that by itself).  The parse is exact for every k: a parallel check of the
 * recurrence at every position decides whether another pass is needed. */
void salz_cuda_set_parse_segments(salz_cuda_ctx *ctx, int k);

/* Run this context's copies and kernels on the caller's CUDA stream
 * (a cudaStream_t passed as void *; NULL restores the context's own stream),
 * so that events the caller records on that stream bracket the work. */
void salz_cuda_set_stream(salz_cuda_ctx *ctx, void *stream);

/* ---- no-allocation entry point --------------------------------------------------
 *
 * No counterpart in the reference; added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid it).
 *
 * PLCP and LCP of ONE text from its suffix array (the arrays salz_cuda_debug_stages
 * returns as plcp / lcp), all buffers the caller's DEVICE memory on `device`:
 *   d_text[n] the text, d_sa[n] its suffix array, d_lcp[n] out (LCP[0] = 0,
 *   LCP[r] = lcp(SA[r-1], SA[r])), d_plcp[n] out or NULL.
 * The work is queued on `stream` (a cudaStream_t as void *, NULL = the default
 * stream); the call does not synchronise.  It calls no cudaMalloc / cudaFree and no
 * allocating library routine: its scratch (text copy, inverse suffix array, PLCP
 * values) is a static device array of SALZ_CUDA_NOALLOC_MAX_N positions, ~3.3 MiB
 * that every GPU which loads the library holds until exit.  n above that limit is
 * rejected (-2, message on stderr); other errors return -1; n == 0 only performs the
 * once-per-device set-up (do that before a graph capture starts).  All calls share
 * the one array, which keeps the last call's contents (what must start at zero is
 * cleared on `stream` at every call): calls that overlap on the same GPU - from
 * different streams or threads - race; serialise them. */
#define SALZ_CUDA_NOALLOC_MAX_N ((size_t)1 << 18)
int salz_cuda_lcp_noalloc(int device, const uint8_t *d_text, size_t n,
                          const int32_t *d_sa, int32_t *d_plcp, int32_t *d_lcp,
                          void *stream);

/* ---- multi-GPU: blocks sharded over the GPUs of one box -------------
 *
 * Replaces nothing in the reference (it is single-threaded); the unit that is
 * sharded is the independent block of its CLI loop (ref programs/salzcli.c:143-179,
 * context per call: ref lib/salz.c:788,821).  Rank r of `world` encodes the contiguous
 * block range salz_cuda_block_range(nblocks, r, world) with its own salz_cuda_ctx; the
 * only exchange is an NCCL all-gather of the per-block encoded sizes, after which every
 * rank scans them into record offsets (framing: ref programs/salzcli.c:162-169).
 * NCCL is bound at run time (libnccl.so.2); without it these calls fail with -1 / NULL
 * and everything above keeps working. */

#define SALZ_CUDA_UNIQUE_ID_BYTES 128

typedef struct salz_cuda_comm salz_cuda_comm;
typedef struct salz_cuda_multi salz_cuda_multi;

/* [lo, hi) of the `nblocks` blocks that rank `rank` of `world` encodes. */
void salz_cuda_block_range(size_t nblocks, int rank, int world, size_t *lo, size_t *hi);

/* One rank per caller (a process or a thread per GPU).  Rank 0 makes the id
 * (ncclGetUniqueId) and ships it to the others by whatever means the launcher
 * has; comm_create blocks until all `world` ranks have joined. */
int salz_cuda_comm_unique_id(uint8_t id[SALZ_CUDA_UNIQUE_ID_BYTES]);
salz_cuda_comm *salz_cuda_comm_create(int device, const uint8_t id[SALZ_CUDA_UNIQUE_ID_BYTES],
                                      int rank, int world);
void salz_cuda_comm_destroy(salz_cuda_comm *comm);
int salz_cuda_comm_rank(const salz_cuda_comm *comm);
int salz_cuda_comm_world(const salz_cuda_comm *comm);
uint64_t salz_cuda_comm_gathers(const salz_cuda_comm *comm);   /* all-gathers executed */

/* Collective: every rank passes the encoded sizes of its own block range
 * (nlocal must equal the length of that range); all ranks receive the sizes of
 * all `nblocks_total` blocks in stream order and, if record_offsets != NULL, the
 * byte offset of every record (u32 length + block) in the assembled stream,
 * nblocks_total + 1 entries, the last being the stream length.  The 8-byte file
 * header is the caller's, as in salz_cuda_encode_stream. */
int salz_cuda_comm_allgather_sizes(salz_cuda_comm *comm, const uint32_t *local_sizes,
                                   size_t nlocal, size_t nblocks_total,
                                   uint32_t *all_sizes, uint64_t *record_offsets);

/* One process driving `ndev` devices (devices == NULL: 0..ndev-1; ndev <= 0: all
 * visible): a host thread, a salz_cuda_ctx and an NCCL rank per device. */
salz_cuda_multi *salz_cuda_multi_create(const int *devices, int ndev,
                                        size_t max_block_len, size_t max_batch_blocks);
void salz_cuda_multi_destroy(salz_cuda_multi *m);
int salz_cuda_multi_devices(const salz_cuda_multi *m);
salz_cuda_ctx *salz_cuda_multi_ctx(salz_cuda_multi *m, int i);   /* rank i's context (stats) */
const char *salz_cuda_multi_last_error(const salz_cuda_multi *m);

/* salz_cuda_encode_stream over all devices of `m`: same arguments, same bytes. */
int salz_cuda_multi_encode_stream(salz_cuda_multi *m, const uint8_t *src,
                                  size_t src_len, size_t block_len, uint8_t *dst,
                                  size_t dst_cap, size_t *dst_len,
                                  uint32_t *block_sizes, size_t *nblocks);

#ifdef __cplusplus
}
#endif
#endif /* SALZ_CUDA_H */
