/*
 * ref_shim.c - builds the UNMODIFIED upstream salz + libsais sources, where
 * they lie under $(REF) (default /root/reference), into
 * oracle/_ref/libsalz_ref.so and adds stage-dump entry points.
 *
 * TEST INFRASTRUCTURE ONLY.  No upstream source is copied into this
 * repository: the translation unit below #includes lib/salz.c from the
 * read-only reference tree so that its static stage functions
 * (lib/salz.c:463-775) can be driven one at a time.  The resulting .so is
 * git-ignored and travels to the GPU box as a prebuilt artefact.
 */
#include "salz.c"      /* found through -I$(REF)/lib */
#include "libsais.h"

/* Dump every intermediate of salz_encode_safe (lib/salz.c:777-823).
 * sa: n, psvnsv: 2n (psv,nsv interleaved), cand: 4n, parse: 3n (offs,len,cost)
 * Any pointer may be NULL. */
int ref_stages(const uint8_t *src, size_t N, int32_t *sa, int32_t *psvnsv,
               int32_t *cand, int32_t *parse, uint8_t *dst, size_t *dst_len)
{
    if (src == NULL || N <= 8)
        return -1;
    size_t cap = dst ? *dst_len : 0;
    uint8_t *tmp = NULL;
    if (dst == NULL) {
        cap = (size_t)salz_encoded_len_max(N);
        tmp = malloc(cap);
        dst = tmp;
    }
    salz_io_ctx *ctx = encode_ctx_create(src, N, dst, cap);
    if (ctx == NULL) {
        free(tmp);
        return -1;
    }
    size_t n = ctx->src_len;
    int rc = -1;
    if (!build_suffix_array(ctx))
        goto out;
    if (sa)
        memcpy(sa, ctx->sa + 1, n * sizeof(int32_t));
    build_psvnsv_array(ctx);
    if (psvnsv)
        for (size_t p = 0; p < n; p++) {
            psvnsv[2 * p] = ctx->aux[4 * p];
            psvnsv[2 * p + 1] = ctx->aux[4 * p + 1];
        }
    factorize(ctx);
    if (cand)
        memcpy(cand, ctx->aux, 4 * n * sizeof(int32_t));
    optimize_factorization(ctx);
    if (parse)
        for (size_t p = 0; p < n; p++) {
            parse[3 * p] = ctx->aux[4 * p];
            parse[3 * p + 1] = ctx->aux[4 * p + 1];
            parse[3 * p + 2] = ctx->aux[4 * p + 2];
        }
    if (!emit_encoding(ctx) || !finalize_encoding(ctx))
        goto out;
    if (dst_len)
        *dst_len = ctx->dst_pos;
    rc = 0;
out:
    encode_ctx_destroy(ctx);
    free(tmp);
    return rc;
}

int ref_suffix_array(const uint8_t *T, int32_t *SA, int32_t n)
{
    return libsais(T, SA, n, 0, NULL);
}

#ifdef LIBSAIS_OPENMP
int ref_suffix_array_omp(const uint8_t *T, int32_t *SA, int32_t n, int32_t threads)
{
    return libsais_omp(T, SA, n, 0, NULL, threads);
}
#endif

int ref_plcp_lcp(const uint8_t *T, const int32_t *SA, int32_t *PLCP,
                 int32_t *LCP, int32_t n)
{
    int rc = libsais_plcp(T, SA, PLCP, n);
    if (rc == 0 && LCP)
        rc = libsais_lcp(PLCP, SA, LCP, n);
    return rc;
}
