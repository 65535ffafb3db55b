/*
 * salz_oracle.h - CPU restatement of the salz compression hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is on the product path:
 * only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this library.  The product (lib/ +
 * salz_b200/csrc) never links or calls it and fails loudly without CUDA.
 *
 * Each function names the reference lines it restates (paths relative to
 * the upstream salz tree).  Parity of this restatement is pinned against the
 * compiled reference (oracle/_ref/libsalz_ref.so, built in place from the
 * upstream sources by oracle/Makefile) and against the committed vectors in
 * tests/golden/.
 */
#ifndef SALZ_ORACLE_H
#define SALZ_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Suffix array of T[0..n): plain lexicographic order, proper prefix first
 * (what lib/salz.c:465 obtains from libsais).  Returns 0 / -1. */
int oracle_suffix_array(const uint8_t *T, int32_t n, int32_t *SA);

/* PLCP[i] = lcp(i, Phi[i]), LCP[r] = PLCP[SA[r]], LCP[0] = 0
 * (libsais.h:319,329 semantics; salz itself never materialises these). */
void oracle_plcp_lcp(const uint8_t *T, int32_t n, const int32_t *SA,
                     int32_t *PLCP, int32_t *LCP);

/* PSV/NSV in text order (lib/salz.c:471-490): psv[p], nsv[p] = text position
 * of the nearest SA neighbour to the left / right with a smaller text
 * position, -1 if none. */
void oracle_psvnsv(const int32_t *SA, int32_t n, int32_t *psv, int32_t *nsv);

/* Candidate table (lib/salz.c:492-560): cand[4p..4p+3] =
 * (p-psv, lcp(psv,p), p-nsv, lcp(nsv,p)); position 0 = (-1,1,-1,1). */
void oracle_candidates(const uint8_t *T, int32_t n, const int32_t *psv,
                       const int32_t *nsv, int32_t *cand);

/* Bit-cost model (lib/salz.c:562-608). */
size_t oracle_vnibble_size(uint32_t v);
int32_t oracle_factor_cost(int32_t offs, int32_t len);

/* Optimal parse (lib/salz.c:610-662).  parse[3p..3p+2] = (offs,len,cost)
 * for p in [1,n); position 0 is the forced literal (0,1,cost as computed
 * by the same recurrence).  cost_n (entry n) is 0. */
void oracle_parse(const int32_t *cand, int32_t n, int32_t *parse);

/* Sequential bit/byte writer + token emission + finalisation
 * (lib/salz.c:258-457, 664-775).  src has N = n + 8 bytes.  Returns 0 and the
 * encoded length in *dst_len, or -1 when dst is too small at any write. */
int oracle_emit(const uint8_t *src, size_t N, const int32_t *parse,
                uint8_t *dst, size_t *dst_len);

/* Whole block (lib/salz.c:777-823).  N <= 8 returns -1 (the reference fails
 * for N < 8 and has undefined behaviour for N == 8). */
int oracle_encode(const uint8_t *src, size_t N, uint8_t *dst, size_t *dst_len);

/* Decoder restatement (lib/salz.c:829-1228). */
int oracle_decode(const uint8_t *src, size_t src_len, uint8_t *dst,
                  size_t *dst_len);

/* Stage dump used by the parity tests: any output pointer may be NULL. */
int oracle_stages(const uint8_t *src, size_t N, int32_t *SA, int32_t *PLCP,
                  int32_t *LCP, int32_t *cand, int32_t *parse, uint8_t *dst,
                  size_t *dst_len);

#ifdef __cplusplus
}
#endif
#endif
