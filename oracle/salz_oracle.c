/*
 * salz_oracle.c - CPU restatement of the salz compression hot path.
 *
 * TEST INFRASTRUCTURE ONLY (see salz_oracle.h).  Written from the behaviour of
 * the upstream sources, not copied from them; every function cites the lines
 * it restates.  Pinned against the compiled reference by
 * tests/test_oracle_vs_ref.py and against tests/golden/.
 */
#include "salz_oracle.h"

#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------------
 * Suffix array.  The reference calls libsais(T, SA, n, 0, NULL)
 * (lib/salz.c:463-469).  The suffix array is a unique mathematical object, so
 * the restatement is a textbook Manber-Myers prefix doubling with two
 * counting-sort passes per round: O(n log n), no sentinel, a suffix that is a
 * proper prefix of another sorts first.
 * ---------------------------------------------------------------------- */
int oracle_suffix_array(const uint8_t *T, int32_t n, int32_t *SA)
{
    if (n < 0)
        return -1;
    if (n == 0)
        return 0;

    int32_t *rank = malloc(sizeof(int32_t) * (size_t)n);
    int32_t *tmp = malloc(sizeof(int32_t) * (size_t)n);
    int32_t *cnt = malloc(sizeof(int32_t) * ((size_t)n + 257));
    int32_t *nrank = malloc(sizeof(int32_t) * (size_t)n);
    if (!rank || !tmp || !cnt || !nrank) {
        free(rank); free(tmp); free(cnt); free(nrank);
        return -1;
    }

    /* Round 0: counting sort on the first byte. */
    memset(cnt, 0, sizeof(int32_t) * 257);
    for (int32_t i = 0; i < n; i++)
        cnt[T[i] + 1]++;
    for (int c = 0; c < 256; c++)
        cnt[c + 1] += cnt[c];
    for (int32_t i = 0; i < n; i++)
        SA[cnt[T[i]]++] = i;
    /* rank = 1 + number of strictly smaller first bytes (0 is "past the end") */
    rank[SA[0]] = 1;
    for (int32_t r = 1; r < n; r++)
        rank[SA[r]] = (T[SA[r]] == T[SA[r - 1]]) ? rank[SA[r - 1]] : r + 1;

    for (int32_t h = 1;; h <<= 1) {
        /* Sort by secondary key rank[i+h] (0 when i+h >= n): the suffixes
         * with i+h >= n come first in increasing... any order, they are
         * already distinguished by the primary key being unique for them. */
        memset(cnt, 0, sizeof(int32_t) * ((size_t)n + 2));
        for (int32_t i = 0; i < n; i++) {
            int32_t k = (i + h < n) ? rank[i + h] : 0;
            cnt[k + 1]++;
        }
        for (int32_t k = 0; k <= n; k++)
            cnt[k + 1] += cnt[k];
        for (int32_t i = 0; i < n; i++) {
            int32_t k = (i + h < n) ? rank[i + h] : 0;
            tmp[cnt[k]++] = i;
        }
        /* Stable sort by primary key rank[i]. */
        memset(cnt, 0, sizeof(int32_t) * ((size_t)n + 2));
        for (int32_t i = 0; i < n; i++)
            cnt[rank[i] + 1]++;
        for (int32_t k = 0; k <= n; k++)
            cnt[k + 1] += cnt[k];
        for (int32_t j = 0; j < n; j++) {
            int32_t i = tmp[j];
            SA[cnt[rank[i]]++] = i;
        }
        /* Re-rank. */
        int distinct = 1;
        nrank[SA[0]] = 1;
        for (int32_t r = 1; r < n; r++) {
            int32_t a = SA[r - 1], b = SA[r];
            int32_t a2 = (a + h < n) ? rank[a + h] : 0;
            int32_t b2 = (b + h < n) ? rank[b + h] : 0;
            if (rank[a] == rank[b] && a2 == b2) {
                nrank[b] = nrank[a];
                distinct = 0;
            } else {
                nrank[b] = r + 1;
            }
        }
        memcpy(rank, nrank, sizeof(int32_t) * (size_t)n);
        if (distinct || h >= n)
            break;
    }

    free(rank); free(tmp); free(cnt); free(nrank);
    return 0;
}

/* ------------------------------------------------------------------------
 * PLCP / LCP (libsais.h:319,329; libsais.c:7649-7833 semantics).  Restated as
 * the classic phi algorithm: Phi[SA[r]] = SA[r-1], amortised left-to-right
 * compare with the "drop one" lower bound.
 * ---------------------------------------------------------------------- */
void oracle_plcp_lcp(const uint8_t *T, int32_t n, const int32_t *SA,
                     int32_t *PLCP, int32_t *LCP)
{
    if (n <= 0)
        return;
    int32_t *phi = malloc(sizeof(int32_t) * (size_t)n);
    phi[SA[0]] = -1;
    for (int32_t r = 1; r < n; r++)
        phi[SA[r]] = SA[r - 1];
    int32_t l = 0;
    for (int32_t i = 0; i < n; i++) {
        int32_t j = phi[i];
        if (j < 0) {
            PLCP[i] = 0;
            l = 0;
            continue;
        }
        while (i + l < n && j + l < n && T[i + l] == T[j + l])
            l++;
        PLCP[i] = l;
        if (l > 0)
            l--;
    }
    if (LCP)
        for (int32_t r = 0; r < n; r++)
            LCP[r] = PLCP[SA[r]];
    free(phi);
}

/* ------------------------------------------------------------------------
 * PSV / NSV (lib/salz.c:471-490).  Monotone stack over SA order; values are
 * text positions, -1 when no smaller neighbour exists on that side.
 * ---------------------------------------------------------------------- */
void oracle_psvnsv(const int32_t *SA, int32_t n, int32_t *psv, int32_t *nsv)
{
    if (n <= 0)
        return;
    int32_t *stack = malloc(sizeof(int32_t) * ((size_t)n + 1));
    int32_t top = 0;
    stack[0] = -1;
    for (int32_t r = 0; r <= n; r++) {
        int32_t v = (r < n) ? SA[r] : -1;
        while (stack[top] > v) {
            int32_t p = stack[top];
            psv[p] = stack[top - 1];
            nsv[p] = v;
            top--;
        }
        stack[++top] = v;
    }
    free(stack);
}

/* Longest common prefix of T[a..n) and T[b..n), plain definition. */
static int32_t naive_lcp(const uint8_t *T, int32_t n, int32_t a, int32_t b,
                         int32_t start)
{
    int32_t hi = (a > b) ? a : b;
    int32_t l = start;
    while (hi + l < n && T[a + l] == T[b + l])
        l++;
    return l;
}

/* ------------------------------------------------------------------------
 * Candidates (lib/salz.c:492-560).  The reference's common_len trick is an
 * amortisation only: lcp(PSV[p],p) >= lcp(PSV[p-1],p-1) - 1, so
 * the compare may start at prev-1.  We keep the same bound so the oracle
 * stays linear on repetitive inputs.
 * ---------------------------------------------------------------------- */
void oracle_candidates(const uint8_t *T, int32_t n, const int32_t *psv,
                       const int32_t *nsv, int32_t *cand)
{
    if (n <= 0)
        return;
    cand[0] = -1; cand[1] = 1; cand[2] = -1; cand[3] = 1;
    int32_t pl = 0, nl = 0;
    for (int32_t p = 1; p < n; p++) {
        int32_t a = psv[p], b = nsv[p];
        int32_t la = 0, lb = 0;
        if (a != -1)
            la = naive_lcp(T, n, a, p, pl > 0 ? pl - 1 : 0);
        if (b != -1)
            lb = naive_lcp(T, n, b, p, nl > 0 ? nl - 1 : 0);
        pl = la; nl = lb;
        cand[4 * (size_t)p + 0] = p - a;
        cand[4 * (size_t)p + 1] = la;
        cand[4 * (size_t)p + 2] = p - b;
        cand[4 * (size_t)p + 3] = lb;
    }
}

/* ------------------------------------------------------------------------
 * Cost model (lib/salz.c:562-608).  vnibble code: 3 payload bits per nibble,
 * bijective numeration, so k nibbles cover [T_k, T_{k+1}) with
 * T_1 = 0, T_{k+1} = T_k + 8^k  (8, 72, 584, 4680, ...).
 * ---------------------------------------------------------------------- */
size_t oracle_vnibble_size(uint32_t v)
{
    uint64_t limit = 8, span = 8;
    size_t k = 1;
    while (k < 11 && (uint64_t)v >= limit) {
        span *= 8;
        limit += span;
        k++;
    }
    return k;
}

/* token bit + offset (8 direct bits + vnibble of (offs-1)>>8) + GR3 length */
int32_t oracle_factor_cost(int32_t offs, int32_t len)
{
    uint32_t hi = ((uint32_t)offs - 1u) >> 8;
    uint32_t l = (uint32_t)len - 3u;
    return (int32_t)(1 + 8 + 4 * oracle_vnibble_size(hi) + (l >> 3) + 1 + 3);
}

/* ------------------------------------------------------------------------
 * Optimal parse (lib/salz.c:610-662).  Right-to-left shortest path; ties are
 * broken by strict '<' in the order literal, PSV candidate, NSV candidate.
 * ---------------------------------------------------------------------- */
void oracle_parse(const int32_t *cand, int32_t n, int32_t *parse)
{
    if (n <= 0)
        return;
    int32_t *cost = malloc(sizeof(int32_t) * ((size_t)n + 1));
    cost[n] = 0;
    for (int32_t p = n - 1; p >= 1; p--) {
        int32_t bo = 0, bl = 1, bc = 9 + cost[p + 1];
        for (int k = 0; k < 2; k++) {
            int32_t o = cand[4 * (size_t)p + 2 * k];
            int32_t l = cand[4 * (size_t)p + 2 * k + 1];
            if (l >= 3) {
                int32_t c = oracle_factor_cost(o, l) + cost[p + l];
                if (c < bc) {
                    bc = c; bo = o; bl = l;
                }
            }
        }
        cost[p] = bc;
        parse[3 * (size_t)p + 0] = bo;
        parse[3 * (size_t)p + 1] = bl;
        parse[3 * (size_t)p + 2] = bc;
    }
    /* position 0: forced literal (lib/salz.c:546-548) */
    parse[0] = 0;
    parse[1] = 1;
    parse[2] = 9 + ((n >= 2) ? cost[1] : 0);
    free(cost);
}

/* ------------------------------------------------------------------------
 * Sequential writer (lib/salz.c:258-350): 64-bit MSB-first bit words whose
 * 8-byte slot in the output is reserved at the moment their first bit is
 * produced, interleaved with directly written bytes.
 * ---------------------------------------------------------------------- */
struct writer {
    uint8_t *dst;
    size_t cap;
    size_t pos;       /* next free output byte */
    uint64_t acc;     /* bits of the open word, right-aligned */
    unsigned room;    /* bits still free in the open word */
    size_t slot;      /* where the open word will be stored */
    int ok;
};

static void store_word(struct writer *w)
{
    /* the reference stores before checking capacity (lib/salz.c:272-275);
     * the very first call stores 8 zero bytes at slot 0, later covered by
     * the header, so only the capacity behaviour matters here */
    if (w->slot + 8 <= w->cap)
        memcpy(w->dst + w->slot, &w->acc, 8);
}

static void open_word(struct writer *w)
{
    store_word(w);
    if (w->pos + 8 > w->cap) {
        w->ok = 0;
        return;
    }
    w->acc = 0;
    w->room = 64;
    w->slot = w->pos;
    w->pos += 8;
}

static void put_bits(struct writer *w, uint64_t v, unsigned count)
{
    while (count && w->ok) {
        if (w->room == 0) {
            open_word(w);
            if (!w->ok)
                return;
        }
        unsigned take = count < w->room ? count : w->room;
        uint64_t chunk = (v >> (count - take)) &
                         ((take == 64) ? ~0ull : ((1ull << take) - 1));
        w->acc = (take == 64) ? chunk : ((w->acc << take) | chunk);
        w->room -= take;
        count -= take;
    }
}

static void put_zeros(struct writer *w, size_t count)
{
    while (count && w->ok) {
        unsigned take = count > 32 ? 32 : (unsigned)count;
        put_bits(w, 0, take);
        count -= take;
    }
}

static void put_byte(struct writer *w, uint8_t b)
{
    if (!w->ok)
        return;
    if (w->pos >= w->cap) {
        w->ok = 0;
        return;
    }
    w->dst[w->pos++] = b;
}

/* vnibble bits, first-written nibble first (lib/salz.c:352-457): the last
 * nibble carries the 0x8 stop flag and the low 3 bits; each earlier nibble
 * holds the next 3 bits of the bijective base-8 expansion. */
static void put_vnibble(struct writer *w, uint32_t v)
{
    uint8_t nib[12];
    int k = 0;
    int64_t x = v;
    nib[k++] = (uint8_t)((x & 7) | 8);
    x = (x >> 3) - 1;
    while (x >= 0) {
        nib[k++] = (uint8_t)(x & 7);
        x = (x >> 3) - 1;
    }
    while (k--)
        put_bits(w, nib[k], 4);
}

int oracle_emit(const uint8_t *src, size_t N, const int32_t *parse,
                uint8_t *dst, size_t *dst_len)
{
    if (N <= 8 || *dst_len < 4)
        return -1;
    size_t n = N - 8;
    struct writer w = { dst, *dst_len, 4, 0, 0, 0, 1 };

    /* lib/salz.c:701-732 */
    size_t p = 0;
    while (p < n && w.ok) {
        uint32_t len = (uint32_t)parse[3 * p + 1];
        if (len == 1) {
            put_bits(&w, 0, 1);
            put_byte(&w, src[p]);
            p += 1;
        } else {
            uint32_t offs = (uint32_t)parse[3 * p + 0];
            put_bits(&w, 1, 1);
            put_vnibble(&w, (offs - 1) >> 8);
            put_byte(&w, (uint8_t)((offs - 1) & 0xff));
            put_zeros(&w, (len - 3) >> 3);
            put_bits(&w, 1, 1);
            put_bits(&w, (len - 3) & 7, 3);
            p += len;
        }
    }
    /* lib/salz.c:734-775 */
    for (size_t i = 0; i < 8 && w.ok; i++) {
        put_bits(&w, 0, 1);
        put_byte(&w, src[n + i]);
    }
    if (!w.ok)
        return -1;
    w.acc = (w.room >= 64) ? 0 : (w.acc << w.room);
    store_word(&w);

    uint32_t hdr;
    if (w.pos > N + 4) {
        if (N + 4 > w.cap)
            return -1;
        hdr = (0u << 24) | (uint32_t)(N & 0xffffff);
        memcpy(dst + 4, src, N);
        w.pos = N + 4;
    } else {
        hdr = (1u << 24) | (uint32_t)((w.pos - 4) & 0xffffff);
    }
    memcpy(dst, &hdr, 4);
    *dst_len = w.pos;
    return 0;
}

int oracle_stages(const uint8_t *src, size_t N, int32_t *SA_out,
                  int32_t *PLCP_out, int32_t *LCP_out, int32_t *cand_out,
                  int32_t *parse_out, uint8_t *dst, size_t *dst_len)
{
    if (src == NULL || N <= 8)
        return -1;
    int32_t n = (int32_t)(N - 8);
    int rc = -1;
    int32_t *SA = malloc(sizeof(int32_t) * (size_t)n);
    int32_t *psv = malloc(sizeof(int32_t) * (size_t)n);
    int32_t *nsv = malloc(sizeof(int32_t) * (size_t)n);
    int32_t *cand = malloc(sizeof(int32_t) * 4 * (size_t)n);
    int32_t *parse = malloc(sizeof(int32_t) * 3 * (size_t)n);
    if (!SA || !psv || !nsv || !cand || !parse)
        goto out;
    if (oracle_suffix_array(src, n, SA) != 0)
        goto out;
    if (SA_out)
        memcpy(SA_out, SA, sizeof(int32_t) * (size_t)n);
    if (PLCP_out)
        oracle_plcp_lcp(src, n, SA, PLCP_out, LCP_out);
    oracle_psvnsv(SA, n, psv, nsv);
    oracle_candidates(src, n, psv, nsv, cand);
    if (cand_out)
        memcpy(cand_out, cand, sizeof(int32_t) * 4 * (size_t)n);
    oracle_parse(cand, n, parse);
    if (parse_out)
        memcpy(parse_out, parse, sizeof(int32_t) * 3 * (size_t)n);
    rc = 0;
    if (dst)
        rc = oracle_emit(src, N, parse, dst, dst_len);
out:
    free(SA); free(psv); free(nsv); free(cand); free(parse);
    return rc;
}

int oracle_encode(const uint8_t *src, size_t N, uint8_t *dst, size_t *dst_len)
{
    if (src == NULL || dst == NULL || dst_len == NULL)
        return -1;
    return oracle_stages(src, N, NULL, NULL, NULL, NULL, NULL, dst, dst_len);
}

/* ------------------------------------------------------------------------
 * Decoder restatement (lib/salz.c:829-1228).  Header = native u32:
 * type in the top byte (0 PLAIN, 1 SALZ), payload length in the low 24 bits.
 * The SALZ payload is consumed until exhausted; bit words are fetched lazily
 * (8 bytes, little-endian, MSB-first) when the first bit is needed.
 * Differences from the reference that cannot be observed through the API:
 * factors are copied byte-wise instead of in 8-byte strides, so nothing is
 * ever written past dst_pos + factor_len.
 * ---------------------------------------------------------------------- */
struct reader {
    const uint8_t *src;
    size_t len, pos;
    uint64_t acc;
    unsigned have;
    int ok;
};

static unsigned get_bit(struct reader *r)
{
    if (r->have == 0) {
        if (r->pos + 8 > r->len) {
            r->ok = 0;
            return 0;
        }
        memcpy(&r->acc, r->src + r->pos, 8);
        r->pos += 8;
        r->have = 64;
    }
    unsigned b = (unsigned)(r->acc >> 63);
    r->acc <<= 1;
    r->have--;
    return b;
}

int oracle_decode(const uint8_t *src, size_t src_len, uint8_t *dst,
                  size_t *dst_len)
{
    if (src == NULL || dst == NULL || dst_len == NULL || src_len < 4)
        return -1;
    uint32_t hdr;
    memcpy(&hdr, src, 4);
    uint32_t type = hdr >> 24;
    size_t plen = hdr & 0xffffff;
    if (type >= 2 || plen > src_len - 4)
        return -1;
    size_t cap = *dst_len;
    if (type == 0) {
        if (plen > cap)
            return -1;
        memcpy(dst, src + 4, plen);
        *dst_len = plen;
        return 0;
    }
    struct reader r = { src + 4, plen, 0, 0, 0, 1 };
    size_t out = 0;
    while (r.pos < r.len) {
        unsigned tok = get_bit(&r);
        if (!r.ok)
            return -1;
        if (tok == 0) {
            if (r.pos >= r.len || out >= cap)
                return -1;
            dst[out++] = r.src[r.pos++];
            continue;
        }
        /* offset: vnibble (up to 11 nibbles) then one direct byte */
        uint32_t v = 0;
        for (int k = 0; k < 11; k++) {
            unsigned nib = 0;
            for (int b = 0; b < 4; b++)
                nib = (nib << 1) | get_bit(&r);
            if (!r.ok)
                return -1;
            v = (k == 0) ? (nib & 7) : (((v + 1) << 3) | (nib & 7));
            if (nib >= 8)
                break;
        }
        if (r.pos >= r.len)
            return -1;
        uint32_t offs = ((v << 8) | r.src[r.pos++]) + 1;
        /* length: GR3 */
        uint32_t q = 0;
        for (;;) {
            unsigned b = get_bit(&r);
            if (!r.ok)
                return -1;
            if (b)
                break;
            q++;
        }
        uint32_t low = 0;
        for (int b = 0; b < 3; b++)
            low = (low << 1) | get_bit(&r);
        if (!r.ok)
            return -1;
        size_t len = (size_t)((q << 3) | low) + 3;
        if (out + len > cap || offs > out)
            return -1;
        for (size_t i = 0; i < len; i++, out++)
            dst[out] = dst[out - offs];
    }
    *dst_len = out;
    return 0;
}
