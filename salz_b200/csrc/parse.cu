// parse.cu - PSV/NSV with match lengths, and the cost-optimal LZ parse.
//
//   k_ansv : for every SA slot r (value v = SA[r]) the nearest slots to the left /
//            right holding a smaller text position (ref lib/salz.c:471-490), and
//            the match lengths as range minima of LCP between them
//            (== lcp_cmp's result, ref lib/salz.c:492-538), written in TEXT order
//            as (p-psv, psv_len, p-nsv, nsv_len) exactly like the reference's aux
//            table after factorize() (ref lib/salz.c:540-560).
//   k_parse: right-to-left shortest-path DP with the reference's cost model and
//            strict '<' tie-break order literal, PSV, NSV (ref lib/salz.c:562-662).
#include "salz_dev.cuh"

namespace {

struct Pyr {
    const int32_t *sa[6];   // sa[0] = SA (as int32), sa[l] = level-l minima
    const int32_t *lcp[6];
    int top;                // coarsest level
    int logS;
};

// ---------------------------------------------------------------------------
// ANSV by walking the min-pyramids.  Level l entry j of block b lives at
// (b << (logS - 5l)) + j.  Padding entries hold INT_MAX in both pyramids.
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_ansv(Pyr py, const uint32_t *__restrict__ blkN, int4 *__restrict__ cand)
{
    const int logS = py.logS;
    const uint32_t Smask = (1u << logS) - 1;
    const uint32_t s = blockIdx.x * 256 + threadIdx.x;
    const uint32_t b = s >> logS, r = s & Smask;
    const uint32_t n = blkN[b] - 8;
    if (r >= n) return;

    const int32_t *sa0 = py.sa[0] + ((size_t)b << logS);
    const int32_t *lcp0 = py.lcp[0] + ((size_t)b << logS);
    const int32_t v = sa0[r];

    // level-0 accessors honour the n_b bound (slots >= n are uninitialised)
    auto SAat = [&](int lvl, uint32_t j) -> int32_t {
        if (lvl == 0) return (j < n) ? sa0[j] : 0x7fffffff;
        return py.sa[lvl][((size_t)b << (logS - 5 * lvl)) + j];
    };
    auto LCPat = [&](int lvl, uint32_t j) -> int32_t {
        if (lvl == 0) return (j < n) ? lcp0[j] : 0x7fffffff;
        return py.lcp[lvl][((size_t)b << (logS - 5 * lvl)) + j];
    };

    // ---- PSV: nearest smaller to the left, min LCP over (r', r] ----
    int32_t psv = -1, psvLen = 0;
    {
        int32_t m = lcp0[r];
        int lvl = 0;
        uint32_t pos = r;
        int fl = -1;
        uint32_t fj = 0;
        for (;;) {
            uint32_t start = pos & ~31u;
            for (uint32_t j = pos; j-- > start;) {
                if (SAat(lvl, j) < v) { fl = lvl; fj = j; break; }
                m = min(m, LCPat(lvl, j));
            }
            if (fl >= 0 || lvl == py.top) break;
            pos >>= 5;
            lvl++;
        }
        if (fl >= 0) {
            while (fl > 0) {
                fl--;
                uint32_t c0 = fj << 5;
                for (uint32_t c = c0 + 32; c-- > c0;) {
                    if (SAat(fl, c) < v) { fj = c; break; }
                    m = min(m, LCPat(fl, c));
                }
            }
            psv = sa0[fj];
            psvLen = m;
        }
    }
    // ---- NSV: nearest smaller to the right, min LCP over (r, r''] ----
    int32_t nsv = -1, nsvLen = 0;
    {
        int32_t m = 0x7fffffff;
        int lvl = 0;
        uint32_t pos = r;
        int fl = -1;
        uint32_t fj = 0;
        for (;;) {
            uint32_t levelSize = 1u << (logS - 5 * lvl);
            uint32_t end = min(pos | 31u, levelSize - 1);
            for (uint32_t j = pos + 1; j <= end; j++) {
                if (SAat(lvl, j) < v) { fl = lvl; fj = j; break; }
                m = min(m, LCPat(lvl, j));
            }
            if (fl >= 0 || lvl == py.top) break;
            pos >>= 5;
            lvl++;
        }
        if (fl >= 0) {
            while (fl > 0) {
                fl--;
                uint32_t c0 = fj << 5;
                for (uint32_t c = c0; c < c0 + 32; c++) {
                    if (SAat(fl, c) < v) { fj = c; break; }
                    m = min(m, LCPat(fl, c));
                }
            }
            nsv = sa0[fj];
            nsvLen = min(m, lcp0[fj]);
        }
    }
    int4 o;
    if (v == 0) {
        o = make_int4(-1, 1, -1, 1);  // position 0 is a forced literal (ref lib/salz.c:546-548)
    } else {
        o.x = v - psv;   // psv == -1 gives v + 1 like the reference
        o.y = psvLen;
        o.z = v - nsv;
        o.w = nsvLen;
    }
    cand[((size_t)b << logS) + (uint32_t)v] = o;
}

// ---------------------------------------------------------------------------
// Optimal parse.  One warp per block walks 32-position windows right to left.
// Inside a window, lane l owns position q - l (q = right end).  Costs of
// targets beyond the window are already final and are gathered in parallel;
// the literal chain inside the window is a prefix-min scan in (min,+); factor
// edges that land inside the window are relaxed to a fixpoint (at most ~10
// iterations because factors are >= 3 long).  Costs are exact, so the choice
// can be recomputed per lane with the reference's comparison order.
//
// Also produced: jump[p] = first position >= the end of p's PARSE_CHUNK that
// the chosen parse reaches from p (used by the token walk in emit.cu).
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(32) k_parse(const int4 *__restrict__ cand, const uint32_t *__restrict__ blkN,
                                              int logS, int32_t *cost, int2 *__restrict__ choice,
                                              uint32_t *__restrict__ jump)
{
    __shared__ uint32_t ring[PARSE_CHUNK];
    const uint32_t b = blockIdx.x;
    const int lane = threadIdx.x;
    const uint32_t N = blkN[b];
    const uint32_t n = N - 8;
    const size_t gb = (size_t)b << logS;
    const int4 *cd = cand + gb;
    int32_t *cs = cost + gb + b;  // n_b + 1 entries per block (stride S + 1)
    int2 *ch = choice + gb;
    uint32_t *jp = jump + gb;

    int32_t costRight = SALZ_INF;  // cost[q+1] of the window being processed
    const uint32_t lastWin = (N - 1) >> 5;
    for (uint32_t w = lastWin + 1; w-- > 0;) {
        const uint32_t q = (w << 5) + 31;
        const uint32_t p = q - lane;
        const uint32_t chunkEnd = ((p >> PARSE_CHUNK_LOG) + 1) << PARSE_CHUNK_LOG;  // same for all lanes

        // --- candidates and far costs ---
        int32_t a1 = SALZ_INF, a2 = SALZ_INF;     // edge weights
        uint32_t l1 = 0, l2 = 0, o1 = 0, o2 = 0;
        int32_t f1 = SALZ_INF, f2 = SALZ_INF;     // edge weight + far target cost
        bool near1 = false, near2 = false;
        int32_t x;                                // best non-literal, non-near option so far
        if (p >= 1 && p < n) {
            int4 c = cd[p];
            if (c.y >= 3) {
                o1 = (uint32_t)c.x; l1 = (uint32_t)c.y;
                a1 = factor_cost_bits(o1, l1);
                if (l1 <= (uint32_t)lane) near1 = true;
                else f1 = a1 + __ldcg(&cs[p + l1]);
            }
            if (c.w >= 3) {
                o2 = (uint32_t)c.z; l2 = (uint32_t)c.w;
                a2 = factor_cost_bits(o2, l2);
                if (l2 <= (uint32_t)lane) near2 = true;
                else f2 = a2 + __ldcg(&cs[p + l2]);
            }
        }
        if (p == n) x = 0;                    // cost[n] = 0 (ref lib/salz.c:621)
        else x = min(f1, f2);                 // INF for p > n and p == 0
        const int32_t xfar = x;

        // --- fixpoint over in-window edges ---
        int32_t cst;
        const unsigned anyNear = __ballot_sync(0xffffffffu, near1 || near2);
        for (;;) {
            // cost[l] = min( min_{k<=l} x[k] + 9(l-k), costRight + 9(l+1) )
            int32_t y = x - 9 * lane;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                int32_t t = __shfl_up_sync(0xffffffffu, y, o);
                if (lane >= o) y = min(y, t);
            }
            cst = min(y + 9 * lane, costRight + 9 * (lane + 1));
            if (p > n) cst = SALZ_INF;
            if (!anyNear) break;
            int32_t t1 = __shfl_sync(0xffffffffu, cst, near1 ? (lane - (int)l1) : 0);
            int32_t t2 = __shfl_sync(0xffffffffu, cst, near2 ? (lane - (int)l2) : 0);
            int32_t nx = xfar;
            if (near1) nx = min(nx, a1 + t1);
            if (near2) nx = min(nx, a2 + t2);
            bool changed = nx < x;
            x = min(x, nx);
            if (!__any_sync(0xffffffffu, changed)) break;
        }

        // --- choice with the reference's tie-break ---
        int32_t right = __shfl_up_sync(0xffffffffu, cst, 1);
        if (lane == 0) right = costRight;
        int32_t t1 = __shfl_sync(0xffffffffu, cst, near1 ? (lane - (int)l1) : 0);
        int32_t t2 = __shfl_sync(0xffffffffu, cst, near2 ? (lane - (int)l2) : 0);
        int32_t c1 = near1 ? a1 + t1 : f1;
        int32_t c2 = near2 ? a2 + t2 : f2;
        int32_t best = 9 + right;
        uint32_t bo = 0, bl = 1;
        if (l1 >= 3 && c1 < best) { best = c1; bo = o1; bl = l1; }
        if (l2 >= 3 && c2 < best) { best = c2; bo = o2; bl = l2; }
        if (p == 0 || p >= n) { bo = 0; bl = 1; }  // forced literals: position 0 and the 8-byte tail

        // --- chunk exit pointers ---
        uint32_t t = p + bl;  // next token start
#pragma unroll
        for (int s = 0; s < 5; s++) {
            bool inwin = (t <= q) && (t < N);
            uint32_t tt = __shfl_sync(0xffffffffu, t, inwin ? (int)(q - t) : 0);
            if (inwin) t = tt;
        }
        uint32_t j = t;
        if (t < chunkEnd && t < N) j = ring[t & (PARSE_CHUNK - 1)];
        __syncwarp();
        ring[p & (PARSE_CHUNK - 1)] = j;
        __syncwarp();

        if (p < N) {
            if (p <= n) cs[p] = cst;
            ch[p] = make_int2((int)bo, (int)bl);
            jp[p] = j;
        }
        costRight = __shfl_sync(0xffffffffu, cst, 31);
        __syncwarp();
    }
}

}  // namespace

int ansv_build(Workspace &ws, const BatchDesc &bd, cudaStream_t st, LaunchCounter &lc)
{
    Pyr py;
    py.logS = bd.logS;
    py.sa[0] = (const int32_t *)ws.SA;
    py.lcp[0] = ws.lcp;
    int top = 0;
    while ((bd.S >> (5 * top)) > 32) top++;
    for (int l = 1; l < 6; l++) {
        py.sa[l] = ws.pyrSA[l];
        py.lcp[l] = ws.pyrLCP[l];
    }
    py.top = top;
    const size_t total = (size_t)bd.B * bd.S;
    k_ansv<<<(uint32_t)(total / 256), 256, 0, st>>>(py, ws.blkN, ws.cand);
    lc.launches += 1;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        fprintf(stderr, "salz_cuda: ansv_build: %s\n", cudaGetErrorString(e));
        return -1;
    }
    return 0;
}

int parse_build(Workspace &ws, const BatchDesc &bd, cudaStream_t st, LaunchCounter &lc)
{
    k_parse<<<bd.B, 32, 0, st>>>(ws.cand, ws.blkN, bd.logS, ws.cost, ws.choice, ws.jump);
    lc.launches += 1;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        fprintf(stderr, "salz_cuda: parse_build: %s\n", cudaGetErrorString(e));
        return -1;
    }
    return 0;
}
