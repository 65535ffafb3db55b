// ansv.cu - PSV/NSV with match lengths (all nearest smaller values over the
// suffix array, values = text positions).
//
// For every SA slot r (value v = SA[r]): the nearest slots to the left / right
// holding a smaller text position (ref lib/salz.c:471-490) and the match
// lengths as range minima of LCP between them (== lcp_cmp's result,
// ref lib/salz.c:492-538), written in TEXT order as
// (p-psv, psv_len, p-nsv, nsv_len), exactly the reference's aux table after
// factorize() (ref lib/salz.c:540-560).
//
// Three levels, each resolving what the previous one could not:
//   1. k_ansv_tile, segments: one thread runs the sequential stack algorithm
//      over 32 consecutive slots held in shared memory.  The stack is threaded
//      through the elements themselves (every element remembers the index of its
//      in-segment PSV), and carries the running LCP minimum, so PSV, NSV and both
//      lengths fall out of the same pass.  Unresolved to the left: the segment's
//      prefix minima; to the right: its suffix minima (what is left on the stack).
//   2. k_ansv_tile, tile: those elements search the neighbouring segments of the
//      4096-slot tile: nearest segment whose minimum is smaller, then the
//      suffix-minima (left search) / prefix-minima (right search) chain inside it.
//   3. k_ansv_far: the few elements whose answer lies outside their tile walk the
//      32-ary min pyramids over SA and LCP built by lcp.cu.
#include "salz_dev.cuh"

namespace {

struct Pyr {
    const int32_t *sa[6];   // sa[0] = SA (as int32), sa[l] = level-l minima
    const int32_t *lcp[6];
    int top;                // coarsest level
    int logS;
};

#ifndef SALZ_AL
#define SALZ_AL 32
#endif
constexpr int AL = SALZ_AL;         // slots per segment (one thread): 32 or 16
constexpr int AT = 4096 / AL;       // threads = segments per tile
constexpr int APR = 32 / AL;        // segments per group of the level-1 pyramid (32 slots)
constexpr int ATS = AL * AT;        // slots per tile
constexpr int APAD = AL + 1;        // padded segment stride in shared memory
constexpr int32_t AINF = 0x7fffffff;
constexpr uint32_t LINK_PM = 0x80;  // element is a prefix minimum; low 7 bits = next prefix minimum
constexpr uint32_t LINK_NONE = 0x7f;

// far item: slot (30 bits) | side (1 bit, 1 = right) | partial LCP minimum (31 bits)
__device__ __forceinline__ uint64_t far_item(uint32_t slot, uint32_t right, int32_t acc)
{
    return ((uint64_t)slot << 32) | ((uint64_t)right << 31) | (uint32_t)acc;
}

__global__ void __launch_bounds__(AT) k_ansv_tile(const uint32_t *__restrict__ SA, const int32_t *__restrict__ LCP,
                                                  const uint32_t *__restrict__ blkN, int logS,
                                                  int2 *__restrict__ cand2, uint64_t *__restrict__ farItems,
                                                  uint32_t *__restrict__ farCount, int32_t *__restrict__ tileMinV,
                                                  int32_t *__restrict__ tileMinL, int32_t *__restrict__ pyrSA1,
                                                  int32_t *__restrict__ pyrLCP1)
{
    __shared__ uint32_t sv[AT * APAD];   // (value << 8) | link
    __shared__ int32_t sm[AT * APAD];    // LCP, then the element's running minimum
    __shared__ int32_t segMin[AT], segLcp[AT];
    __shared__ int32_t preV[AT], preL[AT], sufV[AT], sufL[AT];  // minima over the segments before / after
    __shared__ uint8_t segLen[AT];

    const int tid = threadIdx.x;
    const uint32_t tilesPerBlock = 1u << (logS - 12);
    const uint32_t b = blockIdx.x / tilesPerBlock, tb = blockIdx.x % tilesPerBlock;
    const uint32_t n = blkN[b] - 8;
    const uint32_t t0 = tb * ATS;  // first slot of the tile (local)
    const size_t gb = (size_t)b << logS;
    const size_t pyr1 = ((gb + t0) >> 5) + tid / APR;  // APR segments make one group of the level-1 pyramid
    if (t0 >= n) {
        if (tid == 0) { tileMinV[blockIdx.x] = AINF; tileMinL[blockIdx.x] = AINF; }
        if (tid % APR == 0) {
            pyrSA1[pyr1] = AINF;
            pyrLCP1[pyr1] = AINF;
        }
        return;
    }
    int2 *cd = cand2 + 2 * gb;
    int4 *cd4 = (int4 *)cd;
    // candidates leave as 8- and 16-byte pieces of 32-byte sectors, in text order: they are
    // kept in L2 (evict_last) until the other pieces arrive; SA and LCP stream through
    const uint64_t keep = l2_evict_last(), stream = l2_evict_first();

    // tile load, eight slots per thread at a time: the loads of a batch are all in flight before
    // the first of them is used (one load per loop iteration exposed the DRAM latency 32 times:
    // 23 % of this kernel's stall samples sat on the first use, profiles/r01_v15_ncu_stalls.txt)
    constexpr int LDB = 8;
#pragma unroll 1
    for (int base = 0; base < AL; base += LDB) {
        uint32_t a[LDB];
        int32_t l[LDB];
#pragma unroll
        for (int u = 0; u < LDB; u++) {
            const uint32_t r = t0 + (uint32_t)((base + u) * AT + tid);
            a[u] = (r < n) ? ld_stream(SA + gb + r, stream) : 0u;
            l[u] = (r < n) ? ld_stream(LCP + gb + r, stream) : 0;
        }
#pragma unroll
        for (int u = 0; u < LDB; u++) {
            const int j = (base + u) * AT + tid;
            const int k = (j / AL) * APAD + (j % AL);
            sv[k] = a[u] << 8;
            sm[k] = l[u];
        }
    }
    __syncthreads();

    // ---- level 1: sequential stack over my segment ----
    const int sb = tid * APAD;
    const int len = (int)min((uint32_t)AL, (t0 + tid * AL < n) ? n - (t0 + tid * AL) : 0u);
    {
        int top = -1, lastPM = -1;
        int32_t topVal = 0, lmin = AINF;
        for (int i = 0; i < len; i++) {
            const int32_t v = (int32_t)(sv[sb + i] >> 8);
            int32_t acc = sm[sb + i];
            lmin = min(lmin, acc);
            while (top >= 0 && topVal > v) {       // i is the NSV of top
                const uint32_t lk = sv[sb + top] & 0xffu;
                const int32_t tm = sm[sb + top];   // min LCP over (PSV of top, top]
                if (lk & LINK_PM) {                // its PSV lies outside the segment: NSV half only
                    st_hint(&cd[2 * topVal + 1], make_int2(topVal - v, acc), keep);
                    top = -1;
                } else {                           // both known: one 16-byte store instead of two 8-byte ones
                    const int32_t pv = (int32_t)(sv[sb + (int)lk] >> 8);
                    st_hint(&cd4[topVal], make_int4(topVal - pv, tm, topVal - v, acc), keep);
                    top = (int)lk;
                    topVal = pv;
                }
                acc = min(acc, tm);
            }
            sm[sb + i] = acc;
            if (top >= 0) {                        // top is the PSV of i: written when i leaves the stack
                sv[sb + i] = ((uint32_t)v << 8) | (uint32_t)top;
            } else {                               // prefix minimum: PSV lies further left
                sv[sb + i] = ((uint32_t)v << 8) | LINK_PM | LINK_NONE;
                if (lastPM >= 0) sv[sb + lastPM] = (sv[sb + lastPM] & ~0x7fu) | (uint32_t)i;
                lastPM = i;
            }
            top = i;
            topVal = v;
        }
        segLen[tid] = (uint8_t)len;
        segMin[tid] = (lastPM >= 0) ? (int32_t)(sv[sb + lastPM] >> 8) : AINF;
        segLcp[tid] = lmin;
        // minima over 32 consecutive SA slots (AINF past n): level 1 of the pyramids k_ansv_far walks
        int32_t gv = segMin[tid], gl = lmin;
#pragma unroll
        for (int o = 1; o < APR; o <<= 1) {
            gv = min(gv, __shfl_xor_sync(0xffffffffu, gv, o));
            gl = min(gl, __shfl_xor_sync(0xffffffffu, gl, o));
        }
        if (tid % APR == 0) {
            pyrSA1[pyr1] = gv;
            pyrLCP1[pyr1] = gl;
        }
    }
    __syncthreads();
    // exclusive prefix / suffix minima over the tile's segments: an element smaller
    // than everything on one side skips the search (monotone SA: every element)
    if (tid < 32) {
        constexpr int K = AT / 32;   // segments per lane
        int32_t v[K], l[K];
        int32_t av = AINF, al = AINF;
#pragma unroll
        for (int k = 0; k < K; k++) {
            v[k] = segMin[tid * K + k]; l[k] = segLcp[tid * K + k];
            av = min(av, v[k]); al = min(al, l[k]);
        }
        int32_t pv = av, pl = al, sv_ = av, sl = al;  // inclusive scans over lanes
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int32_t a = __shfl_up_sync(0xffffffffu, pv, o), b2 = __shfl_up_sync(0xffffffffu, pl, o);
            int32_t c = __shfl_down_sync(0xffffffffu, sv_, o), d = __shfl_down_sync(0xffffffffu, sl, o);
            if (tid >= o) { pv = min(pv, a); pl = min(pl, b2); }
            if (tid + o < 32) { sv_ = min(sv_, c); sl = min(sl, d); }
        }
        if (tid == 31) { tileMinV[blockIdx.x] = pv; tileMinL[blockIdx.x] = pl; }
        int32_t ev = __shfl_up_sync(0xffffffffu, pv, 1), el = __shfl_up_sync(0xffffffffu, pl, 1);
        if (tid == 0) { ev = AINF; el = AINF; }
#pragma unroll
        for (int k = 0; k < K; k++) {
            preV[tid * K + k] = ev; preL[tid * K + k] = el;
            ev = min(ev, v[k]); el = min(el, l[k]);
        }
        ev = __shfl_down_sync(0xffffffffu, sv_, 1); el = __shfl_down_sync(0xffffffffu, sl, 1);
        if (tid == 31) { ev = AINF; el = AINF; }
#pragma unroll
        for (int k = K - 1; k >= 0; k--) {
            sufV[tid * K + k] = ev; sufL[tid * K + k] = el;
            ev = min(ev, v[k]); el = min(el, l[k]);
        }
    }
    __syncthreads();
    if (len == 0) return;

    const bool firstTile = (tb == 0), lastTile = (t0 + ATS >= n);
    // ---- level 2, left: my prefix minima look for their PSV in earlier segments ----
    for (int i = 0; i >= 0;) {
        const uint32_t w = sv[sb + i];
        const int32_t v = (int32_t)(w >> 8);
        int32_t acc = sm[sb + i];
        int s = tid - 1;
        if (v < preV[tid]) {
            acc = min(acc, preL[tid]);
            s = -1;
        } else {
            for (; s >= 0; s--) {
                if (segMin[s] < v) break;
                acc = min(acc, segLcp[s]);
            }
        }
        if (s >= 0) {
            int e = s * APAD + segLen[s] - 1;      // top of that segment's final stack
            while ((int32_t)(sv[e] >> 8) > v) {
                acc = min(acc, sm[e]);
                e = s * APAD + (int)(sv[e] & 0x7fu);
            }
            st_hint(&cd[2 * v], make_int2(v - (int32_t)(sv[e] >> 8), acc), keep);
        } else if (firstTile) {
            st_hint(&cd[2 * v], make_int2(v + 1, 0), keep);       // no PSV: offs = p - (-1), len 0 (ref lib/salz.c:546-560)
        } else {
            farItems[atomicAdd(farCount, 1u)] = far_item((uint32_t)gb + t0 + tid * AL + i, 0, acc);
        }
        const uint32_t nx = w & 0x7fu;
        i = (nx == LINK_NONE) ? -1 : (int)nx;
    }
    // ---- level 2, right: what is left on my stack looks for its NSV in later segments ----
    {
        int32_t accR = AINF;  // min LCP over (e, end of segment)
        for (int e = len - 1; e >= 0;) {
            const uint32_t w = sv[sb + e];
            const int32_t v = (int32_t)(w >> 8);
            int32_t acc = accR;
            int s = tid + 1;
            if (v < sufV[tid]) {
                acc = min(acc, sufL[tid]);
                s = AT;
            } else {
                for (; s < AT; s++) {
                    if (segMin[s] < v) break;
                    acc = min(acc, segLcp[s]);
                }
            }
            // what is still on the stack has not had its in-segment PSV written yet
            const bool havePsv = !(w & LINK_PM);
            const int2 ps = havePsv ? make_int2(v - (int32_t)(sv[sb + (int)(w & 0x7fu)] >> 8), sm[sb + e]) : make_int2(0, 0);
            if (s < AT) {
                int f = s * APAD;                  // first prefix minimum of that segment
                while ((int32_t)(sv[f] >> 8) > v) f = s * APAD + (int)(sv[f] & 0x7fu);
                const int2 ns = make_int2(v - (int32_t)(sv[f] >> 8), min(acc, sm[f]));
                if (havePsv) st_hint(&cd4[v], make_int4(ps.x, ps.y, ns.x, ns.y), keep);
                else st_hint(&cd[2 * v + 1], ns, keep);
            } else if (lastTile) {
                if (havePsv) st_hint(&cd4[v], make_int4(ps.x, ps.y, v + 1, 0), keep);
                else st_hint(&cd[2 * v + 1], make_int2(v + 1, 0), keep);
            } else {
                if (havePsv) st_hint(&cd[2 * v], ps, keep);
                farItems[atomicAdd(farCount, 1u)] = far_item((uint32_t)gb + t0 + tid * AL + e, 1, acc);
            }
            accR = min(accR, sm[sb + e]);
            e = (w & LINK_PM) ? -1 : (int)(w & 0x7fu);
        }
    }
}

// ---------------------------------------------------------------------------
// Level 3: walk the min pyramids from the tile boundary.  Level l entry j of
// block b lives at (b << (logS - 5l)) + j.  Padding entries hold INT_MAX.
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_ansv_far(Pyr py, const uint32_t *__restrict__ blkN,
                                                  const uint64_t *__restrict__ farItems,
                                                  const uint32_t *__restrict__ farCount, int2 *__restrict__ cand2,
                                                  const int32_t *__restrict__ tpreV, const int32_t *__restrict__ tsufV)
{
    const int logS = py.logS;
    const uint32_t Smask = (1u << logS) - 1;
    const uint32_t count = *farCount;
    for (uint32_t it = blockIdx.x * 256 + threadIdx.x; it < count; it += gridDim.x * 256) {
        const uint64_t item = farItems[it];
        const uint32_t slot = (uint32_t)(item >> 32);
        const bool right = (item >> 31) & 1;
        int32_t m = (int32_t)(item & 0x7fffffffu);
        const uint32_t b = slot >> logS, r = slot & Smask;
        const uint32_t n = blkN[b] - 8;
        const int32_t *sa0 = py.sa[0] + ((size_t)b << logS);
        const int32_t *lcp0 = py.lcp[0] + ((size_t)b << logS);
        const int32_t v = sa0[r];
        auto SAat = [&](int lvl, uint32_t j) -> int32_t {
            if (lvl == 0) return (j < n) ? sa0[j] : AINF;
            return py.sa[lvl][((size_t)b << (logS - 5 * lvl)) + j];
        };
        auto LCPat = [&](int lvl, uint32_t j) -> int32_t {
            if (lvl == 0) return (j < n) ? lcp0[j] : AINF;
            return py.lcp[lvl][((size_t)b << (logS - 5 * lvl)) + j];
        };
        int fl = -1;
        uint32_t fj = 0;
        int2 o = make_int2(v + 1, 0);  // none found
        const uint32_t tileG = slot / ATS;
        if (!right && v < tpreV[tileG]) {
            cand2[2 * (((size_t)b << logS) + (uint32_t)v)] = o;      // smaller than every slot before the tile
        } else if (right && v < tsufV[tileG]) {
            cand2[2 * (((size_t)b << logS) + (uint32_t)v) + 1] = o;  // smaller than every slot after the tile
        } else if (!right) {
            // scan slots < tile start; m already covers (tile start - 1, r]
            int lvl = 0;
            uint32_t pos = r & ~(uint32_t)(ATS - 1);
            for (;;) {
                const uint32_t start = pos & ~31u;
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
                    const uint32_t c0 = fj << 5;
                    for (uint32_t c = c0 + 32; c-- > c0;) {
                        if (SAat(fl, c) < v) { fj = c; break; }
                        m = min(m, LCPat(fl, c));
                    }
                }
                o = make_int2(v - sa0[fj], m);
            }
            cand2[2 * (((size_t)b << logS) + (uint32_t)v)] = o;
        } else {
            // scan slots >= tile end; m already covers (r, tile end)
            int lvl = 0;
            uint32_t pos = r | (uint32_t)(ATS - 1);
            for (;;) {
                const uint32_t levelSize = 1u << (logS - 5 * lvl);
                const uint32_t end = min(pos | 31u, levelSize - 1);
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
                    const uint32_t c0 = fj << 5;
                    for (uint32_t c = c0; c < c0 + 32; c++) {
                        if (SAat(fl, c) < v) { fj = c; break; }
                        m = min(m, LCPat(fl, c));
                    }
                }
                o = make_int2(v - sa0[fj], min(m, lcp0[fj]));
            }
            cand2[2 * (((size_t)b << logS) + (uint32_t)v) + 1] = o;
        }
    }
}

// Per block: exclusive prefix / suffix minima of the tile minima (values only).
__global__ void __launch_bounds__(32) k_ansv_tilescan(const int32_t *__restrict__ tileMinV, uint32_t tilesPerBlock,
                                                      int32_t *__restrict__ tpreV, int32_t *__restrict__ tsufV)
{
    const size_t o = (size_t)blockIdx.x * tilesPerBlock;
    if (threadIdx.x == 0) {
        int32_t m = AINF;
        for (uint32_t t = 0; t < tilesPerBlock; t++) { tpreV[o + t] = m; m = min(m, tileMinV[o + t]); }
    } else if (threadIdx.x == 1) {
        int32_t m = AINF;
        for (uint32_t t = tilesPerBlock; t-- > 0;) { tsufV[o + t] = m; m = min(m, tileMinV[o + t]); }
    }
}

// position 0 is a forced literal (ref lib/salz.c:546-548)
__global__ void k_cand0(int4 *__restrict__ cand, int logS, int B)
{
    int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b < B) cand[(size_t)b << logS] = make_int4(-1, 1, -1, 1);
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
    uint32_t *farCount = ws.counters + 8;
    uint64_t *farItems = ws.keys[0];  // the sort buffers are free once the suffix array is complete
    cudaMemsetAsync(farCount, 0, 4, st);
    k_ansv_tile<<<(uint32_t)(total / ATS), AT, 0, st>>>(ws.SA, ws.lcp, ws.blkN, bd.logS, (int2 *)ws.cand, farItems,
                                                        farCount, ws.tileMinV, ws.tileMinL, ws.pyrSA[1], ws.pyrLCP[1]);
    if (pyramid_upper_build(ws, bd, st, lc)) return -1;
    k_ansv_tilescan<<<bd.B, 32, 0, st>>>(ws.tileMinV, bd.S / ATS, ws.tilePreV, ws.tileSufV);
    k_ansv_far<<<148 * 8, 256, 0, st>>>(py, ws.blkN, farItems, farCount, (int2 *)ws.cand, ws.tilePreV, ws.tileSufV);
    k_cand0<<<(bd.B + 127) / 128, 128, 0, st>>>(ws.cand, bd.logS, bd.B);
    lc.launches += 4;
    lc.alg_bytes += (uint64_t)bd.M * 24;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        fprintf(stderr, "salz_cuda: ansv_build: %s\n", cudaGetErrorString(e));
        return -1;
    }
    return 0;
}
