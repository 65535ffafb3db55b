// sa.cu - batched suffix-array construction by prefix doubling.
//
// Replaces the single libsais() call of the reference (ref lib/salz.c:463-469,
// 3rdparty/libsais/src/libsais.c:6597-6611).  The suffix array is unique, so
// any exact suffix sorter yields the same bytes downstream.
//
// All blocks of a batch are sorted by the same launches, but every block is
// its own segment of the sort list: tiles never straddle blocks, digit
// histograms and look-back chains are per block, so no block-id bits travel
// in the keys.  The list holds the suffixes whose final position is not yet
// known; it is always ordered by group, and a group id is the SA slot of the
// group's first member.  One round:
//
//   k_build    key = (group id, 1 + group id of suffix p+h, 0 past the end)
//   k_scatter  stable LSD radix passes over the key bits (8 bits per pass):
//              per-warp shared-memory digit counters, ranks from warp-wide
//              peer masks (shared-memory atomicOr, or match_any), decoupled
//              look-back over per-tile digit counts, keys staged through
//              shared memory in digit order and written back as coalesced runs
//   k_rerank_* adjacent-key compare + scan: new group ids; suffixes whose
//              group became a singleton are written to SA and dropped
//
// Element formats.  PACKED (block capacity S <= 2^21): one u64 per element,
//   (group:logS | second:logS | p:logS); a pass moves 16 B per element.
// Otherwise: u64 key (group | second) + u32 value p; 24 B per element.
#include "salz_dev.cuh"

#include <algorithm>

namespace {

#ifndef SALZ_TT
#define SALZ_TT 384
#endif
#ifndef SALZ_ITEMS
#define SALZ_ITEMS 12
#endif
constexpr int TT = SALZ_TT;         // threads per tile CTA (>= 256)
constexpr int ITEMS = SALZ_ITEMS;
constexpr int TILE = TT * ITEMS;    // elements per tile
constexpr int WARPS = TT / 32;
constexpr int MAXP = 8;             // radix passes per sort, at most
#ifndef SALZ_LBW
#define SALZ_LBW 8
#endif
constexpr int LBW = SALZ_LBW;            // look-back window (predecessor tiles fetched per round trip)

constexpr uint32_t LB_AGG = 1u << 30;   // tile aggregate published
constexpr uint32_t LB_INCL = 2u << 30;  // inclusive prefix published
constexpr uint32_t LB_MASK = (1u << 30) - 1;

__device__ __forceinline__ uint32_t ld_relaxed(const uint32_t *p)
{
    uint32_t v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed(uint32_t *p, uint32_t v)
{
    asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// The sort list of one round: per-block segments cut into tiles.
struct ListView {
    const uint32_t *cnt;        // [B]   elements of block b in the list
    const uint32_t *base;       // [B+1] list offset of block b
    const uint32_t *tileBase;   // [B+1] first tile of block b
    const uint4 *tileDesc;      // [tiles] (block, tile index inside the block, first element, elements)
};

struct TileLoc { uint32_t b, lt, first, valid; };

__device__ __forceinline__ TileLoc locate(const ListView &lv, uint32_t tile)
{
    const uint4 d = lv.tileDesc[tile];
    TileLoc t;
    t.b = d.x;
    t.lt = d.y;
    t.first = d.z;
    t.valid = d.w;
    return t;
}

// Exclusive sum scan over the values held by threads 0..255 (others pass 0).
// Must be called by every thread of the CTA.  scratch: >= 9 uint32 in smem.
__device__ __forceinline__ uint32_t excl_scan_first256(uint32_t v, uint32_t *scratch)
{
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    uint32_t inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    if (warp < 8 && lane == 31) scratch[warp] = inc;
    __syncthreads();
    if (tid == 0) {
        uint32_t s = 0;
        for (int w = 0; w < 8; w++) { uint32_t c = scratch[w]; scratch[w] = s; s += c; }
    }
    __syncthreads();
    uint32_t r = (warp < 8) ? (scratch[warp] + inc - v) : 0;
    __syncthreads();
    return r;
}

// ---- list bookkeeping --------------------------------------------------------

// cnt[b] = n_b for the first round.
__global__ void k_first_counts(const uint32_t *__restrict__ blkN, int B, uint32_t *__restrict__ cnt)
{
    int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b < B) cnt[b] = blkN[b] - 8;
}

// Single CTA: base / tileBase = exclusive prefixes of cnt / tiles(cnt);
// counters[0] = list length, counters[1] = tiles.
__global__ void __launch_bounds__(1024) k_tilemap(const uint32_t *__restrict__ cnt, int B, uint32_t *__restrict__ base,
                                                  uint32_t *__restrict__ tileBase, uint32_t *__restrict__ counters)
{
    __shared__ uint32_t pc[1024], pt[1024];
    const int tid = threadIdx.x;
    const int per = (B + 1023) / 1024;
    const int lo = min(B, tid * per), hi = min(B, lo + per);
    uint32_t sc = 0, stl = 0;
    for (int b = lo; b < hi; b++) { uint32_t c = cnt[b]; sc += c; stl += (c + TILE - 1) / TILE; }
    pc[tid] = sc;
    pt[tid] = stl;
    __syncthreads();
    if (tid < 32) {  // 1024 partials: each lane scans 32 of them, then a warp scan
        uint32_t a = 0, t = 0;
        for (int i = 0; i < 32; i++) { a += pc[tid * 32 + i]; t += pt[tid * 32 + i]; }
        uint32_t ia = a, it = t;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t x = __shfl_up_sync(0xffffffffu, ia, o), y = __shfl_up_sync(0xffffffffu, it, o);
            if (tid >= o) { ia += x; it += y; }
        }
        uint32_t ra = ia - a, rt = it - t;
        for (int i = 0; i < 32; i++) {
            uint32_t x = pc[tid * 32 + i], y = pt[tid * 32 + i];
            pc[tid * 32 + i] = ra; pt[tid * 32 + i] = rt;
            ra += x; rt += y;
        }
        if (tid == 31) { counters[0] = ia; counters[1] = it; base[B] = ia; tileBase[B] = it; }
    }
    __syncthreads();
    sc = pc[tid];
    stl = pt[tid];
    for (int b = lo; b < hi; b++) {
        uint32_t c = cnt[b];
        base[b] = sc;
        tileBase[b] = stl;
        sc += c;
        stl += (c + TILE - 1) / TILE;
    }
}

__global__ void __launch_bounds__(128) k_tilefill(const uint32_t *__restrict__ cnt, const uint32_t *__restrict__ base,
                                                  const uint32_t *__restrict__ tileBase, uint4 *__restrict__ tileDesc)
{
    const uint32_t b = blockIdx.x;
    const uint32_t lo = tileBase[b], hi = tileBase[b + 1], c = cnt[b], first = base[b];
    for (uint32_t t = lo + threadIdx.x; t < hi; t += blockDim.x) {
        const uint32_t lt = t - lo;
        tileDesc[t] = make_uint4(b, lt, first + lt * TILE, min((uint32_t)TILE, c - lt * TILE));
    }
}

// ---- key construction -----------------------------------------------------------

__device__ __forceinline__ void hist_flush(const uint32_t *sh, int passes, uint32_t *hist, uint32_t b)
{
    for (int i = threadIdx.x; i < passes * 256; i += blockDim.x)
        if (sh[i]) atomicAdd(&hist[(size_t)b * MAXP * 256 + i], sh[i]);
}

// Round 0: key field = first `nchars` bytes of the suffix, big-endian, zero padded
// past n_b.  Elements of a block are emitted in DESCENDING text position so that
// the stable sort leaves zero-padded ties (only the last < nchars suffixes of a
// block can be involved) in their correct order: shorter suffix first.
template <bool PACKED>
__global__ void __launch_bounds__(TT) k_init(const uint8_t *__restrict__ T, size_t Tstride,
                                             const uint32_t *__restrict__ blkN, ListView lv, int logS, int nchars,
                                             uint64_t *__restrict__ keys, uint32_t *__restrict__ vals,
                                             uint32_t *__restrict__ hist)
{
    __shared__ uint32_t sh[MAXP * 256];
    for (int i = threadIdx.x; i < nchars * 256; i += TT) sh[i] = 0;
    __syncthreads();
    const TileLoc tl = locate(lv, blockIdx.x);
    const uint32_t n = blkN[tl.b] - 8;
    const uint8_t *Tb = T + (size_t)tl.b * Tstride;
#pragma unroll
    for (int i = 0; i < ITEMS; i++) {
        const uint32_t idx = threadIdx.x + i * TT;
        if (idx < tl.valid) {
            const uint32_t p = n - 1 - (tl.lt * TILE + idx);
            uint64_t le = load_u64_unaligned(Tb + p);
            uint64_t be = __byte_perm((uint32_t)(le >> 32), 0, 0x0123) |
                          ((uint64_t)__byte_perm((uint32_t)le, 0, 0x0123) << 32);
            const uint32_t have = n - p;  // bytes of the suffix
            if (have < 8) be &= ~0ull << (8 * (8 - have));
            const uint64_t kf = be >> (8 * (8 - nchars));
            if (PACKED) {
                keys[tl.first + idx] = (kf << logS) | p;
            } else {
                keys[tl.first + idx] = kf;
                vals[tl.first + idx] = p;
            }
            for (int ps = 0; ps < nchars; ps++) atomicAdd(&sh[ps * 256 + ((uint32_t)(kf >> (8 * ps)) & 255u)], 1u);
        }
    }
    __syncthreads();
    hist_flush(sh, nchars, hist, tl.b);
}

// Doubling round: the list holds (group, -, p); second = 1 + local group id of
// suffix p+h, 0 past the end.
template <bool PACKED>
__global__ void __launch_bounds__(TT) k_build(const uint32_t *__restrict__ rank, const uint32_t *__restrict__ blkN,
                                              ListView lv, int logS, uint32_t h, int passes,
                                              uint64_t *__restrict__ keys, const uint32_t *__restrict__ vals,
                                              uint32_t *__restrict__ hist)
{
    __shared__ uint32_t sh[MAXP * 256];
    for (int i = threadIdx.x; i < passes * 256; i += TT) sh[i] = 0;
    __syncthreads();
    const TileLoc tl = locate(lv, blockIdx.x);
    const uint32_t n = blkN[tl.b] - 8;
    const uint32_t Smask = (1u << logS) - 1;
    const uint32_t *rk = rank + ((size_t)tl.b << logS);
#pragma unroll
    for (int i = 0; i < ITEMS; i++) {
        const uint32_t idx = threadIdx.x + i * TT;
        if (idx < tl.valid) {
            const uint64_t e = keys[tl.first + idx];
            uint32_t p, prim;
            if (PACKED) { p = (uint32_t)e & Smask; prim = (uint32_t)(e >> (2 * logS)); }
            else { p = vals[tl.first + idx]; prim = (uint32_t)(e >> logS); }
            uint32_t sec = 0;
            if (p + h < n) sec = (rk[p + h] & Smask) + 1;
            const uint64_t kf = ((uint64_t)prim << logS) | sec;
            keys[tl.first + idx] = PACKED ? ((kf << logS) | p) : kf;
            for (int ps = 0; ps < passes; ps++) atomicAdd(&sh[ps * 256 + ((uint32_t)(kf >> (8 * ps)) & 255u)], 1u);
        }
    }
    __syncthreads();
    hist_flush(sh, passes, hist, tl.b);
}

// digitBase[b][pass][d] = exclusive scan of hist[b][pass][*].  grid = (passes, B)
__global__ void __launch_bounds__(256) k_scan_hist(const uint32_t *__restrict__ hist, uint32_t *__restrict__ digitBase)
{
    __shared__ uint32_t scratch[16];
    const size_t o = ((size_t)blockIdx.y * MAXP + blockIdx.x) * 256 + threadIdx.x;
    digitBase[o] = excl_scan_first256(hist[o], scratch);
}

// ---- the radix scatter pass -------------------------------------------------------
// Tile ids are handed out by an atomic ticket so that every predecessor in the
// look-back chain is already running.  tileState[0] is the ticket, the states
// start at tileState + 256.
constexpr size_t SCATTER_SMEM_PACKED = (size_t)TILE * 8 + 2 * (size_t)WARPS * 256 * 4 + 2 * 256 * 4 + 64;
constexpr size_t SCATTER_SMEM_PAIRS = SCATTER_SMEM_PACKED + (size_t)TILE * 4;
constexpr int SCATTER_MIN_CTAS = (TT * ITEMS <= 2048) ? 4 : 2;

template <bool PACKED, bool MATCH_ANY>
__global__ void __launch_bounds__(TT, SCATTER_MIN_CTAS)
k_scatter(const uint64_t *__restrict__ keysIn, const uint32_t *__restrict__ valsIn, uint64_t *__restrict__ keysOut,
          uint32_t *__restrict__ valsOut, ListView lv, uint32_t numTiles, int shift, int pass,
          const uint32_t *__restrict__ digitBase, uint32_t *tileState)
{
    extern __shared__ __align__(16) uint8_t smem[];
    uint64_t *skeys = (uint64_t *)smem;                 // [TILE] keys in digit order
    uint32_t *wcnt = (uint32_t *)(skeys + TILE);        // [WARPS][256] per-warp digit counts -> running offsets
    uint32_t *masks = wcnt + WARPS * 256;               // [WARPS][256] peer masks (self-cleaning)
    uint32_t *dstart = masks + WARPS * 256;             // [256]
    uint32_t *gbase = dstart + 256;                     // [256]
    uint32_t *scratch = gbase + 256;                    // [16]
    uint32_t *svals = scratch + 16;                     // [TILE] (pairs only)
    __shared__ uint32_t s_tile[2];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    uint32_t *state = tileState + 256;
    uint32_t *mycnt = wcnt + warp * 256;
    uint32_t *mymask = masks + warp * 256;
    const uint32_t lt_mask = (1u << lane) - 1u;

    // Persistent CTA: tiles are claimed by ticket (so every predecessor in a
    // look-back chain is already running), and the next tile's descriptor and
    // keys are fetched while the current tile is ranked and written.
    if (tid == 0) s_tile[0] = atomicAdd(tileState, 1u);
    for (int i = tid; i < WARPS * 256; i += TT) masks[i] = 0;
    __syncthreads();
    uint32_t tile = s_tile[0];
    int par = 0;
    TileLoc nl = {0, 0, 0, 0};
    uint64_t nkey[ITEMS];
    uint32_t nval[ITEMS];
    if (tile < numTiles) {
        nl = locate(lv, tile);
#pragma unroll
        for (int i = 0; i < ITEMS; i++) {
            const uint32_t idx = warp * (32 * ITEMS) + i * 32 + lane;
            nkey[i] = (idx < nl.valid) ? keysIn[nl.first + idx] : ~0ull;
            if (!PACKED) nval[i] = (idx < nl.valid) ? valsIn[nl.first + idx] : 0u;
        }
    }

    while (tile < numTiles) {
        const TileLoc tl = nl;
        uint64_t key[ITEMS];
        uint32_t val[ITEMS];
#pragma unroll
        for (int i = 0; i < ITEMS; i++) {
            key[i] = nkey[i];
            if (!PACKED) val[i] = nval[i];
        }
        for (int i = tid; i < WARPS * 256; i += TT) wcnt[i] = 0;
        if (tid == 0) s_tile[par ^ 1] = atomicAdd(tileState, 1u);
        __syncthreads();
        const uint32_t nextTile = s_tile[par ^ 1];
        uint4 nd = make_uint4(0, 0, 0, 0);
        if (nextTile < numTiles) nd = lv.tileDesc[nextTile];

        // early counts: the tile's digit histogram is published before the ranking so
        // that successors find an inclusive prefix by the time they look back
#pragma unroll
        for (int i = 0; i < ITEMS; i++) {
            const uint32_t d = (uint32_t)(key[i] >> shift) & 255u;
            // a warp-uniform digit (the rule in the passes over the group bits) would
            // serialise 32 same-address atomics: one lane adds 32 instead
            if (__all_sync(0xffffffffu, d == __shfl_sync(0xffffffffu, d, 0))) {
                if (lane == 0) mycnt[d] += 32;
            } else {
                atomicAdd(&mycnt[d], 1u);
            }
        }
        __syncthreads();

        uint32_t total = 0;
        if (tid < 256) {
            uint32_t sum = 0;
#pragma unroll
            for (int w = 0; w < WARPS; w++) {
                uint32_t c = wcnt[w * 256 + tid];
                wcnt[w * 256 + tid] = sum;
                sum += c;
            }
            total = sum;
            const uint32_t cnt = total - ((tid == 255) ? (TILE - tl.valid) : 0u);
            st_relaxed(&state[(size_t)tile * 256 + tid], (tl.lt == 0 ? LB_INCL : LB_AGG) | cnt);
        }
        const uint32_t ex = excl_scan_first256(total, scratch);
        if (tid < 256) dstart[tid] = ex;
        __syncthreads();

        // next tile's keys: in flight during ranking, look-back and write-out
        nl.b = nd.x; nl.lt = nd.y; nl.first = nd.z; nl.valid = nd.w;
#pragma unroll
        for (int i = 0; i < ITEMS; i++) {
            const uint32_t idx = warp * (32 * ITEMS) + i * 32 + lane;
            nkey[i] = (idx < nl.valid) ? keysIn[nl.first + idx] : ~0ull;
            if (!PACKED) nval[i] = (idx < nl.valid) ? valsIn[nl.first + idx] : 0u;
        }

        // ranking: peers of a lane = lanes of the warp holding the same digit
#pragma unroll
        for (int i = 0; i < ITEMS; i++) {
            const uint32_t d = (uint32_t)(key[i] >> shift) & 255u;
            uint32_t m;
            const bool uniform = __all_sync(0xffffffffu, d == __shfl_sync(0xffffffffu, d, 0));
            if (uniform) {
                m = 0xffffffffu;
            } else if (MATCH_ANY) {
                m = __match_any_sync(0xffffffffu, d);
            } else {
                atomicOr(&mymask[d], 1u << lane);
                __syncwarp();
                m = mymask[d];
                __syncwarp();
            }
            const int leader = __ffs(m) - 1;
            uint32_t pre = 0;
            if (lane == leader) {
                pre = mycnt[d];
                mycnt[d] = pre + __popc(m);
                if (!MATCH_ANY && !uniform) mymask[d] = 0;
            }
            pre = __shfl_sync(0xffffffffu, pre, leader);
            const uint32_t pos = dstart[d] + pre + __popc(m & lt_mask);
            skeys[pos] = key[i];
            if (!PACKED) svals[pos] = val[i];
            __syncwarp();
        }

        if (tid < 256) {
            uint32_t prev = 0;
            if (tl.lt != 0) {
                // windowed look-back: LBW predecessors are fetched at once so that a
                // chain of aggregates costs one L2 round trip, not one per hop
                uint32_t t = tile - 1;
                for (bool done = false; !done;) {
                    const uint32_t w = min((uint32_t)LBW, tl.lt - (tile - 1 - t));  // stay inside the block
                    uint32_t sv[LBW];
#pragma unroll
                    for (int k = 0; k < LBW; k++)
                        sv[k] = ((uint32_t)k < w) ? ld_relaxed(&state[(size_t)(t - k) * 256 + tid]) : 0u;
#pragma unroll
                    for (int k = 0; k < LBW; k++) {
                        if (done || (uint32_t)k >= w) break;
                        const uint32_t f = sv[k] >> 30;
                        if (f == 0) break;             // not published yet: poll again from here
                        prev += sv[k] & LB_MASK;
                        t--;
                        if (f == 2) done = true;
                    }
                }
                const uint32_t cnt = total - ((tid == 255) ? (TILE - tl.valid) : 0u);
                st_relaxed(&state[(size_t)tile * 256 + tid], LB_INCL | (prev + cnt));
            }
            gbase[tid] = lv.base[tl.b] + digitBase[((size_t)tl.b * MAXP + pass) * 256 + tid] + prev - ex;
        }
        __syncthreads();

        for (uint32_t j = tid; j < tl.valid; j += TT) {
            const uint64_t k = skeys[j];
            const uint32_t dst = gbase[(uint32_t)(k >> shift) & 255u] + j;
            keysOut[dst] = k;
            if (!PACKED) valsOut[dst] = svals[j];
        }
        tile = nextTile;
        par ^= 1;
    }
}

// ---- re-rank + compaction -------------------------------------------------------------
// Element k of a block's segment starts a new group ("head") when its key field
// differs from its predecessor's, or when it is the first of the segment.  From
// round 1 on a zero second key (suffix ends within h characters) is always a
// head: two different suffixes of the same group cannot both end inside the
// compared prefix.  "pst" marks the first element of an OLD group (group field
// changes); the SA slot of element k is  group(k) + (k - index of its pst),
// and the new group id is the slot of the element's head.

struct HK { uint32_t head, pst, keep; };  // head/pst: list index + 1 of the last flagged element (0 none)

__device__ __forceinline__ HK hk_combine(HK a, HK b)
{
    HK r;
    r.head = b.head ? b.head : a.head;
    r.pst = b.pst ? b.pst : a.pst;
    r.keep = a.keep + b.keep;
    return r;
}

__device__ __forceinline__ HK hk_block_excl(HK v, HK *agg, HK *scratch /*[WARPS+1]*/)
{
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    HK inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        HK t;
        t.head = __shfl_up_sync(0xffffffffu, inc.head, o);
        t.pst = __shfl_up_sync(0xffffffffu, inc.pst, o);
        t.keep = __shfl_up_sync(0xffffffffu, inc.keep, o);
        if (lane >= o) inc = hk_combine(t, inc);
    }
    if (lane == 31) scratch[warp] = inc;
    __syncthreads();
    if (tid == 0) {
        HK s = {0, 0, 0};
        for (int w = 0; w < WARPS; w++) {
            HK c = scratch[w];
            scratch[w] = s;
            s = hk_combine(s, c);
        }
        scratch[WARPS] = s;
    }
    __syncthreads();
    HK ex;
    ex.head = __shfl_up_sync(0xffffffffu, inc.head, 1);
    ex.pst = __shfl_up_sync(0xffffffffu, inc.pst, 1);
    ex.keep = __shfl_up_sync(0xffffffffu, inc.keep, 1);
    if (lane == 0) ex = HK{0, 0, 0};
    ex = hk_combine(scratch[warp], ex);
    *agg = scratch[WARPS];
    __syncthreads();
    return ex;
}

// Flags of the ITEMS elements a thread owns (consecutive), plus head of the next one.
template <bool PACKED, bool ROUND0>
__device__ __forceinline__ void load_flags(const uint64_t *__restrict__ keys, const TileLoc &tl, uint32_t segRemain,
                                           uint32_t i0, int logS, uint64_t e[ITEMS], bool head[ITEMS + 1],
                                           bool pst[ITEMS])
{
    // key field and group of an element
    const uint64_t secMask = ((uint64_t)1 << logS) - 1;
    auto kf = [&](uint64_t x) -> uint64_t { return PACKED ? (x >> logS) : x; };
    const bool havePrev = (tl.lt > 0 || i0 > 0) && i0 < tl.valid;
    uint64_t prev = havePrev ? kf(keys[tl.first + i0 - 1]) : 0;
#pragma unroll
    for (int i = 0; i <= ITEMS; i++) {
        const uint32_t idx = i0 + i;
        // idx may be one past the tile: still inside the segment when segRemain allows
        const bool exists = (i < ITEMS) ? (idx < tl.valid) : (idx < tl.valid || (idx == TILE && segRemain > TILE));
        if (exists) {
            const uint64_t raw = keys[tl.first + idx];
            const uint64_t cur = kf(raw);
            const bool first = (tl.lt == 0 && idx == 0);
            bool p = first, hd = first || prev != cur;
            if (!ROUND0) {
                p = p || (prev >> logS) != (cur >> logS);
                hd = hd || (cur & secMask) == 0;
            }
            head[i] = hd;
            if (i < ITEMS) { pst[i] = p; e[i] = raw; }
            prev = cur;
        } else {
            head[i] = true;  // end of the segment closes the last group
            if (i < ITEMS) { pst[i] = false; e[i] = 0; }
        }
    }
}

template <bool PACKED, bool ROUND0>
__global__ void __launch_bounds__(TT) k_rerank_reduce(const uint64_t *__restrict__ keys, ListView lv, int logS,
                                                      uint32_t *__restrict__ aggHead, uint32_t *__restrict__ aggPst,
                                                      uint32_t *__restrict__ aggKeep)
{
    __shared__ HK scratch[WARPS + 1];
    const TileLoc tl = locate(lv, blockIdx.x);
    const uint32_t segRemain = lv.cnt[tl.b] - tl.lt * TILE;
    const uint32_t i0 = threadIdx.x * ITEMS;
    uint64_t e[ITEMS];
    bool head[ITEMS + 1], pst[ITEMS];
    load_flags<PACKED, ROUND0>(keys, tl, segRemain, i0, logS, e, head, pst);
    HK v = {0, 0, 0};
#pragma unroll
    for (int i = 0; i < ITEMS; i++) {
        if (i0 + i < tl.valid) {
            const uint32_t k = tl.first + i0 + i;
            if (head[i]) v.head = k + 1;
            if (pst[i]) v.pst = k + 1;
            if (!(head[i] && head[i + 1])) v.keep++;
        }
    }
    HK agg;
    hk_block_excl(v, &agg, scratch);
    if (threadIdx.x == 0) {
        aggHead[blockIdx.x] = agg.head;
        aggPst[blockIdx.x] = agg.pst;
        aggKeep[blockIdx.x] = agg.keep;
    }
}

// Single CTA: exclusive scan of the tile aggregates, in place.
__global__ void __launch_bounds__(1024) k_rerank_scan(uint32_t *__restrict__ aggHead, uint32_t *__restrict__ aggPst,
                                                      uint32_t *__restrict__ aggKeep, uint32_t tiles)
{
    __shared__ HK part[1024];
    const int tid = threadIdx.x;
    const uint32_t per = (tiles + 1023) / 1024;
    const uint32_t lo = min(tiles, tid * per), hi = min(tiles, lo + per);
    HK s = {0, 0, 0};
    for (uint32_t t = lo; t < hi; t++) s = hk_combine(s, HK{aggHead[t], aggPst[t], aggKeep[t]});
    part[tid] = s;
    __syncthreads();
    if (tid < 32) {
        HK a = {0, 0, 0};
        for (int i = 0; i < 32; i++) a = hk_combine(a, part[tid * 32 + i]);
        HK inc = a;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            HK t;
            t.head = __shfl_up_sync(0xffffffffu, inc.head, o);
            t.pst = __shfl_up_sync(0xffffffffu, inc.pst, o);
            t.keep = __shfl_up_sync(0xffffffffu, inc.keep, o);
            if (tid >= o) inc = hk_combine(t, inc);
        }
        HK run;
        run.head = __shfl_up_sync(0xffffffffu, inc.head, 1);
        run.pst = __shfl_up_sync(0xffffffffu, inc.pst, 1);
        run.keep = __shfl_up_sync(0xffffffffu, inc.keep, 1);
        if (tid == 0) run = HK{0, 0, 0};
        for (int i = 0; i < 32; i++) {
            HK c = part[tid * 32 + i];
            part[tid * 32 + i] = run;
            run = hk_combine(run, c);
        }
    }
    __syncthreads();
    s = part[tid];
    for (uint32_t t = lo; t < hi; t++) {
        HK c = {aggHead[t], aggPst[t], aggKeep[t]};
        aggHead[t] = s.head;
        aggPst[t] = s.pst;
        aggKeep[t] = s.keep;
        s = hk_combine(s, c);
    }
}

template <bool PACKED, bool ROUND0>
__global__ void __launch_bounds__(TT) k_rerank_apply(const uint64_t *__restrict__ keys, const uint32_t *__restrict__ vals,
                                                     ListView lv, int logS, const uint32_t *__restrict__ aggHead,
                                                     const uint32_t *__restrict__ aggPst,
                                                     const uint32_t *__restrict__ aggKeep, uint32_t *__restrict__ rank,
                                                     uint32_t *__restrict__ SA, uint64_t *__restrict__ keysOut,
                                                     uint32_t *__restrict__ valsOut, uint32_t *__restrict__ cntNext)
{
    __shared__ HK scratch[WARPS + 1];
    const TileLoc tl = locate(lv, blockIdx.x);
    const uint32_t segRemain = lv.cnt[tl.b] - tl.lt * TILE;
    const uint32_t i0 = threadIdx.x * ITEMS;
    const uint32_t Smask = (1u << logS) - 1;
    uint64_t e[ITEMS];
    bool head[ITEMS + 1], pst[ITEMS];
    load_flags<PACKED, ROUND0>(keys, tl, segRemain, i0, logS, e, head, pst);
    HK v = {0, 0, 0};
#pragma unroll
    for (int i = 0; i < ITEMS; i++) {
        if (i0 + i < tl.valid) {
            const uint32_t k = tl.first + i0 + i;
            if (head[i]) v.head = k + 1;
            if (pst[i]) v.pst = k + 1;
            if (!(head[i] && head[i + 1])) v.keep++;
        }
    }
    HK agg;
    HK ex = hk_block_excl(v, &agg, scratch);
    ex = hk_combine(HK{aggHead[blockIdx.x], aggPst[blockIdx.x], aggKeep[blockIdx.x]}, ex);
    if (threadIdx.x == 0 && agg.keep) atomicAdd(&cntNext[tl.b], agg.keep);
    uint32_t curHead = ex.head, curPst = ex.pst, outPos = ex.keep;
    const size_t gb = (size_t)tl.b << logS;
#pragma unroll
    for (int i = 0; i < ITEMS; i++) {
        if (i0 + i < tl.valid) {
            const uint32_t k = tl.first + i0 + i;
            if (head[i]) curHead = k + 1;
            if (pst[i]) curPst = k + 1;
            uint32_t p, prim;
            if (PACKED) { p = (uint32_t)e[i] & Smask; prim = ROUND0 ? 0u : (uint32_t)(e[i] >> (2 * logS)); }
            else { p = vals[k]; prim = ROUND0 ? 0u : (uint32_t)(e[i] >> logS); }
            const uint32_t r = prim + (curHead - curPst);  // SA slot of the group head
            rank[gb + p] = (uint32_t)gb | r;
            if (head[i] && head[i + 1]) {
                SA[gb + r] = p;
            } else {
                if (PACKED) {
                    keysOut[outPos] = ((uint64_t)r << (2 * logS)) | p;
                } else {
                    keysOut[outPos] = (uint64_t)r << logS;
                    valsOut[outPos] = p;
                }
                outPos++;
            }
        }
    }
}

}  // namespace

#define CK(call)                                                                          \
    do {                                                                                  \
        cudaError_t e_ = (call);                                                          \
        if (e_ != cudaSuccess) {                                                          \
            fprintf(stderr, "salz_cuda: %s failed at %s:%d: %s\n", #call, __FILE__,       \
                    __LINE__, cudaGetErrorString(e_));                                    \
            return -1;                                                                    \
        }                                                                                 \
    } while (0)

namespace {

struct SortCtx {
    Workspace &ws;
    const BatchDesc &bd;
    cudaStream_t st;
    LaunchCounter &lc;
    bool packed;
    int matchAny;
    int cur;       // keys[cur] holds the list
    int lcur;      // list view index
    uint32_t count, tiles;
};

ListView view(const Workspace &ws, int i)
{
    ListView lv;
    lv.cnt = ws.listCnt[i];
    lv.base = ws.listBase[i];
    lv.tileBase = ws.tileBase[i];
    lv.tileDesc = ws.tileDesc[i];
    return lv;
}

template <bool PACKED>
int radix_sort(SortCtx &s, int passes, int shift0)
{
    Workspace &ws = s.ws;
    const size_t smem = PACKED ? SCATTER_SMEM_PACKED : SCATTER_SMEM_PAIRS;
    const uint32_t grid = std::min<uint32_t>(s.tiles, (uint32_t)(s.ws.numSMs * SCATTER_MIN_CTAS));
    for (int ps = 0; ps < passes; ps++) {
        CK(cudaMemsetAsync(ws.tileState, 0, ((size_t)s.tiles + 1) * 256 * 4, s.st));
        if (s.matchAny)
            k_scatter<PACKED, true><<<grid, TT, smem, s.st>>>(ws.keys[s.cur], ws.vals[s.cur], ws.keys[s.cur ^ 1],
                                                              ws.vals[s.cur ^ 1], view(ws, s.lcur), s.tiles,
                                                              shift0 + 8 * ps, ps, ws.digitBase, ws.tileState);
        else
            k_scatter<PACKED, false><<<grid, TT, smem, s.st>>>(ws.keys[s.cur], ws.vals[s.cur], ws.keys[s.cur ^ 1],
                                                               ws.vals[s.cur ^ 1], view(ws, s.lcur), s.tiles,
                                                               shift0 + 8 * ps, ps, ws.digitBase, ws.tileState);
        s.lc.launches += 1;
        s.lc.sort_launches += 1;
        s.lc.sort_elems += s.count;
        s.lc.sort_bytes += (uint64_t)s.count * (PACKED ? 16 : 24);
        s.cur ^= 1;
    }
    CK(cudaGetLastError());
    return 0;
}

// Re-rank the sorted list, write finished suffixes to SA, compact the rest into
// the other key buffer and build the next round's list map.
template <bool PACKED, bool ROUND0>
int rerank(SortCtx &s)
{
    Workspace &ws = s.ws;
    const int logS = s.bd.logS;
    const int nxt = s.lcur ^ 1;
    const ListView lv = view(ws, s.lcur);
    k_rerank_reduce<PACKED, ROUND0><<<s.tiles, TT, 0, s.st>>>(ws.keys[s.cur], lv, logS, ws.aggHead, ws.aggPst, ws.aggKeep);
    k_rerank_scan<<<1, 1024, 0, s.st>>>(ws.aggHead, ws.aggPst, ws.aggKeep, s.tiles);
    CK(cudaMemsetAsync(ws.listCnt[nxt], 0, (size_t)s.bd.B * 4, s.st));
    k_rerank_apply<PACKED, ROUND0><<<s.tiles, TT, 0, s.st>>>(ws.keys[s.cur], ws.vals[s.cur], lv, logS, ws.aggHead,
                                                             ws.aggPst, ws.aggKeep, ws.rank, ws.SA, ws.keys[s.cur ^ 1],
                                                             ws.vals[s.cur ^ 1], ws.listCnt[nxt]);
    k_tilemap<<<1, 1024, 0, s.st>>>(ws.listCnt[nxt], s.bd.B, ws.listBase[nxt], ws.tileBase[nxt], ws.counters);
    k_tilefill<<<s.bd.B, 128, 0, s.st>>>(ws.listCnt[nxt], ws.listBase[nxt], ws.tileBase[nxt], ws.tileDesc[nxt]);
    s.lc.launches += 5;
    CK(cudaMemcpyAsync(ws.h_counters, ws.counters, 8, cudaMemcpyDeviceToHost, s.st));
    CK(cudaStreamSynchronize(s.st));
    s.count = ws.h_counters[0];
    s.tiles = ws.h_counters[1];
    s.cur ^= 1;  // compacted list was written to the other buffer
    s.lcur = nxt;
    return 0;
}

template <bool PACKED>
int sa_build_t(SortCtx &s, uint64_t &algBytes, int &rounds, float *msSort, cudaEvent_t evA, cudaEvent_t evB)
{
    Workspace &ws = s.ws;
    const BatchDesc &bd = s.bd;
    const int logS = bd.logS;
    static bool attrSet = false;
    if (!attrSet) {
        CK(cudaFuncSetAttribute(k_scatter<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SCATTER_SMEM_PACKED));
        CK(cudaFuncSetAttribute(k_scatter<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SCATTER_SMEM_PACKED));
        CK(cudaFuncSetAttribute(k_scatter<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SCATTER_SMEM_PAIRS));
        CK(cudaFuncSetAttribute(k_scatter<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SCATTER_SMEM_PAIRS));
        attrSet = true;
    }
    const int elemBytes = PACKED ? 8 : 12;
    int nchars = (2 * logS) / 8;
    if (nchars > 8) nchars = 8;
    const int passesD = (2 * logS + 7) / 8;
    const int shift0 = PACKED ? logS : 0;
    const size_t histBytes = (size_t)bd.B * MAXP * 256 * 4;

    // round 0
    k_first_counts<<<(bd.B + 255) / 256, 256, 0, s.st>>>(ws.blkN, bd.B, ws.listCnt[0]);
    k_tilemap<<<1, 1024, 0, s.st>>>(ws.listCnt[0], bd.B, ws.listBase[0], ws.tileBase[0], ws.counters);
    k_tilefill<<<bd.B, 128, 0, s.st>>>(ws.listCnt[0], ws.listBase[0], ws.tileBase[0], ws.tileDesc[0]);
    CK(cudaMemsetAsync(ws.hist, 0, histBytes, s.st));
    k_init<PACKED><<<s.tiles, TT, 0, s.st>>>(ws.T, bd.Tstride, ws.blkN, view(ws, 0), logS, nchars, ws.keys[0], ws.vals[0],
                                             ws.hist);
    k_scan_hist<<<dim3(nchars, bd.B), 256, 0, s.st>>>(ws.hist, ws.digitBase);
    s.lc.launches += 5;
    algBytes += (uint64_t)s.count * (1 + elemBytes);
    if (msSort) CK(cudaEventRecord(evA, s.st));
    if (radix_sort<PACKED>(s, nchars, shift0)) return -1;
    if (msSort) CK(cudaEventRecord(evB, s.st));
    algBytes += (uint64_t)s.count * 2 * elemBytes * nchars;
    uint32_t before = s.count;
    if (rerank<PACKED, true>(s)) return -1;
    if (msSort) { float ms; CK(cudaEventElapsedTime(&ms, evA, evB)); *msSort += ms; }
    algBytes += (uint64_t)before * (2 * elemBytes + 4) + (uint64_t)s.count * elemBytes;

    rounds = 0;
    uint64_t h = (uint64_t)nchars;
    while (s.count > 0) {
        if (h >= bd.S) {
            fprintf(stderr, "salz_cuda: prefix doubling failed to converge (h=%llu)\n", (unsigned long long)h);
            return -1;
        }
        CK(cudaMemsetAsync(ws.hist, 0, histBytes, s.st));
        k_build<PACKED><<<s.tiles, TT, 0, s.st>>>(ws.rank, ws.blkN, view(ws, s.lcur), logS, (uint32_t)h, passesD,
                                                  ws.keys[s.cur], ws.vals[s.cur], ws.hist);
        k_scan_hist<<<dim3(passesD, bd.B), 256, 0, s.st>>>(ws.hist, ws.digitBase);
        s.lc.launches += 2;
        algBytes += (uint64_t)s.count * (2 * elemBytes + 4);
        if (msSort) CK(cudaEventRecord(evA, s.st));
        if (radix_sort<PACKED>(s, passesD, shift0)) return -1;
        if (msSort) CK(cudaEventRecord(evB, s.st));
        algBytes += (uint64_t)s.count * 2 * elemBytes * passesD;
        before = s.count;
        if (rerank<PACKED, false>(s)) return -1;
        if (msSort) { float ms; CK(cudaEventElapsedTime(&ms, evA, evB)); *msSort += ms; }
        algBytes += (uint64_t)before * (2 * elemBytes + 4) + (uint64_t)s.count * elemBytes;
        h *= 2;
        rounds++;
    }
    CK(cudaGetLastError());
    return 0;
}

}  // namespace

int sa_tile_elems() { return TILE; }

int sa_build(Workspace &ws, const BatchDesc &bd, int sortMode, cudaStream_t st, LaunchCounter &lc,
             uint64_t &algBytes, int &rounds, float *msSort)
{
    if (bd.M == 0) return 0;
    cudaEvent_t evA = nullptr, evB = nullptr;
    if (msSort) {
        CK(cudaEventCreate(&evA));
        CK(cudaEventCreate(&evB));
        *msSort = 0.f;
    }
    uint32_t tiles = 0;
    for (int b = 0; b < bd.B; b++) tiles += (ws.h_blkN[b] - 8 + TILE - 1) / TILE;
    SortCtx s = {ws, bd, st, lc, 3 * bd.logS <= 64, sortMode == 0, 0, 0, bd.M, tiles};
    int rc = s.packed ? sa_build_t<true>(s, algBytes, rounds, msSort, evA, evB)
                      : sa_build_t<false>(s, algBytes, rounds, msSort, evA, evB);
    if (msSort) {
        cudaEventDestroy(evA);
        cudaEventDestroy(evB);
    }
    return rc;
}
