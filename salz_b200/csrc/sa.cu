// sa.cu - batched suffix-array construction by prefix doubling.
//
// Replaces the single libsais() call of the reference (ref lib/salz.c:463-469,
// 3rdparty/libsais/src/libsais.c:6597-6611).  The suffix array is unique, so
// any exact suffix sorter yields the same bytes downstream.
//
// All blocks of a batch are sorted together.  The sort list holds the suffixes
// whose final position is not yet known ("unresolved"); it is always ordered by
// the group they belong to, and group ids are padded global SA slots
// (b*S + index of the group head), so one global stable LSD radix sort of the
// packed key (group id, rank of the suffix h positions further) re-sorts every
// group of every block at once.  Suffixes whose group became a singleton are
// written to SA and dropped from the list (the list shrinks every round).
//
// Radix pass = one kernel: per-warp shared-memory digit counters filled with
// __match_any_sync ranking, decoupled look-back over per-tile digit counts for
// the global offsets, keys/values staged through shared memory in digit order
// and written back as coalesced runs.
#include "salz_dev.cuh"

namespace {

constexpr int SORT_THREADS = 512;
constexpr int SORT_ITEMS = 8;
constexpr int SORT_TILE = SORT_THREADS * SORT_ITEMS;
constexpr int SORT_WARPS = SORT_THREADS / 32;
constexpr size_t SORT_SMEM = (size_t)SORT_TILE * 12 + (size_t)SORT_WARPS * 256 * 4 + 2 * 256 * 4 + 64;

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

// Exclusive sum scan over the values held by threads 0..255 (others pass 0).
// Must be called by every thread of the CTA.  scratch: >= 9 uint32 in smem.
__device__ __forceinline__ uint32_t excl_scan_first256(uint32_t v, uint32_t *scratch, uint32_t *total)
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
        scratch[8] = s;
    }
    __syncthreads();
    uint32_t r = (warp < 8) ? (scratch[warp] + inc - v) : 0;
    if (total) *total = scratch[8];
    __syncthreads();
    return r;
}

// Adds the digits of `key` for passes [0,passes) to shared histograms, one
// shared atomic per distinct digit per warp.
__device__ __forceinline__ void hist_add(uint32_t *sh, uint64_t key, int passes, bool valid)
{
    const int lane = threadIdx.x & 31;
    unsigned act = __ballot_sync(0xffffffffu, valid);
    if (!valid) return;
    for (int ps = 0; ps < passes; ps++) {
        uint32_t d = (uint32_t)(key >> (8 * ps)) & 255u;
        unsigned m = __match_any_sync(act, d);
        if (lane == __ffs(m) - 1) atomicAdd(&sh[ps * 256 + d], (uint32_t)__popc(m));
    }
}

__device__ __forceinline__ int find_block(const uint32_t *elemStart, int B, uint32_t e)
{
    int lo = 0, hi = B;  // largest b with elemStart[b] <= e
    while (hi - lo > 1) {
        int mid = (lo + hi) >> 1;
        if (elemStart[mid] <= e) lo = mid; else hi = mid;
    }
    return lo;
}

// Round 0: key = (block id, first h0 bytes big-endian, zero padded past n_b).
// Elements of a block are emitted in DESCENDING text position so that the
// stable sort leaves zero-padded ties (only the last < h0 suffixes of a block
// can be involved) in their correct order: shorter suffix first.
__global__ void __launch_bounds__(256) k_init_keys(const uint8_t *__restrict__ T, size_t Tstride,
                                                   const uint32_t *__restrict__ blkN,
                                                   const uint32_t *__restrict__ elemStart, int B, int logS,
                                                   int h0, int passes, uint32_t M,
                                                   uint64_t *__restrict__ keys, uint32_t *__restrict__ vals,
                                                   uint32_t *__restrict__ slots, uint32_t *__restrict__ hist)
{
    __shared__ uint32_t sh[8 * 256];
    for (int i = threadIdx.x; i < passes * 256; i += blockDim.x) sh[i] = 0;
    __syncthreads();
    uint32_t e = blockIdx.x * blockDim.x + threadIdx.x;
    bool valid = e < M;
    uint64_t key = 0;
    if (valid) {
        int b = find_block(elemStart, B, e);
        uint32_t r = e - elemStart[b];
        uint32_t n = blkN[b] - 8;
        uint32_t p = n - 1 - r;
        uint64_t le = load_u64_unaligned(T + (size_t)b * Tstride + p);
        uint64_t be = __byte_perm((uint32_t)(le >> 32), 0, 0x0123) |
                      ((uint64_t)__byte_perm((uint32_t)le, 0, 0x0123) << 32);
        uint32_t have = n - p;  // bytes of the suffix
        if (have < 8) be &= ~0ull << (8 * (8 - have));
        uint64_t chars = (h0 == 8) ? be : (be >> (8 * (8 - h0)));
        key = chars | ((h0 == 8) ? 0ull : ((uint64_t)b << (8 * h0)));
        keys[e] = key;
        vals[e] = ((uint32_t)b << logS) | p;
        slots[e] = ((uint32_t)b << logS) | r;
    }
    hist_add(sh, key, passes, valid);
    __syncthreads();
    for (int i = threadIdx.x; i < passes * 256; i += blockDim.x)
        if (sh[i]) atomicAdd(&hist[i], sh[i]);
}

// Doubling round: key = (group id << logS) | (1 + local rank of suffix p+h, 0 past the end).
__global__ void __launch_bounds__(256) k_build_keys(const uint32_t *__restrict__ vals,
                                                    const uint32_t *__restrict__ prim,
                                                    const uint32_t *__restrict__ rank,
                                                    const uint32_t *__restrict__ blkN, int logS, uint32_t h,
                                                    int passes, uint32_t count, uint64_t *__restrict__ keys,
                                                    uint32_t *__restrict__ hist)
{
    __shared__ uint32_t sh[8 * 256];
    for (int i = threadIdx.x; i < passes * 256; i += blockDim.x) sh[i] = 0;
    __syncthreads();
    uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    bool valid = k < count;
    uint64_t key = 0;
    if (valid) {
        uint32_t g = vals[k];
        uint32_t Smask = (1u << logS) - 1;
        uint32_t b = g >> logS, p = g & Smask;
        uint32_t n = blkN[b] - 8;
        uint32_t sec = 0;
        if (p + h < n) sec = (rank[g + h] & Smask) + 1;
        key = ((uint64_t)prim[k] << logS) | sec;
        keys[k] = key;
    }
    hist_add(sh, key, passes, valid);
    __syncthreads();
    for (int i = threadIdx.x; i < passes * 256; i += blockDim.x)
        if (sh[i]) atomicAdd(&hist[i], sh[i]);
}

// digitBase[pass][d] = exclusive scan of hist[pass][*]
__global__ void __launch_bounds__(256) k_scan_hist(const uint32_t *__restrict__ hist, uint32_t *__restrict__ digitBase)
{
    __shared__ uint32_t scratch[16];
    uint32_t v = hist[blockIdx.x * 256 + threadIdx.x];
    uint32_t ex = excl_scan_first256(v, scratch, nullptr);
    digitBase[blockIdx.x * 256 + threadIdx.x] = ex;
}

// ---- fallback (mode 0): per-tile histograms + serial scan -------------------
__global__ void __launch_bounds__(256) k_tile_hist(const uint64_t *__restrict__ keys, uint32_t count, int shift,
                                                   uint32_t *__restrict__ tileHist)
{
    __shared__ uint32_t sh[256];
    sh[threadIdx.x] = 0;
    __syncthreads();
    uint32_t base = blockIdx.x * SORT_TILE;
    for (int i = threadIdx.x; i < SORT_TILE; i += 256) {
        uint32_t k = base + i;
        if (k < count) atomicAdd(&sh[(uint32_t)(keys[k] >> shift) & 255u], 1u);
    }
    __syncthreads();
    tileHist[(size_t)blockIdx.x * 256 + threadIdx.x] = sh[threadIdx.x];
}

// tileHist[t][d] -> exclusive prefix over tiles; digitBase[d] = exclusive scan of totals
__global__ void __launch_bounds__(256) k_scan_tile_hist(uint32_t *__restrict__ tileHist, uint32_t tiles,
                                                        uint32_t *__restrict__ digitBase)
{
    __shared__ uint32_t scratch[16];
    uint32_t sum = 0;
    for (uint32_t t = 0; t < tiles; t++) {
        uint32_t c = tileHist[(size_t)t * 256 + threadIdx.x];
        tileHist[(size_t)t * 256 + threadIdx.x] = sum;
        sum += c;
    }
    uint32_t ex = excl_scan_first256(sum, scratch, nullptr);
    digitBase[threadIdx.x] = ex;
}

// ---- the radix scatter pass --------------------------------------------------
// tileOffsets == nullptr : decoupled look-back over tileState (tile ids handed
//                          out by an atomic ticket so predecessors are running)
// tileOffsets != nullptr : offsets precomputed by k_scan_tile_hist
__global__ void __launch_bounds__(SORT_THREADS, 2)
k_radix_scatter(const uint64_t *__restrict__ keysIn, const uint32_t *__restrict__ valsIn,
                uint64_t *__restrict__ keysOut, uint32_t *__restrict__ valsOut, uint32_t count, int shift,
                const uint32_t *__restrict__ digitBase, uint32_t *tileState, uint32_t *tileCounter,
                const uint32_t *__restrict__ tileOffsets)
{
    extern __shared__ __align__(16) uint8_t smem[];
    uint64_t *skeys = (uint64_t *)smem;
    uint32_t *svals = (uint32_t *)(skeys + SORT_TILE);
    uint32_t *wcnt = svals + SORT_TILE;           // [SORT_WARPS][256]
    uint32_t *dstart = wcnt + SORT_WARPS * 256;   // [256]
    uint32_t *gbase = dstart + 256;               // [256]
    uint32_t *scratch = gbase + 256;              // [16]
    __shared__ uint32_t s_tile;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    uint32_t tile;
    if (tileOffsets == nullptr) {
        if (tid == 0) s_tile = atomicAdd(tileCounter, 1u);
        __syncthreads();
        tile = s_tile;
    } else {
        tile = blockIdx.x;
    }
    const uint32_t base = tile * SORT_TILE;
    const uint32_t valid = min((uint32_t)SORT_TILE, count - base);

    for (int i = tid; i < SORT_WARPS * 256; i += SORT_THREADS) wcnt[i] = 0;

    uint64_t key[SORT_ITEMS];
    uint32_t val[SORT_ITEMS];
#pragma unroll
    for (int i = 0; i < SORT_ITEMS; i++) {
        uint32_t idx = warp * (32 * SORT_ITEMS) + i * 32 + lane;
        if (idx < valid) {
            key[i] = keysIn[base + idx];
            val[i] = valsIn[base + idx];
        } else {
            key[i] = ~0ull;
            val[i] = 0;
        }
    }
    __syncthreads();

    uint32_t rnk[SORT_ITEMS];
    uint32_t *mycnt = wcnt + warp * 256;
#pragma unroll
    for (int i = 0; i < SORT_ITEMS; i++) {
        uint32_t d = (uint32_t)(key[i] >> shift) & 255u;
        unsigned m = __match_any_sync(0xffffffffu, d);
        int leader = __ffs(m) - 1;
        uint32_t pre = 0;
        if (lane == leader) {
            pre = mycnt[d];
            mycnt[d] = pre + __popc(m);
        }
        pre = __shfl_sync(0xffffffffu, pre, leader);
        rnk[i] = pre + __popc(m & ((1u << lane) - 1u));
        __syncwarp();
    }
    __syncthreads();

    uint32_t total = 0;
    if (tid < 256) {
        uint32_t sum = 0;
#pragma unroll
        for (int w = 0; w < SORT_WARPS; w++) {
            uint32_t c = wcnt[w * 256 + tid];
            wcnt[w * 256 + tid] = sum;
            sum += c;
        }
        total = sum;
    }
    uint32_t ex = excl_scan_first256(total, scratch, nullptr);
    if (tid < 256) {
        dstart[tid] = ex;
        uint32_t cnt = total - ((tid == 255) ? (SORT_TILE - valid) : 0u);
        uint32_t prev = 0;
        if (tileOffsets != nullptr) {
            prev = tileOffsets[(size_t)tile * 256 + tid];
        } else if (tile == 0) {
            st_relaxed(&tileState[tid], LB_INCL | cnt);
        } else {
            st_relaxed(&tileState[(size_t)tile * 256 + tid], LB_AGG | cnt);
            for (uint32_t t = tile - 1;; --t) {
                uint32_t s;
                do { s = ld_relaxed(&tileState[(size_t)t * 256 + tid]); } while ((s >> 30) == 0);
                prev += s & LB_MASK;
                if ((s >> 30) == 2) break;
            }
            st_relaxed(&tileState[(size_t)tile * 256 + tid], LB_INCL | (prev + cnt));
        }
        gbase[tid] = digitBase[tid] + prev - ex;
    }
    __syncthreads();

#pragma unroll
    for (int i = 0; i < SORT_ITEMS; i++) {
        uint32_t d = (uint32_t)(key[i] >> shift) & 255u;
        uint32_t pos = dstart[d] + mycnt[d] + rnk[i];
        skeys[pos] = key[i];
        svals[pos] = val[i];
    }
    __syncthreads();

    for (uint32_t j = tid; j < valid; j += SORT_THREADS) {
        uint64_t k = skeys[j];
        uint32_t d = (uint32_t)(k >> shift) & 255u;
        uint32_t dst = gbase[d] + j;
        keysOut[dst] = k;
        valsOut[dst] = svals[j];
    }
}

}  // namespace

// ---- re-rank + compaction ----------------------------------------------------
namespace {

constexpr int RR_THREADS = 256;
constexpr int RR_ITEMS = 8;
constexpr int RR_TILE = RR_THREADS * RR_ITEMS;

// A sorted list element k starts a new group ("head") when its key differs
// from its predecessor's.  From round 1 on a zero secondary key (suffix ends
// within h characters) is always a head: two different suffixes of the same
// group cannot both end inside the compared prefix.
__device__ __forceinline__ bool is_head(uint64_t prev, uint64_t cur, uint32_t k, uint64_t secMask)
{
    return k == 0 || prev != cur || (secMask != 0 && (cur & secMask) == 0);
}

struct HeadKeep { uint32_t head; uint32_t keep; };  // head: slot+1 of last head (0 none)

__device__ __forceinline__ HeadKeep hk_combine(HeadKeep a, HeadKeep b)
{
    HeadKeep r;
    r.head = b.head ? b.head : a.head;
    r.keep = a.keep + b.keep;
    return r;
}

// Block-wide exclusive scan of HeadKeep over RR_THREADS threads; returns the
// exclusive prefix for this thread and the block aggregate in *agg.
__device__ __forceinline__ HeadKeep hk_block_excl(HeadKeep v, HeadKeep *agg, HeadKeep *scratch /*[8+1]*/)
{
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    HeadKeep inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        HeadKeep t;
        t.head = __shfl_up_sync(0xffffffffu, inc.head, o);
        t.keep = __shfl_up_sync(0xffffffffu, inc.keep, o);
        if (lane >= o) inc = hk_combine(t, inc);
    }
    if (lane == 31) scratch[warp] = inc;
    __syncthreads();
    if (tid == 0) {
        HeadKeep s; s.head = 0; s.keep = 0;
        for (int w = 0; w < RR_THREADS / 32; w++) {
            HeadKeep c = scratch[w];
            scratch[w] = s;
            s = hk_combine(s, c);
        }
        scratch[RR_THREADS / 32] = s;
    }
    __syncthreads();
    HeadKeep ex;
    ex.head = __shfl_up_sync(0xffffffffu, inc.head, 1);
    ex.keep = __shfl_up_sync(0xffffffffu, inc.keep, 1);
    if (lane == 0) { ex.head = 0; ex.keep = 0; }
    ex = hk_combine(scratch[warp], ex);
    *agg = scratch[RR_THREADS / 32];
    __syncthreads();
    return ex;
}

// Loads keys[k0-1 .. k0+RR_ITEMS] and returns per-item flags.
__device__ __forceinline__ void load_flags(const uint64_t *__restrict__ keys, uint32_t count, uint32_t k0,
                                           uint64_t secMask, bool head[RR_ITEMS + 1])
{
    uint64_t prev = (k0 > 0 && k0 - 1 < count) ? keys[k0 - 1] : 0;
#pragma unroll
    for (int i = 0; i <= RR_ITEMS; i++) {
        uint32_t k = k0 + i;
        if (k < count) {
            uint64_t cur = keys[k];
            head[i] = is_head(prev, cur, k, secMask);
            prev = cur;
        } else {
            head[i] = true;  // end of list closes the last group
        }
    }
}

__global__ void __launch_bounds__(RR_THREADS) k_rerank_reduce(const uint64_t *__restrict__ keys,
                                                              const uint32_t *__restrict__ slots, uint32_t count,
                                                              uint64_t secMask, uint32_t *__restrict__ aggKeep,
                                                              uint32_t *__restrict__ aggHead)
{
    __shared__ HeadKeep scratch[RR_THREADS / 32 + 1];
    uint32_t k0 = blockIdx.x * RR_TILE + threadIdx.x * RR_ITEMS;
    bool head[RR_ITEMS + 1];
    load_flags(keys, count, k0, secMask, head);
    HeadKeep v; v.head = 0; v.keep = 0;
#pragma unroll
    for (int i = 0; i < RR_ITEMS; i++) {
        uint32_t k = k0 + i;
        if (k < count) {
            if (head[i]) v.head = slots[k] + 1;
            if (!(head[i] && head[i + 1])) v.keep++;
        }
    }
    HeadKeep agg;
    hk_block_excl(v, &agg, scratch);
    if (threadIdx.x == 0) {
        aggKeep[blockIdx.x] = agg.keep;
        aggHead[blockIdx.x] = agg.head;
    }
}

// Single CTA: exclusive scan of the tile aggregates, in place.
__global__ void __launch_bounds__(1024) k_rerank_scan(uint32_t *__restrict__ aggKeep, uint32_t *__restrict__ aggHead,
                                                      uint32_t tiles, uint32_t *__restrict__ counters)
{
    __shared__ HeadKeep part[1024];
    const int tid = threadIdx.x;
    uint32_t per = (tiles + 1023) / 1024;
    uint32_t lo = tid * per, hi = min(tiles, lo + per);
    HeadKeep s; s.head = 0; s.keep = 0;
    for (uint32_t t = lo; t < hi; t++) {
        HeadKeep c; c.head = aggHead[t]; c.keep = aggKeep[t];
        s = hk_combine(s, c);
    }
    part[tid] = s;
    __syncthreads();
    if (tid == 0) {
        HeadKeep run; run.head = 0; run.keep = 0;
        for (int i = 0; i < 1024; i++) {
            HeadKeep c = part[i];
            part[i] = run;
            run = hk_combine(run, c);
        }
        counters[0] = run.keep;
    }
    __syncthreads();
    s = part[tid];
    for (uint32_t t = lo; t < hi; t++) {
        HeadKeep c; c.head = aggHead[t]; c.keep = aggKeep[t];
        aggHead[t] = s.head;
        aggKeep[t] = s.keep;
        s = hk_combine(s, c);
    }
}

__global__ void __launch_bounds__(RR_THREADS) k_rerank_apply(
    const uint64_t *__restrict__ keys, const uint32_t *__restrict__ vals, const uint32_t *__restrict__ slots,
    uint32_t count, uint64_t secMask, uint32_t Smask, const uint32_t *__restrict__ aggKeep,
    const uint32_t *__restrict__ aggHead, uint32_t *__restrict__ rank, uint32_t *__restrict__ SA,
    uint32_t *__restrict__ valsOut, uint32_t *__restrict__ slotsOut, uint32_t *__restrict__ primOut)
{
    __shared__ HeadKeep scratch[RR_THREADS / 32 + 1];
    uint32_t k0 = blockIdx.x * RR_TILE + threadIdx.x * RR_ITEMS;
    bool head[RR_ITEMS + 1];
    load_flags(keys, count, k0, secMask, head);
    uint32_t sl[RR_ITEMS], vv[RR_ITEMS];
    HeadKeep v; v.head = 0; v.keep = 0;
#pragma unroll
    for (int i = 0; i < RR_ITEMS; i++) {
        uint32_t k = k0 + i;
        sl[i] = 0; vv[i] = 0;
        if (k < count) {
            sl[i] = slots[k];
            vv[i] = vals[k];
            if (head[i]) v.head = sl[i] + 1;
            if (!(head[i] && head[i + 1])) v.keep++;
        }
    }
    HeadKeep agg;
    HeadKeep ex = hk_block_excl(v, &agg, scratch);
    HeadKeep carry; carry.head = aggHead[blockIdx.x]; carry.keep = aggKeep[blockIdx.x];
    ex = hk_combine(carry, ex);
    uint32_t curHead = ex.head;  // slot+1 of the head governing the next element
    uint32_t outPos = ex.keep;
#pragma unroll
    for (int i = 0; i < RR_ITEMS; i++) {
        uint32_t k = k0 + i;
        if (k < count) {
            if (head[i]) curHead = sl[i] + 1;
            uint32_t r = curHead - 1;
            rank[vv[i]] = r;
            if (head[i] && head[i + 1]) {
                SA[sl[i]] = vv[i] & Smask;
            } else {
                valsOut[outPos] = vv[i];
                slotsOut[outPos] = sl[i];
                primOut[outPos] = r;
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

static int radix_sort(Workspace &ws, uint32_t count, int passes, int &cur, int mode, cudaStream_t st,
                      LaunchCounter &lc, cudaEvent_t evA, cudaEvent_t evB, float *msSort)
{
    static bool attrSet = false;
    if (!attrSet) {
        CK(cudaFuncSetAttribute(k_radix_scatter, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SORT_SMEM));
        attrSet = true;
    }
    uint32_t tiles = (count + SORT_TILE - 1) / SORT_TILE;
    if (msSort) CK(cudaEventRecord(evA, st));
    for (int ps = 0; ps < passes; ps++) {
        int shift = 8 * ps;
        if (mode == 1) {
            CK(cudaMemsetAsync(ws.tileState, 0, (size_t)tiles * 256 * 4, st));
            CK(cudaMemsetAsync(ws.tileCounter, 0, 4, st));
            k_radix_scatter<<<tiles, SORT_THREADS, SORT_SMEM, st>>>(
                ws.keys[cur], ws.vals[cur], ws.keys[cur ^ 1], ws.vals[cur ^ 1], count, shift,
                ws.digitBase + ps * 256, ws.tileState, ws.tileCounter, nullptr);
            lc.launches += 1;
        } else {
            k_tile_hist<<<tiles, 256, 0, st>>>(ws.keys[cur], count, shift, ws.tileHist);
            k_scan_tile_hist<<<1, 256, 0, st>>>(ws.tileHist, tiles, ws.digitBase + ps * 256);
            k_radix_scatter<<<tiles, SORT_THREADS, SORT_SMEM, st>>>(
                ws.keys[cur], ws.vals[cur], ws.keys[cur ^ 1], ws.vals[cur ^ 1], count, shift,
                ws.digitBase + ps * 256, nullptr, nullptr, ws.tileHist);
            lc.launches += 3;
        }
        lc.sort_launches += 1;
        lc.sort_elems += count;
        cur ^= 1;
    }
    if (msSort) {
        CK(cudaEventRecord(evB, st));
    }
    CK(cudaGetLastError());
    return 0;
}

static int rerank(Workspace &ws, uint32_t count, int cur, int slotCur, uint64_t secMask, uint32_t Smask,
                  cudaStream_t st, LaunchCounter &lc, uint32_t *kept)
{
    uint32_t tiles = (count + RR_TILE - 1) / RR_TILE;
    k_rerank_reduce<<<tiles, RR_THREADS, 0, st>>>(ws.keys[cur], ws.slot[slotCur], count, secMask, ws.aggKeep,
                                                  ws.aggHead);
    k_rerank_scan<<<1, 1024, 0, st>>>(ws.aggKeep, ws.aggHead, tiles, ws.counters);
    // compacted values go to the other vals buffer (the sort's scratch side)
    k_rerank_apply<<<tiles, RR_THREADS, 0, st>>>(ws.keys[cur], ws.vals[cur], ws.slot[slotCur], count, secMask,
                                                 Smask, ws.aggKeep, ws.aggHead, ws.rank, ws.SA, ws.vals[cur ^ 1],
                                                 ws.slot[slotCur ^ 1], ws.prim);
    lc.launches += 3;
    CK(cudaMemcpyAsync(ws.h_counters, ws.counters, 4, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    *kept = ws.h_counters[0];
    return 0;
}

static int ceil_log2_u64(uint64_t v)
{
    int l = 0;
    while ((1ull << l) < v) l++;
    return l;
}

int sa_build(Workspace &ws, const BatchDesc &bd, int sortMode, cudaStream_t st, LaunchCounter &lc,
             uint64_t &algBytes, int &rounds, float *msSort)
{
    const uint32_t M = bd.M;
    if (M == 0) return 0;
    cudaEvent_t evA = nullptr, evB = nullptr;
    if (msSort) {
        CK(cudaEventCreate(&evA));
        CK(cudaEventCreate(&evB));
        *msSort = 0.f;
    }
    int bbits = ceil_log2_u64((uint64_t)bd.B);
    int h0 = (64 - bbits) / 8;
    if (h0 > 8) h0 = 8;
    int passes0 = (h0 * 8 + bbits + 7) / 8;
    int passesD = (2 * bd.logS + bbits + 7) / 8;
    const uint32_t Smask = bd.S - 1;

    CK(cudaMemsetAsync(ws.hist, 0, 8 * 256 * 4, st));
    k_init_keys<<<(M + 255) / 256, 256, 0, st>>>(ws.T, bd.Tstride, ws.blkN, ws.elemStart, bd.B, bd.logS, h0,
                                                  passes0, M, ws.keys[0], ws.vals[0], ws.slot[0], ws.hist);
    k_scan_hist<<<passes0, 256, 0, st>>>(ws.hist, ws.digitBase);
    lc.launches += 2;
    algBytes += (uint64_t)M * (1 + 8 + 4 + 4);
    int cur = 0, slotCur = 0;
    if (radix_sort(ws, M, passes0, cur, sortMode, st, lc, evA, evB, msSort)) return -1;
    algBytes += (uint64_t)M * 24 * passes0;
    uint32_t kept = 0;
    if (rerank(ws, M, cur, slotCur, 0, Smask, st, lc, &kept)) return -1;
    if (msSort) { float ms; CK(cudaEventElapsedTime(&ms, evA, evB)); *msSort += ms; }
    algBytes += (uint64_t)M * (8 + 16 + 8) + (uint64_t)kept * 12;
    cur ^= 1;      // compacted vals were written to the other buffer
    slotCur ^= 1;
    rounds = 0;
    uint64_t h = (uint64_t)h0;
    const uint64_t secMask = ((uint64_t)1 << bd.logS) - 1;
    while (kept > 0) {
        uint32_t count = kept;
        if (h >= bd.S) {
            fprintf(stderr, "salz_cuda: prefix doubling failed to converge (h=%llu)\n", (unsigned long long)h);
            return -1;
        }
        CK(cudaMemsetAsync(ws.hist, 0, 8 * 256 * 4, st));
        k_build_keys<<<(count + 255) / 256, 256, 0, st>>>(ws.vals[cur], ws.prim, ws.rank, ws.blkN, bd.logS,
                                                          (uint32_t)h, passesD, count, ws.keys[cur], ws.hist);
        k_scan_hist<<<passesD, 256, 0, st>>>(ws.hist, ws.digitBase);
        lc.launches += 2;
        algBytes += (uint64_t)count * 20;
        if (radix_sort(ws, count, passesD, cur, sortMode, st, lc, evA, evB, msSort)) return -1;
        algBytes += (uint64_t)count * 24 * passesD;
        if (rerank(ws, count, cur, slotCur, secMask, Smask, st, lc, &kept)) return -1;
        if (msSort) { float ms; CK(cudaEventElapsedTime(&ms, evA, evB)); *msSort += ms; }
        algBytes += (uint64_t)count * (8 + 16 + 8) + (uint64_t)kept * 12;
        cur ^= 1;
        slotCur ^= 1;
        h *= 2;
        rounds++;
    }
    if (msSort) {
        cudaEventDestroy(evA);
        cudaEventDestroy(evB);
    }
    CK(cudaGetLastError());
    return 0;
}
