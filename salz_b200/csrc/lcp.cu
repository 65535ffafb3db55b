// lcp.cu - PLCP by the phi method restricted to irreducible positions, LCP in
// suffix-array order, and the min-pyramids the ANSV/range-min stage walks.
//
// salz itself never materialises an LCP array (it compares text directly,
// ref lib/salz.c:492-538); this stage exists because the B200 design obtains
// the match lengths as range minima of LCP.  Oracle: libsais_plcp/libsais_lcp
// (ref 3rdparty/libsais/src/libsais.h:319,329).
//
//   Phi[i]  = SA[ISA[i]-1]           (none for the smallest suffix)
//   i is irreducible  <=>  i == 0 or Phi[i] == 0 or T[i-1] != T[Phi[i]-1]
//   irreducible: PLCP[i] by direct comparison (sum over a block is O(n log n))
//   reducible:   PLCP[i] = PLCP[i-1] - 1
// PLCP[i] + i is non-decreasing in i, so the fill is an inclusive max-scan of
// V[i] = (irreducible ? PLCP[i] + i : 0) followed by "- i".
#include "salz_dev.cuh"

namespace {

constexpr int LT = 1024;  // threads = positions per tile

// Length of the common prefix of T[a..) and T[b..) limited to `cap` bytes,
// thread-private part: at most `budget` bytes are examined; returns the
// number of matching bytes found so far (== budget means "still equal").
__device__ __forceinline__ uint32_t lcp_private(const uint8_t *Ta, const uint8_t *Tb, uint32_t cap, uint32_t budget)
{
    uint32_t l = 0;
    uint32_t lim = min(cap, budget);
    while (l < lim) {
        uint64_t x = load_u64_unaligned(Ta + l) ^ load_u64_unaligned(Tb + l);
        if (x) {
            l += (uint32_t)(__ffsll((long long)x) - 1) >> 3;
            return min(l, lim);
        }
        l += 8;
    }
    return lim;
}

// One thread per position, CTAs of IT threads (a tile of LT positions is IT/LT CTAs: a
// CTA is gone as soon as its own warps are, and the tile maximum is an atomicMax per warp;
// tileMax is zeroed before the launch).
constexpr int IT = 256;
#ifndef SALZ_PLCP_CTAS
#define SALZ_PLCP_CTAS 6
#endif
// PI positions per thread (IT apart), their three dependent gathers (rank -> SA -> text)
// issued stage by stage so that PI chains are in flight per thread: the kernel is bound by
// the latency of those gathers, not by any throughput.
#ifndef SALZ_PLCP_ITEMS
#define SALZ_PLCP_ITEMS 2
#endif
constexpr int PI = SALZ_PLCP_ITEMS;
static_assert(LT % (IT * PI) == 0, "a CTA's positions must lie in one tile");
__global__ void __launch_bounds__(IT, SALZ_PLCP_CTAS) k_plcp_irreducible(const uint8_t *__restrict__ T, size_t Tstride,
                                                         const uint32_t *__restrict__ blkN,
                                                         const uint64_t *__restrict__ rank,
                                                         const uint32_t *__restrict__ SA, int logS,
                                                         int32_t *__restrict__ V, int32_t *__restrict__ tileMax)
{
    const uint32_t Smask = (1u << logS) - 1;
    const uint32_t g0 = blockIdx.x * (IT * PI) + threadIdx.x;
    const uint32_t b = g0 >> logS;   // S is a multiple of IT * PI: one block per CTA
    const uint32_t n = blkN[b] - 8;  // blkN >= 9 for every block of a batch
    const uint8_t *Tb = T + (size_t)b * Tstride;
    const int lane = threadIdx.x & 31;

    uint32_t p[PI], q[PI], cap[PI], l[PI], r[PI];
    int32_t v[PI];
    bool pending[PI], live[PI], cmp[PI];
    uint64_t rk[PI], xa[PI], xb[PI];
#pragma unroll
    for (int k = 0; k < PI; k++) {
        p[k] = (g0 + k * IT) & Smask;
        live[k] = p[k] < n;
        rk[k] = live[k] ? rank[g0 + k * IT] : 0;
    }
#pragma unroll
    for (int k = 0; k < PI; k++) {
        r[k] = (uint32_t)rk[k] & 0xffffffu;   // block-local slot
        live[k] = live[k] && r[k] != 0;       // slot 0 = smallest suffix of the block: PLCP = 0
        q[k] = live[k] ? SA[(g0 & ~Smask) + r[k] - 1] : 0;
    }
#pragma unroll
    for (int k = 0; k < PI; k++) {
        // one load per side: the preceding byte (irreducible if it differs) and the first
        // 7 bytes of the comparison
        cmp[k] = live[k] && p[k] != 0 && q[k] != 0;
        xa[k] = cmp[k] ? load_u64_unaligned(Tb + p[k] - 1) : 0;
        xb[k] = cmp[k] ? load_u64_unaligned(Tb + q[k] - 1) : 0;
    }
#pragma unroll
    for (int k = 0; k < PI; k++) {
        v[k] = 0;
        l[k] = 0;
        cap[k] = 0;
        pending[k] = false;
        if (p[k] < n && r[k] == 0) v[k] = (int32_t)p[k];
        if (live[k]) {
            cap[k] = n - max(p[k], q[k]);
            if (!cmp[k]) {
                l[k] = lcp_private(Tb + p[k], Tb + q[k], cap[k], 32);
                pending[k] = (l[k] == 32 && cap[k] > 32);
                v[k] = (int32_t)(l[k] + p[k]);
            } else {
                const uint64_t x = xa[k] ^ xb[k];
                if (x & 0xffu) {
                    const uint64_t y = x >> 8;
                    if (y) {
                        l[k] = min((uint32_t)(__ffsll((long long)y) - 1) >> 3, cap[k]);
                    } else {
                        l[k] = 7;
                        if (cap[k] > 7) l[k] += lcp_private(Tb + p[k] + 7, Tb + q[k] + 7, cap[k] - 7, 32);
                        else l[k] = cap[k];
                        pending[k] = (l[k] == 39 && cap[k] > 39);
                    }
                    v[k] = (int32_t)(l[k] + p[k]);
                }
            }
        }
    }
    // Long matches: the whole warp compares 32 x 16 bytes per step.
#pragma unroll
    for (int k = 0; k < PI; k++) {
        unsigned todo = __ballot_sync(0xffffffffu, pending[k]);
        while (todo) {
            int src = __ffs(todo) - 1;
            todo &= todo - 1;
            uint32_t p0 = __shfl_sync(0xffffffffu, p[k], src);
            uint32_t q0 = __shfl_sync(0xffffffffu, q[k], src);
            uint32_t cap0 = __shfl_sync(0xffffffffu, cap[k], src);
            uint32_t l0 = __shfl_sync(0xffffffffu, l[k], src);
            for (;;) {
                uint32_t off = l0 + lane * 16;
                uint32_t mine = 16;  // matching bytes in my 16-byte window
                if (off < cap0) {
                    uint64_t x = load_u64_unaligned(Tb + p0 + off) ^ load_u64_unaligned(Tb + q0 + off);
                    if (x) {
                        mine = (uint32_t)(__ffsll((long long)x) - 1) >> 3;
                    } else {
                        uint64_t y = load_u64_unaligned(Tb + p0 + off + 8) ^ load_u64_unaligned(Tb + q0 + off + 8);
                        if (y) mine = 8 + ((uint32_t)(__ffsll((long long)y) - 1) >> 3);
                    }
                    mine = min(mine, cap0 - off);
                } else {
                    mine = 0;
                }
                unsigned stop = __ballot_sync(0xffffffffu, mine < 16);
                if (stop) {
                    int first = __ffs(stop) - 1;
                    uint32_t m = __shfl_sync(0xffffffffu, mine, first);
                    l0 = l0 + first * 16 + m;
                    break;
                }
                l0 += 512;
            }
            l0 = min(l0, cap0);
            if (lane == src) v[k] = (int32_t)(l0 + p[k]);
        }
    }
    int32_t mx = 0;
#pragma unroll
    for (int k = 0; k < PI; k++) {
        V[g0 + k * IT] = v[k];
        mx = max(mx, v[k]);
    }
    // tile maximum
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if (lane == 0 && mx > 0) atomicMax(&tileMax[g0 / LT], mx);
}

// One CTA per block: exclusive max-scan of the tile maxima, in place.
__global__ void __launch_bounds__(256) k_plcp_scan_tiles(int32_t *__restrict__ tileMax, uint32_t tilesPerBlock)
{
    __shared__ int32_t part[256];
    int32_t *tm = tileMax + (size_t)blockIdx.x * tilesPerBlock;
    uint32_t per = (tilesPerBlock + 255) / 256;
    uint32_t lo = threadIdx.x * per, hi = min(tilesPerBlock, lo + per);
    int32_t s = 0;
    for (uint32_t t = lo; t < hi; t++) s = max(s, tm[t]);
    part[threadIdx.x] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        int32_t run = 0;
        for (int i = 0; i < 256; i++) {
            int32_t c = part[i];
            part[i] = run;
            run = max(run, c);
        }
    }
    __syncthreads();
    s = part[threadIdx.x];
    for (uint32_t t = lo; t < hi; t++) {
        int32_t c = tm[t];
        tm[t] = s;
        s = max(s, c);
    }
}

// PLCP[g] = inclusive-max(V)[g] - p, also scattered to LCP[ISA[g]] (gathering
// LCP[i] = PLCP[SA[i]] in SA order instead measured the same).
__global__ void __launch_bounds__(LT) k_plcp_apply(const uint32_t *__restrict__ blkN,
                                                   const uint64_t *__restrict__ rank, int logS,
                                                   const int32_t *__restrict__ tileMax, int32_t *__restrict__ V,
                                                   int32_t *__restrict__ lcp)
{
    __shared__ int32_t wmax[LT / 32];
    const uint32_t Smask = (1u << logS) - 1;
    const uint32_t g = blockIdx.x * LT + threadIdx.x;
    const uint32_t b = g >> logS, p = g & Smask;
    const uint32_t n = blkN[b] - 8;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int32_t v = (p < n) ? V[g] : 0;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        int32_t t = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o) v = max(v, t);
    }
    if (lane == 31) wmax[warp] = v;
    __syncthreads();
    if (threadIdx.x < 32) {
        int32_t w = wmax[threadIdx.x];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int32_t t = __shfl_up_sync(0xffffffffu, w, o);
            if (lane >= o) w = max(w, t);
        }
        wmax[threadIdx.x] = w;  // inclusive over warps
    }
    __syncthreads();
    int32_t carry = tileMax[blockIdx.x];
    if (warp > 0) carry = max(carry, wmax[warp - 1]);
    v = max(v, carry);
    if (p < n) {
        int32_t pl = v - (int32_t)p;
        V[g] = pl;
        st_hint(&lcp[(g & ~Smask) | ((uint32_t)rank[g] & 0xffffffu)], pl, l2_evict_last());   // 4-byte scatter by rank: sectors complete in L2
    }
}

// level l+1 from level l (count = entries of the coarser level, all blocks).
__global__ void __launch_bounds__(256) k_pyramid_up(const int32_t *__restrict__ inSA,
                                                    const int32_t *__restrict__ inLCP, uint32_t outCount,
                                                    int32_t *__restrict__ outSA, int32_t *__restrict__ outLCP)
{
    const uint32_t s = blockIdx.x * 256 + threadIdx.x;  // index into the finer level
    const bool ok = (s >> 5) < outCount;
    int32_t a = ok ? inSA[s] : 0x7fffffff;
    int32_t l = ok ? inLCP[s] : 0x7fffffff;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        a = min(a, __shfl_xor_sync(0xffffffffu, a, o));
        l = min(l, __shfl_xor_sync(0xffffffffu, l, o));
    }
    if (ok && (threadIdx.x & 31) == 0) {
        outSA[s >> 5] = a;
        outLCP[s >> 5] = l;
    }
}

}  // namespace

int lcp_build(Workspace &ws, const BatchDesc &bd, cudaStream_t st, LaunchCounter &lc)
{
    const size_t total = (size_t)bd.B * bd.S;  // padded index space of this batch
    const uint32_t tiles = (uint32_t)(total / LT);
    const uint32_t tilesPerBlock = bd.S / LT;
    if (cudaMemsetAsync(ws.tileMax, 0, (size_t)tiles * sizeof(int32_t), st) != cudaSuccess) return -1;
    k_plcp_irreducible<<<tiles * (LT / (IT * PI)), IT, 0, st>>>(ws.T, bd.Tstride, ws.blkN, ws.rank, ws.SA, bd.logS, ws.plcp,
                                             ws.tileMax);
    k_plcp_scan_tiles<<<bd.B, 256, 0, st>>>(ws.tileMax, tilesPerBlock);
    k_plcp_apply<<<tiles, LT, 0, st>>>(ws.blkN, ws.rank, bd.logS, ws.tileMax, ws.plcp, ws.lcp);
    lc.launches += 3;
    lc.alg_bytes += (uint64_t)bd.M * (17 + 20);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        fprintf(stderr, "salz_cuda: lcp_build: %s\n", cudaGetErrorString(e));
        return -1;
    }
    return 0;
}

// Pyramid levels 2.. from level 1 (which k_ansv_tile writes: its per-thread segments are
// the groups of 32 slots): level l has (S >> 5l) entries per block; stop once <= 32.
int pyramid_upper_build(Workspace &ws, const BatchDesc &bd, cudaStream_t st, LaunchCounter &lc)
{
    int l = 1;
    while ((bd.S >> (5 * l)) > 32) {
        size_t fine = (size_t)bd.B * (bd.S >> (5 * l));
        uint32_t outCount = (uint32_t)(fine >> 5);
        k_pyramid_up<<<(uint32_t)((fine + 255) / 256), 256, 0, st>>>(ws.pyrSA[l], ws.pyrLCP[l], outCount,
                                                                     ws.pyrSA[l + 1], ws.pyrLCP[l + 1]);
        lc.launches += 1;
        l++;
    }
    return cudaGetLastError() == cudaSuccess ? 0 : -1;
}

// ---- no-allocation entry point -------------------------------------------------------------
// salz_cuda_lcp_noalloc (include/salz_cuda.h): PLCP / LCP of one text from its suffix array with
// the kernels above, for callers that cannot call cudaMalloc (code under CUDA graph capture,
// tight serving loops, harnesses that forbid it).  No counterpart in the reference.
// Its scratch - a zero-padded copy of the text, the inverse
// suffix array, the PLCP values, the tile maxima, the block descriptor - lives in the file-scope
// array below: SALZ_CUDA_NOALLOC_MAX_N positions at most, ~3.3 MiB that every GPU which loads
// the library holds until exit.  In host code the array's NAME is a host-side stand-in, not a
// device address; the address comes from cudaGetSymbolAddress, once per device with that device
// current.  All calls share the array and it keeps the last call's contents: what must start at
// zero is cleared on the caller's stream at every call, and calls that overlap on one GPU race.
#include "../../include/salz_cuda.h"

#include <mutex>

namespace {

constexpr size_t NA_MAXN = SALZ_CUDA_NOALLOC_MAX_N;          // positions; also the largest S
static_assert((NA_MAXN & (NA_MAXN - 1)) == 0 && NA_MAXN >= 4096, "power of two, at least one ANSV tile");
constexpr size_t NA_OFF_TEXT = 0;                              // u8  [S + 256], zero behind the text
constexpr size_t NA_OFF_RANK = NA_OFF_TEXT + NA_MAXN + 256;    // u64 [S]  inverse suffix array
constexpr size_t NA_OFF_V = NA_OFF_RANK + 8 * NA_MAXN;         // i32 [S]  V, then PLCP
constexpr size_t NA_OFF_TMAX = NA_OFF_V + 4 * NA_MAXN;         // i32 [S / LT] tile maxima (start at zero)
constexpr size_t NA_OFF_BLKN = NA_OFF_TMAX + 4 * (NA_MAXN / LT);   // u32 [1] N = n + 8
constexpr size_t NA_BYTES = NA_OFF_BLKN + 256;

static __device__ __align__(256) unsigned char g_noalloc_scratch[NA_BYTES];

// rank[SA[r]] = r (the low 24 bits are what the PLCP kernels read); thread 0 leaves the block
// descriptor.  A suffix-array entry outside [0, n) is skipped and counted.
__global__ void __launch_bounds__(256) k_noalloc_isa(const int32_t *__restrict__ SA, uint32_t n, uint64_t *__restrict__ rank,
                                                     uint32_t *__restrict__ blkN)
{
    const uint32_t r = blockIdx.x * 256 + threadIdx.x;
    if (r == 0) blkN[0] = n + 8;
    if (r < n) {
        const uint32_t p = (uint32_t)SA[r];
        if (p < n) rank[p] = r;
    }
}

std::mutex g_noalloc_mu;
unsigned char *g_noalloc_ptr[64];   // per device: device address of g_noalloc_scratch

}  // namespace

extern "C" int salz_cuda_lcp_noalloc(int device, const uint8_t *d_text, size_t n, const int32_t *d_sa, int32_t *d_plcp,
                                     int32_t *d_lcp, void *stream)
{
    if (device < 0 || device >= 64) {
        fprintf(stderr, "salz_cuda_lcp_noalloc: device %d out of range\n", device);
        return -1;
    }
    if (n > NA_MAXN) {
        fprintf(stderr, "salz_cuda_lcp_noalloc: %zu positions exceed the static scratch (SALZ_CUDA_NOALLOC_MAX_N = %zu); "
                        "use a salz_cuda_ctx for larger texts\n", n, NA_MAXN);
        return -2;
    }
    if (n > 0 && (!d_text || !d_sa || !d_lcp)) {
        fprintf(stderr, "salz_cuda_lcp_noalloc: NULL text, suffix array or LCP buffer\n");
        return -1;
    }
    if (cudaSetDevice(device) != cudaSuccess) {
        fprintf(stderr, "salz_cuda_lcp_noalloc: no CUDA device %d; there is no CPU fallback\n", device);
        return -1;
    }
    unsigned char *base;
    {
        // set-up, once per device, with that device current
        std::lock_guard<std::mutex> lock(g_noalloc_mu);
        if (!g_noalloc_ptr[device]) {
            void *p = nullptr;
            const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
            if (e != cudaSuccess || !p) {
                fprintf(stderr, "salz_cuda_lcp_noalloc: cudaGetSymbolAddress: %s\n", cudaGetErrorString(e));
                return -1;
            }
            g_noalloc_ptr[device] = (unsigned char *)p;
        }
        base = g_noalloc_ptr[device];
    }
    if (n == 0) return 0;   // set-up only (e.g. before a graph capture starts)

    cudaStream_t st = (cudaStream_t)stream;
    int logS = 12;
    while (((size_t)1 << logS) < n) logS++;
    const size_t S = (size_t)1 << logS;
    uint8_t *T = base + NA_OFF_TEXT;
    uint64_t *rank = (uint64_t *)(base + NA_OFF_RANK);
    int32_t *V = (int32_t *)(base + NA_OFF_V);
    int32_t *tileMax = (int32_t *)(base + NA_OFF_TMAX);
    uint32_t *blkN = (uint32_t *)(base + NA_OFF_BLKN);
    const uint32_t tiles = (uint32_t)(S / LT);

    // the array keeps the last call's contents: the text padding and the tile maxima must start at
    // zero; so does the inverse suffix array, so that a malformed d_sa (not a permutation) cannot
    // leave a stale slot behind that indexes past d_sa
    cudaError_t e = cudaMemsetAsync(T, 0, S + 256, st);
    if (e == cudaSuccess) e = cudaMemsetAsync(rank, 0, S * sizeof(uint64_t), st);
    if (e == cudaSuccess) e = cudaMemsetAsync(tileMax, 0, (size_t)tiles * sizeof(int32_t), st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(T, d_text, n, cudaMemcpyDeviceToDevice, st);
    if (e != cudaSuccess) {
        fprintf(stderr, "salz_cuda_lcp_noalloc: %s\n", cudaGetErrorString(e));
        return -1;
    }
    k_noalloc_isa<<<(uint32_t)((n + 255) / 256), 256, 0, st>>>(d_sa, (uint32_t)n, rank, blkN);
    k_plcp_irreducible<<<tiles * (LT / (IT * PI)), IT, 0, st>>>(T, S, blkN, rank, (const uint32_t *)d_sa, logS, V, tileMax);
    k_plcp_scan_tiles<<<1, 256, 0, st>>>(tileMax, tiles);
    k_plcp_apply<<<tiles, LT, 0, st>>>(blkN, rank, logS, tileMax, V, d_lcp);
    e = cudaGetLastError();
    if (e == cudaSuccess && d_plcp) e = cudaMemcpyAsync(d_plcp, V, n * sizeof(int32_t), cudaMemcpyDeviceToDevice, st);
    if (e != cudaSuccess) {
        fprintf(stderr, "salz_cuda_lcp_noalloc: %s\n", cudaGetErrorString(e));
        return -1;
    }
    return 0;
}
