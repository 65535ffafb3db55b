// parse.cu - the cost-optimal LZ parse.
//
//   k_parse: right-to-left shortest-path DP with the reference's cost model and
//            strict '<' tie-break order literal, PSV, NSV (ref lib/salz.c:562-662).
//   (k_choose also leaves the chunk-exit pointers of the chosen parse for the token walk in emit.cu.)
#include "salz_dev.cuh"

#include <algorithm>
#include <stdlib.h>

namespace {

// ---------------------------------------------------------------------------
// Optimal parse.  The cost chain
//     cost[p] = min(9 + cost[p+1], w1 + cost[p+l1], w2 + cost[p+l2])
// is serial, so ONE warp (the consumer) walks it position by position, right
// to left, and does nothing else; every lane runs the same scalar recurrence.
// In the shifted space d[p] = cost[p] + 9p the literal edge costs nothing:
//     d[p] = min(d[p+1], F[p], d[t1] + w1 - 9 l1, d[t2] + w2 - 9 l2)
// so the chain is one min per position; F (edges whose targets are at least
// two chunks to the right: final, read from global memory) and the two ring
// addresses/weights of the nearer targets are prepared by the stager warps one
// chunk (PC positions) ahead, in shared memory.  The consumer's ring loads for
// position p-2 are issued before position p is finished (a factor is at least
// 3 long, so they never wait on a value still being computed).
//   flush(c+1)  ring -> global d array, by the stager warps
//   stage(c-1)  consumer records of chunk c-1
// k_choose then derives the choice per position from the d array, fully parallel.
//
// Segments.  With few blocks in a batch one chain per block leaves the GPU
// idle, so a block is cut into K segments of 2^segLog positions, one CTA each.
// Every segment keeps d relative to its own right end (dr[p] = d[p] - base[j]);
// what lies to the right of a segment is read as an ESTIMATE (the values other
// CTAs have in dr[] at that moment plus the bases of the last scan).  Passes
// are repeated until k_choose finds the recurrence satisfied at every position
// (or a pass changes nothing): the recurrence has exactly one solution, so that
// is the reference's.  Pass i leaves at least the i rightmost segments exact (at
// most K+1 passes); text-like inputs are exact after two because only relative
// costs matter and the optimal paths of neighbouring positions merge quickly.
// K = 1 is the plain single pass (nothing is estimated).  The first of the passes
// only has to supply the left end of every segment to the second one, so it walks
// just WARM_CHUNKS chunks per segment (see parse_build).
// ---------------------------------------------------------------------------
#ifndef SALZ_PC
#define SALZ_PC 256
#endif
#ifndef SALZ_HELPERS
#define SALZ_HELPERS 32
#endif
#ifndef SALZ_PARSE_CTAS
#define SALZ_PARSE_CTAS 4
#endif
constexpr int PC = SALZ_PC;        // positions per staged chunk (long enough to hide the helpers' memory latency)
constexpr int HELPERS = SALZ_HELPERS;  // stager threads
#ifndef SALZ_PER
#define SALZ_PER 8
#endif
constexpr int PER = SALZ_PER;  // positions a helper thread has in flight
constexpr int SUBS = PC / (HELPERS * PER);  // batches of PER positions per helper thread and chunk
#ifndef SALZ_NCH
#define SALZ_NCH 4
#endif
constexpr int NCH = SALZ_NCH;        // chains per CTA (lanes of the consumer warp are split between them)
constexpr int PARSE_WARPS = 1 + NCH * HELPERS / 32;
constexpr int PARSE_THREADS = 32 * PARSE_WARPS;  // consumer warp + stagers of every chain
constexpr int NOT_FAR = -1;
constexpr int RING = 2 * PC;       // ring entries; entry RING is a constant INF (absent edge)
constexpr int KMAX = 64;           // segments per block, at most
#ifndef SALZ_WARM
#define SALZ_WARM 8
#endif
constexpr uint32_t WARM_CHUNKS = SALZ_WARM;  // chunks per segment the warm-up pass walks (0: a full first pass)
constexpr size_t PARSE_SMEM = NCH * (2 * PC * sizeof(int4) + (RING + 4) * sizeof(int32_t));

struct SegCtx {
    const int32_t *dr;     // block's d array of this pass (own segment), relative per segment
    const int32_t *drOld;  // the previous pass's array (everything right of the segment)
    const int32_t *sbase;  // [K - j] base[m] - base[j] (shared memory)
    uint32_t R;            // first position right of the segment
    uint32_t j;
    int segLog;
};

// d[t] relative to this segment's base, t anywhere in the block
__device__ __forceinline__ int32_t dval(const SegCtx &sg, uint32_t t)
{
    return (t < sg.R) ? __ldcg(&sg.dr[t]) : __ldcg(&sg.drOld[t]) + sg.sbase[(t >> sg.segLog) - sg.j];
}

// Consumer record: x = F in d-space, y / z = weight - 9 len of the two ring
// edges, w = byte offsets of their ring entries (low / high 16 bits).
__device__ __forceinline__ void stage_chunk(const int4 *__restrict__ cd, const SegCtx &sg, uint32_t n,
                                            uint32_t chunk, int4 *sc, int tid)
{
    const uint32_t farFrom = min((chunk + 2) * PC, sg.R);  // targets >= this are read from global memory
#pragma unroll 1
  for (int sub = 0; sub < SUBS; sub++) {
    const uint32_t base = chunk * PC + sub * (HELPERS * PER);
    sc += (sub ? HELPERS * PER : 0);
    int4 c[PER];
#pragma unroll
    for (int k = 0; k < PER; k++) {
        const uint32_t p = base + tid + k * HELPERS;
        c[k] = (p >= 1 && p < n) ? cd[p] : make_int4(0, 0, 0, 0);
    }
    int32_t t1[PER], t2[PER];
#pragma unroll
    for (int k = 0; k < PER; k++) {  // all far loads in flight together
        const uint32_t p = base + tid + k * HELPERS;
        const uint32_t ta = p + (uint32_t)c[k].y, tb = p + (uint32_t)c[k].w;
        t1[k] = (c[k].y >= 3 && ta >= farFrom) ? dval(sg, ta) : NOT_FAR;
        t2[k] = (c[k].w >= 3 && tb >= farFrom) ? dval(sg, tb) : NOT_FAR;
    }
#pragma unroll
    for (int k = 0; k < PER; k++) {
        const uint32_t p = base + tid + k * HELPERS;
        int32_t F = (p == n) ? 0 : SALZ_INF;  // cost[n] = 0 (ref lib/salz.c:621): d[n] = 9n = the last segment's base
        int32_t w1 = 0, w2 = 0;
        uint32_t o1 = RING * 4, o2 = RING * 4;
        if (c[k].y >= 3) {
            const int w = factor_cost_bits((uint32_t)c[k].x, (uint32_t)c[k].y) - 9 * c[k].y;
            if (p + (uint32_t)c[k].y >= farFrom) F = min(F, w + t1[k]);
            else { w1 = w; o1 = ((p + (uint32_t)c[k].y) & (RING - 1)) * 4; }
        }
        if (c[k].w >= 3) {
            const int w = factor_cost_bits((uint32_t)c[k].z, (uint32_t)c[k].w) - 9 * c[k].w;
            if (p + (uint32_t)c[k].w >= farFrom) F = min(F, w + t2[k]);
            else { w2 = w; o2 = ((p + (uint32_t)c[k].w) & (RING - 1)) * 4; }
        }
        sc[tid + k * HELPERS] = make_int4(F, w1, w2, (int)(o1 | (o2 << 16)));
    }
  }
}

// ring -> global d array for chunk c (needed by stage(c-2) and k_choose); ht in [0, HELPERS).
__device__ __forceinline__ void flush_chunk(const int32_t *ring, int32_t *drb, uint32_t n, uint32_t c, int ht)
{
#pragma unroll
    for (int k = 0; k < PC / HELPERS; k++) {
        const int i = ht + k * HELPERS;
        const uint32_t p = c * PC + i;
        if (p <= n) drb[p] = ring[(c & 1u) * PC + i];
    }
}
__device__ __forceinline__ void helper_barrier(int id) { asm volatile("bar.sync %0, %1;" ::"r"(id), "n"(HELPERS) : "memory"); }

__device__ __forceinline__ int32_t lds_i32(uint32_t addr)
{
    int32_t v;
    asm volatile("ld.shared.s32 %0, [%1];" : "=r"(v) : "r"(addr));
    return v;
}
__device__ __forceinline__ void sts_i32(uint32_t addr, int32_t v)
{
    asm volatile("st.shared.s32 [%0], %1;" ::"r"(addr), "r"(v));
}

// One CTA runs NCH chains: the consumer warp gives each chain 32 / NCH lanes (the lanes
// of a chain run the same scalar recurrence; different chains share the instruction
// stream, which is what the SM runs out of first), and every chain has its own
// HELPERS stager threads, ring and record buffers.  Chain q = b * K + j is segment j of
// block b; grid = ceil(B * K / NCH).  dr: d values by g, relative per segment;
// segBase: [B][KMAX].
struct Chain {
    uint32_t b, j, N, n, R, cBeg, cEnd;
    bool act, last;
};

// warm > 0: only the leftmost `warm` chunks of the segment (the warm-up pass).
__device__ __forceinline__ Chain chain_of(uint32_t q, uint32_t chains, int kLog, int segLog, uint32_t warm,
                                          const uint32_t *__restrict__ blkN)
{
    Chain ch;
    ch.act = false;
    ch.b = ch.j = ch.N = ch.n = ch.R = ch.cBeg = ch.cEnd = 0;
    ch.last = false;
    if (q >= chains) return ch;
    ch.b = q >> kLog;
    ch.j = q & ((1u << kLog) - 1);
    ch.N = blkN[ch.b];
    ch.n = ch.N - 8;
    const uint32_t L = ch.j << segLog;
    if (L >= ch.N) return ch;
    ch.R = min(ch.N, L + (warm ? min(1u << segLog, warm * PC) : (1u << segLog)));
    ch.last = ch.R == ch.N;
    ch.cBeg = L / PC;
    ch.cEnd = (ch.R + PC - 1) / PC;  // segment lengths are multiples of PC
    ch.act = true;
    return ch;
}

__global__ void __launch_bounds__(PARSE_THREADS, SALZ_PARSE_CTAS) k_parse(const int4 *__restrict__ cand,
                                                            const uint32_t *__restrict__ blkN, int logS, int segLog,
                                                            int kLog, uint32_t chains, uint32_t warm, int numSMs, int32_t *dr,
                                                            const int32_t *drOld,
                                                            const int32_t *__restrict__ segBase)
{
    extern __shared__ __align__(16) uint8_t psmem[];
    int4 *stC = (int4 *)psmem;                          // [NCH][2][PC] consumer records
    int32_t *rings = (int32_t *)(stC + NCH * 2 * PC);   // [NCH][RING + 4], d-space
    __shared__ int32_t sbase[NCH][KMAX];
    // co-resident CTAs put their consumer on different warp slots
    const uint32_t lin = blockIdx.x;
    const int cw = (int)((lin + lin / (uint32_t)numSMs) % 4u);
    const int warp = ((int)(threadIdx.x >> 5) + PARSE_WARPS - cw) % PARSE_WARPS;  // role: 0 consumer, 1.. helpers
    const int lane = threadIdx.x & 31;
    // the chain this thread works for
    const int ci = (warp == 0) ? lane / (32 / NCH) : (warp - 1) / (HELPERS / 32);
    const int ht = ((warp - 1) % (HELPERS / 32)) * 32 + lane;  // helper thread index inside the chain
    const Chain ch = chain_of(blockIdx.x * NCH + ci, chains, kLog, segLog, warm, blkN);
    uint32_t maxSteps = 0;
#pragma unroll
    for (int k = 0; k < NCH; k++) {
        const Chain o = chain_of(blockIdx.x * NCH + k, chains, kLog, segLog, warm, blkN);
        if (o.act) maxSteps = max(maxSteps, o.cEnd - o.cBeg);
    }
    if (maxSteps == 0) return;

    const size_t gb = (size_t)ch.b << logS;
    const int4 *cd = cand + gb;
    int32_t *drb = dr + gb;
    int32_t *ring = rings + ci * (RING + 4);
    int4 *myC = stC + ci * 2 * PC;
    const uint32_t ringAddr = (uint32_t)__cvta_generic_to_shared(ring);

    for (int i = threadIdx.x; i < NCH * KMAX; i += PARSE_THREADS) {
        const int k = i / KMAX;
        const Chain o = chain_of(blockIdx.x * NCH + k, chains, kLog, segLog, warm, blkN);
        const uint32_t m = o.j + (uint32_t)(i % KMAX);
        sbase[k][i % KMAX] = (o.act && m < KMAX && ((size_t)m << segLog) < o.N)
                                 ? segBase[o.b * KMAX + m] - segBase[o.b * KMAX + o.j] : 0;
    }
    for (int i = threadIdx.x; i < NCH * (RING + 4); i += PARSE_THREADS) rings[i] = SALZ_INF;
    // lanes of a chain without work idle along with the others: their records must hold loadable addresses
    {
        for (int i = threadIdx.x; i < NCH * 2 * PC; i += PARSE_THREADS) {
            stC[i] = make_int4(SALZ_INF, 0, 0, 0);
        }
    }
    __syncthreads();
    SegCtx sg = {drb, drOld + gb, sbase[ci], ch.R, ch.j, segLog};
    const uint32_t nch = ch.act ? ch.cEnd - ch.cBeg : 0;

    if (warp >= 1 && ch.act) {
        const uint32_t c = ch.cEnd - 1;
        stage_chunk(cd, sg, ch.n, c, myC + (c & 1) * PC, ht);
    }
    __syncthreads();

    int32_t d = ch.last ? SALZ_INF : 0;  // d[p + 1] relative to the segment's base
    // step s: the consumer walks chunk cEnd-1-s of every chain that still has one; the
    // stagers flush the chunk it has just left and prepare the one after
    for (uint32_t s = 0; s <= maxSteps; s++) {
        const uint32_t c = ch.cEnd - 1 - min(s, nch ? nch - 1 : 0u);
        if (warp >= 1) {
            if (ch.act && s >= 1 && s <= nch) {
                flush_chunk(ring, drb, ch.n, ch.cEnd - s, ht);
                helper_barrier(1 + ci);
            }
            if (ch.act && s + 1 < nch)
                stage_chunk(cd, sg, ch.n, c - 1, myC + ((c - 1) & 1) * PC, ht);
        } else {
            const bool live = s < nch;
            if (__any_sync(0xffffffffu, live)) {
                const int4 *rc = myC + (c & 1) * PC;
                // ring slot of position c*PC + i is ((c & 1) * PC + i)
                const uint32_t slot0 = ringAddr + (c & 1u) * PC * 4;
                int4 r0 = rc[PC - 1], r1 = rc[PC - 2];
                int32_t a0 = lds_i32(ringAddr + ((uint32_t)r0.w & 0xffffu)), b0 = lds_i32(ringAddr + ((uint32_t)r0.w >> 16));
                int32_t a1 = lds_i32(ringAddr + ((uint32_t)r1.w & 0xffffu)), b1 = lds_i32(ringAddr + ((uint32_t)r1.w >> 16));
                // Records of positions i-2 .. i-2-HB+1 are fetched one half-body ahead of their
                // use; the ring loads of position j-2 go out before position j is finished (its
                // targets are >= j+1: all stored).
                constexpr int HB = 4;
                int4 X[HB], Y[HB];
#pragma unroll
                for (int k = 0; k < HB; k++) X[k] = rc[(PC - 3 - k) & (PC - 1)];
                auto half = [&](int4 (&use)[HB], int4 (&load)[HB], int i) {
#pragma unroll
                    for (int k = 0; k < HB; k++) load[k] = rc[(i - k - 2 - HB) & (PC - 1)];
#pragma unroll
                    for (int k = 0; k < HB; k++) {
                        const int4 r2 = use[k];
                        const int32_t a2 = lds_i32(ringAddr + ((uint32_t)r2.w & 0xffffu));
                        const int32_t b2 = lds_i32(ringAddr + ((uint32_t)r2.w >> 16));
                        const int32_t nd = min(d, min(r0.x, min(a0 + r0.y, b0 + r0.z)));
                        if (live) {  // a finished chain's lanes idle along (records may be stale)
                            d = nd;
                            sts_i32(slot0 + (uint32_t)(i - k) * 4, d);  // the chain's lanes: same address, same value
                        }
                        r0 = r1; a0 = a1; b0 = b1;
                        r1 = r2; a1 = a2; b1 = b2;
                    }
                };
#pragma unroll 1
                for (int i = PC - 1; i >= 0; i -= 2 * HB) {
                    half(X, Y, i);
                    half(Y, X, i - HB);
                }
            }
        }
        __syncthreads();
    }
}

// Heuristic only (the result never depends on it): every 64th position votes whether one of
// its candidates is longer than `longLen`.  Factors that reach over many segments are what
// keeps the segmented passes from settling.  counters[10] += votes for `longLen`,
// counters[11] += votes for `midLen` (factors longer than the warm-up pass sees).
__global__ void __launch_bounds__(256) k_longcount(const int4 *__restrict__ cand, const uint32_t *__restrict__ blkN,
                                                   int logS, uint32_t total, int longLen, int midLen,
                                                   uint32_t *__restrict__ counters)
{
    const uint32_t i = blockIdx.x * 256 + threadIdx.x;
    bool vote = false, vote2 = false;
    if (i < total) {
        const uint32_t g = i * 64 + 17;
        const uint32_t b = g >> logS, p = g & ((1u << logS) - 1);
        if (p >= 1 && p + 8 < blkN[b]) {
            const int4 c = cand[g];
            vote = c.y >= longLen || c.w >= longLen;
            vote2 = c.y >= midLen || c.w >= midLen;
        }
    }
    const unsigned m = __ballot_sync(0xffffffffu, vote), m2 = __ballot_sync(0xffffffffu, vote2);
    if (m && (threadIdx.x & 31) == 0) atomicAdd(&counters[10], __popc(m));
    if (m2 && (threadIdx.x & 31) == 0) atomicAdd(&counters[11], __popc(m2));
}

// Before the first pass: every base = 9 n_b (the all-literal estimate in d-space,
// and the exact base of the last segment).
__global__ void k_seg_init(const uint32_t *__restrict__ blkN, int B, int32_t *__restrict__ segBase)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= B * KMAX) return;
    segBase[i] = 9 * (int32_t)(blkN[i / KMAX] - 8);
}

// After a pass: bases (absolute d at the segments' right ends) by a suffix sum of the d
// values at the segment starts.
__global__ void k_seg_scan(const uint32_t *__restrict__ blkN, int B, int logS, int segLog,
                           const int32_t *__restrict__ dr, int32_t *__restrict__ segBase)
{
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    const uint32_t N = blkN[b];
    const int Kb = (int)((N + (1u << segLog) - 1) >> segLog);
    const int32_t *drb = dr + ((size_t)b << logS);
    int32_t base = 9 * (int32_t)(N - 8);
    for (int j = Kb - 1; j >= 0; j--) {
        segBase[b * KMAX + j] = base;
        base += drb[(size_t)j << segLog];  // d at the first position of segment j = base of segment j-1
    }
}

// The choice per position from the d values in global memory, with the reference's
// comparisons in its order (strict '<': literal, PSV, NSV; ref lib/salz.c:622-661; a
// common -9p drops out in d-space) - and at the same time the proof that the values
// are final: every position's recurrence must hold on the stored values.  The
// recurrence is evaluated right to left over a DAG, so an array that satisfies it
// everywhere IS its one solution, however it was arrived at.  counters[9] += misses.
__global__ void __launch_bounds__(256) k_choose(const int4 *__restrict__ cand, const uint32_t *__restrict__ blkN,
                                                int logS, int segLog, const int32_t *__restrict__ dr,
                                                const int32_t *__restrict__ segBase, int2 *__restrict__ choice,
                                                uint32_t *__restrict__ jump, uint32_t *__restrict__ counters)
{
    // One CTA per PARSE_CHUNK positions.  Second half: jump[p] = first position >= the end
    // of p's chunk that the chosen parse reaches from p (pointer jumping inside the chunk;
    // used by the token walk in emit.cu).
    __shared__ uint32_t nxt[PARSE_CHUNK];
    const uint32_t chunksPerBlock = 1u << (logS - PARSE_CHUNK_LOG);
    const uint32_t b = blockIdx.x / chunksPerBlock, c = blockIdx.x % chunksPerBlock;
    const uint32_t N = blkN[b];
    const uint32_t base = c << PARSE_CHUNK_LOG;
    if (base >= N) return;
    const uint32_t end = min(N, base + PARSE_CHUNK);
    const uint32_t n = N - 8;
    const size_t gb = (size_t)b << logS;
    const int32_t *drb = dr + gb;
    const int32_t *sb = segBase + b * KMAX;
    constexpr int PER = PARSE_CHUNK / 256;
    uint32_t v[PER];
    int4 cd[PER];
    int32_t mine[PER], lit[PER];
#pragma unroll
    for (int i = 0; i < PER; i++) {
        const uint32_t p = base + threadIdx.x + i * 256;
        cd[i] = (p >= 1 && p < n) ? cand[gb + p] : make_int4(0, 0, 0, 0);  // position 0 is a forced literal
        mine[i] = (p <= n) ? drb[p] + sb[p >> segLog] : 0;
        lit[i] = (p < n) ? drb[p + 1] + sb[(p + 1) >> segLog] : 9 * (int32_t)n;  // d[n] = cost[n] + 9n = 9n (ref lib/salz.c:621)
    }
    bool bad = false;
#pragma unroll
    for (int i = 0; i < PER; i++) {
        const uint32_t p = base + threadIdx.x + i * 256;
        uint32_t bo = 0, bl = 1;
        if (p <= n) {
            int32_t best = lit[i];
            if (cd[i].y >= 3) {
                const uint32_t t = p + (uint32_t)cd[i].y;
                const int32_t w = drb[t] + sb[t >> segLog] + factor_cost_bits((uint32_t)cd[i].x, (uint32_t)cd[i].y) - 9 * cd[i].y;
                if (w < best) { best = w; bo = (uint32_t)cd[i].x; bl = (uint32_t)cd[i].y; }
            }
            if (cd[i].w >= 3) {
                const uint32_t t = p + (uint32_t)cd[i].w;
                const int32_t w = drb[t] + sb[t >> segLog] + factor_cost_bits((uint32_t)cd[i].z, (uint32_t)cd[i].w) - 9 * cd[i].w;
                if (w < best) { best = w; bo = (uint32_t)cd[i].z; bl = (uint32_t)cd[i].w; }
            }
            bad |= mine[i] != best;
        }
        if (p < end) choice[gb + p] = make_int2((int)bo, (int)bl);
        v[i] = (p < end) ? p + bl : end;
        nxt[threadIdx.x + i * 256] = v[i];
    }
    const unsigned m = __ballot_sync(0xffffffffu, bad);
    if (m && (threadIdx.x & 31) == 0) atomicAdd(&counters[9], __popc(m));
    __syncthreads();
    for (int it = 0; it < PARSE_CHUNK_LOG; it++) {
        bool moved = false;
#pragma unroll
        for (int i = 0; i < PER; i++)
            if (v[i] < end) { v[i] = nxt[v[i] - base]; moved = true; }
        if (!__syncthreads_or(moved)) break;
#pragma unroll
        for (int i = 0; i < PER; i++) nxt[threadIdx.x + i * 256] = v[i];
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < PER; i++) {
        const uint32_t p = base + threadIdx.x + i * 256;
        if (p < end) jump[gb + p] = v[i];
    }
}

}  // namespace

// Per-device kernel attributes; called by salz_cuda_ctx_create after cudaSetDevice.
int parse_init_device()
{
    if (cudaFuncSetAttribute(k_parse, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)PARSE_SMEM) != cudaSuccess) return -1;
    if (cudaFuncSetAttribute(k_parse, cudaFuncAttributePreferredSharedMemoryCarveout, 100) != cudaSuccess) return -1;
    return 0;
}

int parse_build(Workspace &ws, const BatchDesc &bd, cudaStream_t st, LaunchCounter &lc)
{
    // Segments per block: several chains per slot (NCH chains per CTA, SALZ_PARSE_CTAS CTAs
    // per SM) so that the slots stay busy to the end of a pass, segments of at least 64 chunks (16384)
    // positions); a cut only pays when it is at least 4-fold (two passes of the chains +
    // the parallel check against one pass).
    int segLog = bd.logS, K = 1;
    uint32_t warm = 0;
    {
        const int slots = ws.numSMs * SALZ_PARSE_CTAS * NCH;
        const int userK = abs(ws.parseSegments);  // < 0: that many segments and the warm-up pass (tests)
        int want = userK ? userK : 8 * slots / std::max(1, bd.B);
        if (!userK && want < 4) want = 1;
        const uint32_t minSeg = userK ? 16u * PC : 64u * PC;
        if (!userK && want >= 4) {
            // inputs with long factors (deep repeats, runs) settle one segment per pass: keep them whole
            const uint32_t samples = (uint32_t)(((size_t)bd.B << bd.logS) / 64);
            if (cudaMemsetAsync(ws.counters + 10, 0, 8, st) != cudaSuccess) return -1;
            k_longcount<<<(samples + 255) / 256, 256, 0, st>>>(ws.cand, ws.blkN, bd.logS, samples, 4096,
                                                               WARM_CHUNKS * PC / 4, ws.counters);
            lc.launches += 1;
            if (cudaMemcpyAsync(ws.h_counters + 10, ws.counters + 10, 8, cudaMemcpyDeviceToHost, st) != cudaSuccess) return -1;
            if (cudaStreamSynchronize(st) != cudaSuccess) return -1;
            if ((uint64_t)ws.h_counters[10] * 2000 > bd.M / 64) want = 1;
            if ((uint64_t)ws.h_counters[11] * 2000 <= bd.M / 64) warm = WARM_CHUNKS;
        }
        while (2 * K <= want && K < KMAX && (bd.S >> 1) >> (bd.logS - segLog) >= minSeg) { K *= 2; segLog--; }
        if ((1u << segLog) < 4u * WARM_CHUNKS * PC) warm = 0;  // short segments: the full pass costs little more
        if (ws.parseSegments < 0) warm = std::max(1u, std::min(WARM_CHUNKS, (1u << segLog) / PC / 2));
    }
    // The d values stay relative to the segment bases: cost[p] = dr[p] + segBase[p >> segLog] - 9p
    // (salz_cuda_debug_stages converts; nothing downstream needs the absolute costs).
    // Two arrays: a pass reads what lies right of a segment from the previous pass's array.
    int32_t *cur = ws.cost, *old = ws.cost2;
    const uint32_t chooseGrid = (uint32_t)bd.B << (bd.logS - PARSE_CHUNK_LOG);
    const size_t arrBytes = ((size_t)bd.B << bd.logS) * sizeof(int32_t);
    bool done = false;
    k_seg_init<<<(bd.B * KMAX + 255) / 256, 256, 0, st>>>(ws.blkN, bd.B, ws.segBase);
    lc.launches += 1;
    if (K > 1) {
        int kLog = 0;
        while ((1 << kLog) < K) kLog++;
        const uint32_t chains = (uint32_t)bd.B << kLog;
        if (cudaMemsetAsync(old, 0, arrBytes, st) != cudaSuccess) return -1;  // all-literal estimate: d = base everywhere
        if (warm && cudaMemsetAsync(cur, 0, arrBytes, st) != cudaSuccess) return -1;  // the warm-up pass writes its windows only
        // Text settles in two passes, and what the second pass reads of the first is the left
        // end of every segment: the first pass walks only the leftmost WARM_CHUNKS chunks of a
        // segment (its values are relative to the end of that window; the bases k_seg_scan
        // derives from them are consistent with that).  An input that has not settled after
        // MAX_PASSES (factors reaching over many segments: one exact segment more per pass) gets
        // the plain one-chain-per-block pass.
        const int MAX_PASSES = warm ? 4 : 3;
        for (int pass = 1; pass <= MAX_PASSES && !done; pass++) {
            k_parse<<<(chains + NCH - 1) / NCH, PARSE_THREADS, PARSE_SMEM, st>>>(ws.cand, ws.blkN, bd.logS, segLog, kLog, chains,
                                                                                 pass == 1 ? warm : 0u, ws.numSMs, cur, old,
                                                                                 ws.segBase);
            k_seg_scan<<<(bd.B + 63) / 64, 64, 0, st>>>(ws.blkN, bd.B, bd.logS, segLog, cur, ws.segBase);
            lc.launches += 2;
            // first pass with a warm-up window walks `warm` chunks of every segment only
            lc.alg_bytes += (pass == 1 && warm) ? (uint64_t)chains * warm * PC * 20 : (uint64_t)bd.M * 20;
            if (pass >= 2) {
                if (cudaMemsetAsync(ws.counters + 9, 0, 4, st) != cudaSuccess) return -1;
                k_choose<<<chooseGrid, 256, 0, st>>>(ws.cand, ws.blkN, bd.logS, segLog, cur, ws.segBase, ws.choice, ws.jump, ws.counters);
                lc.launches += 1;
                lc.alg_bytes += (uint64_t)bd.M * 32;
                if (cudaMemcpyAsync(ws.h_counters + 9, ws.counters + 9, 4, cudaMemcpyDeviceToHost, st) != cudaSuccess) return -1;
                if (cudaStreamSynchronize(st) != cudaSuccess) return -1;
                done = ws.h_counters[9] == 0;
                if (!done && ws.h_counters[9] > 16u * (uint32_t)bd.B) break;  // far from settled: no third pass
            }
            if (!done) std::swap(cur, old);
        }
        if (!done) {
            segLog = bd.logS;
            K = 1;
            k_seg_init<<<(bd.B * KMAX + 255) / 256, 256, 0, st>>>(ws.blkN, bd.B, ws.segBase);
            lc.launches += 1;
        }
    }
    if (!done) {
        k_parse<<<(bd.B + NCH - 1) / NCH, PARSE_THREADS, PARSE_SMEM, st>>>(ws.cand, ws.blkN, bd.logS, segLog, 0, (uint32_t)bd.B,
                                                                           0u, ws.numSMs, cur, old, ws.segBase);
        k_choose<<<chooseGrid, 256, 0, st>>>(ws.cand, ws.blkN, bd.logS, segLog, cur, ws.segBase, ws.choice, ws.jump, ws.counters);
        lc.launches += 2;
        lc.alg_bytes += (uint64_t)bd.M * (20 + 32);
    }
    ws.parseSegLog = segLog;
    ws.parseDr = cur;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        fprintf(stderr, "salz_cuda: parse_build: %s\n", cudaGetErrorString(e));
        return -1;
    }
    return 0;
}
