// parse.cu - the cost-optimal LZ parse.
//
//   k_parse: right-to-left shortest-path DP with the reference's cost model and
//            strict '<' tie-break order literal, PSV, NSV (ref lib/salz.c:562-662).
//   k_jump : chunk-exit pointers of the chosen parse for the token walk in emit.cu.
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
//   stage(c-1)  consumer records of chunk c-1
//   finish(c+1) the choice per position, recomputed from the final costs with
//               the reference's comparison order (strict '<': literal, PSV,
//               NSV; ref lib/salz.c:622-661), and written out
//   flush(c)    ring -> global d array, by the helper warps
//
// Segments.  With few blocks in a batch one chain per block leaves the GPU
// idle, so a block is cut into K segments of 2^segLog positions, one CTA each.
// Every segment keeps d relative to its own right end (dr[p] = d[p] - base[j]);
// what lies to the right of a segment is read as an ESTIMATE (the values other
// CTAs have in dr[] at that moment plus the bases of the last scan).  Passes
// are repeated until a pass changes nothing: then every segment was computed
// from final values, and since the recurrence has exactly one solution that is
// the reference's.  Pass i leaves at least the i rightmost segments exact (at
// most K+1 passes); text-like inputs settle after two because only relative
// costs matter and the optimal paths of neighbouring positions merge quickly.
// K = 1 is the plain single pass (nothing is estimated).
// ---------------------------------------------------------------------------
#ifndef SALZ_PC
#define SALZ_PC 1024
#endif
#ifndef SALZ_HELPERS
#define SALZ_HELPERS 64
#endif
#ifndef SALZ_PARSE_CTAS
#define SALZ_PARSE_CTAS 4
#endif
constexpr int PC = SALZ_PC;        // positions per staged chunk (long enough to hide the helpers' memory latency)
constexpr int HELPERS = SALZ_HELPERS;  // threads per helper group (stagers, finishers)
constexpr int PER = 8;             // positions a helper thread has in flight
constexpr int SUBS = PC / (HELPERS * PER);  // batches of PER positions per helper thread and chunk
constexpr int PARSE_WARPS = 1 + 2 * HELPERS / 32;
constexpr int PARSE_THREADS = 32 * PARSE_WARPS;  // consumer warp + stagers + finishers
constexpr int NOT_FAR = -1;
constexpr int RING = 2 * PC;       // ring entries; entry RING is a constant INF (absent edge)
constexpr int KMAX = 64;           // segments per block, at most
constexpr size_t PARSE_SMEM = 2 * PC * (sizeof(int4) + sizeof(uint32_t)) + (RING + 4) * sizeof(int32_t);

struct SegCtx {
    const int32_t *dr;     // block's d array, relative per segment
    const int32_t *sbase;  // [K - j] base[m] - base[j] (shared memory)
    uint32_t R;            // first position right of the segment
    uint32_t j;
    int segLog;
};

// d[t] relative to this segment's base, t anywhere in the block
__device__ __forceinline__ int32_t dval(const SegCtx &sg, uint32_t t)
{
    const int32_t v = __ldcg(&sg.dr[t]);
    return (t < sg.R) ? v : v + sg.sbase[(t >> sg.segLog) - sg.j];
}

// Consumer record: x = F in d-space, y / z = weight - 9 len of the two ring
// edges, w and sd[] = shared-memory addresses of their ring entries.
__device__ __forceinline__ void stage_chunk(const int4 *__restrict__ cd, const SegCtx &sg, uint32_t n,
                                            uint32_t chunk, int4 *sc, uint32_t *sd, uint32_t ringAddr, int tid)
{
    const uint32_t farFrom = min((chunk + 2) * PC, sg.R);  // targets >= this are read from global memory
#pragma unroll 1
  for (int sub = 0; sub < SUBS; sub++) {
    const uint32_t base = chunk * PC + sub * (HELPERS * PER);
    sc += (sub ? HELPERS * PER : 0);
    sd += (sub ? HELPERS * PER : 0);
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
        sc[tid + k * HELPERS] = make_int4(F, w1, w2, (int)(ringAddr + o1));
        sd[tid + k * HELPERS] = ringAddr + o2;
    }
  }
}

// Choices of a chunk whose d values (and every value to its right in the
// segment) are in global memory: the reference's comparisons, in its order
// (ref lib/salz.c:622-661), on d-space values (a common -9p drops out).
__device__ __forceinline__ void finish_chunk(const int4 *__restrict__ cd, const SegCtx &sg, uint32_t N, uint32_t n,
                                             uint32_t chunk, int2 *__restrict__ ch, int tid)
{
#pragma unroll 1
  for (int sub = 0; sub < SUBS; sub++) {
    const uint32_t base = chunk * PC + sub * (HELPERS * PER);
    int4 c[PER];
    int32_t r0[PER], r1[PER], r2[PER];
#pragma unroll
    for (int k = 0; k < PER; k++) {
        const uint32_t p = base + tid + k * HELPERS;
        c[k] = (p >= 1 && p < n) ? cd[p] : make_int4(0, 0, 0, 0);  // position 0 and the 8-byte tail are forced literals
    }
#pragma unroll
    for (int k = 0; k < PER; k++) {  // all loads in flight together
        const uint32_t p = base + tid + k * HELPERS;
        const bool in = p >= 1 && p < n;
        r0[k] = in ? dval(sg, p + 1) : 0;
        r1[k] = (in && c[k].y >= 3) ? dval(sg, p + (uint32_t)c[k].y) : 0;
        r2[k] = (in && c[k].w >= 3) ? dval(sg, p + (uint32_t)c[k].w) : 0;
    }
#pragma unroll
    for (int k = 0; k < PER; k++) {
        const uint32_t p = base + tid + k * HELPERS;
        if (p >= N) break;
        uint32_t bo = 0, bl = 1;
        if (p >= 1 && p < n) {
            int32_t best = r0[k];
            if (c[k].y >= 3) {
                const int32_t c1 = factor_cost_bits((uint32_t)c[k].x, (uint32_t)c[k].y) - 9 * c[k].y + r1[k];
                if (c1 < best) { best = c1; bo = (uint32_t)c[k].x; bl = (uint32_t)c[k].y; }
            }
            if (c[k].w >= 3) {
                const int32_t c2 = factor_cost_bits((uint32_t)c[k].z, (uint32_t)c[k].w) - 9 * c[k].w + r2[k];
                if (c2 < best) { best = c2; bo = (uint32_t)c[k].z; bl = (uint32_t)c[k].w; }
            }
        }
        ch[p] = make_int2((int)bo, (int)bl);
    }
  }
}

// ring -> global d array for chunk c (needed by stage(c-2) and finish(c)); ht in [0, 2 HELPERS).
// Returns true when a value differs from what the array held.
__device__ __forceinline__ bool flush_chunk(const int32_t *ring, int32_t *drb, uint32_t n, uint32_t c, int ht)
{
    constexpr int FP = PC / (2 * HELPERS);
    int32_t old[FP];
#pragma unroll
    for (int k = 0; k < FP; k++) {
        const uint32_t p = c * PC + ht + k * 2 * HELPERS;
        old[k] = (p <= n) ? drb[p] : 0;
    }
    bool changed = false;
#pragma unroll
    for (int k = 0; k < FP; k++) {
        const int i = ht + k * 2 * HELPERS;
        const uint32_t p = c * PC + i;
        if (p <= n) {
            const int32_t v = ring[(c & 1u) * PC + i];
            if (old[k] != v) { changed = true; drb[p] = v; }
        }
    }
    return changed;
}
__device__ __forceinline__ void helper_barrier() { asm volatile("bar.sync 1, %0;" ::"n"(2 * HELPERS) : "memory"); }

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

// grid = (K, B).  dr: d values by g, relative per segment; segBase / segChanged /
// segDirty: [B][KMAX].
__global__ void __launch_bounds__(PARSE_THREADS, SALZ_PARSE_CTAS) k_parse(const int4 *__restrict__ cand,
                                                            const uint32_t *__restrict__ blkN, int logS, int segLog,
                                                            int numSMs, int32_t *dr, const int32_t *__restrict__ segBase,
                                                            const uint32_t *__restrict__ segDirty,
                                                            uint32_t *__restrict__ segChanged,
                                                            int2 *__restrict__ choice)
{
    extern __shared__ __align__(16) uint8_t psmem[];
    int4 *stC = (int4 *)psmem;            // [2][PC] consumer records
    uint32_t *stD = (uint32_t *)(stC + 2 * PC);  // [2][PC] second ring address
    int32_t *ring = (int32_t *)(stD + 2 * PC);  // [RING + 1], d-space
    __shared__ int32_t sbase[KMAX];
    __shared__ uint32_t s_changed;
    const uint32_t ringAddr = (uint32_t)__cvta_generic_to_shared(ring);
    const uint32_t b = blockIdx.y, j = blockIdx.x;
    const uint32_t N = blkN[b];
    const uint32_t n = N - 8;
    const uint32_t L = j << segLog;
    if (L >= N) return;
    if (!segDirty[b * KMAX + j]) return;  // nothing to the right changed since this segment was computed
    const uint32_t R = min(N, L + (1u << segLog));
    const bool last = R == N;
    // co-resident CTAs put their consumer on different warp slots
    const uint32_t lin = blockIdx.x + blockIdx.y * gridDim.x;
    const int cw = (int)((lin + lin / (uint32_t)numSMs) % 4u);
    const int warp = ((int)(threadIdx.x >> 5) + PARSE_WARPS - cw) % PARSE_WARPS;  // role: 0 consumer, 1.. helpers
    const int lane = threadIdx.x & 31;
    const int tid = warp * 32 + lane;
    const size_t gb = (size_t)b << logS;
    const int4 *cd = cand + gb;
    int32_t *drb = dr + gb;
    int2 *ch = choice + gb;

    if (threadIdx.x < KMAX) {
        const uint32_t m = j + threadIdx.x;
        sbase[threadIdx.x] = (m < KMAX && ((size_t)m << segLog) < N) ? segBase[b * KMAX + m] - segBase[b * KMAX + j] : 0;
    }
    if (threadIdx.x == 0) s_changed = 0;
    for (int i = threadIdx.x; i <= RING; i += PARSE_THREADS) ring[i] = SALZ_INF;
    __syncthreads();
    SegCtx sg = {drb, sbase, R, j, segLog};

    const uint32_t cBeg = L / PC, cEnd = (R + PC - 1) / PC;  // segment lengths are multiples of PC
    if (warp >= 1 && warp <= HELPERS / 32) {
        const uint32_t c = cEnd - 1;
        stage_chunk(cd, sg, n, c, stC + (c & 1) * PC, stD + (c & 1) * PC, ringAddr, tid - 32);
    }
    __syncthreads();

    int32_t d = last ? SALZ_INF : 0;  // d[p + 1] relative to the segment's base
    bool changed = false;
    for (uint32_t c = cEnd; c-- > cBeg;) {
        if (warp >= 1) {
            // helpers: the chunk the consumer has just left goes to global memory first
            if (c + 1 < cEnd) {
                changed |= flush_chunk(ring, drb, n, c + 1, tid - 32);
                helper_barrier();
            }
            if (warp > HELPERS / 32) {
                if (c + 1 < cEnd) finish_chunk(cd, sg, N, n, c + 1, ch, tid - 32 - HELPERS);
            } else if (c > cBeg) {
                stage_chunk(cd, sg, n, c - 1, stC + ((c - 1) & 1) * PC, stD + ((c - 1) & 1) * PC, ringAddr, tid - 32);
            }
        } else {
            const int4 *rc = stC + (c & 1) * PC;
            const uint32_t *rd = stD + (c & 1) * PC;
            // ring slot of position c*PC + i is ((c & 1) * PC + i)
            const uint32_t slot0 = ringAddr + (c & 1u) * PC * 4;
            int4 r0 = rc[PC - 1], r1 = rc[PC - 2];
            int32_t a0 = lds_i32((uint32_t)r0.w), b0 = lds_i32(rd[PC - 1]);
            int32_t a1 = lds_i32((uint32_t)r1.w), b1 = lds_i32(rd[PC - 2]);
            // Records of positions i-2 .. i-9 are fetched one half-body ahead of their use;
            // the ring loads of position j-2 go out before position j is finished (its
            // targets are >= j+1: all stored).
            int4 X[8], Y[8];
            uint32_t XD[8], YD[8];
#pragma unroll
            for (int k = 0; k < 8; k++) { X[k] = rc[(PC - 3 - k) & (PC - 1)]; XD[k] = rd[(PC - 3 - k) & (PC - 1)]; }
            auto half = [&](int4 (&use)[8], uint32_t (&useD)[8], int4 (&load)[8], uint32_t (&loadD)[8], int i) {
#pragma unroll
                for (int k = 0; k < 8; k++) { load[k] = rc[(i - k - 10) & (PC - 1)]; loadD[k] = rd[(i - k - 10) & (PC - 1)]; }
#pragma unroll
                for (int k = 0; k < 8; k++) {
                    const int4 r2 = use[k];
                    const int32_t a2 = lds_i32((uint32_t)r2.w);
                    const int32_t b2 = lds_i32(useD[k]);
                    d = min(d, min(r0.x, min(a0 + r0.y, b0 + r0.z)));
                    sts_i32(slot0 + (uint32_t)(i - k) * 4, d);  // every lane: same address, same value
                    r0 = r1; a0 = a1; b0 = b1;
                    r1 = r2; a1 = a2; b1 = b2;
                }
            };
#pragma unroll 1
            for (int i = PC - 1; i >= 0; i -= 16) {
                half(X, XD, Y, YD, i);
                half(Y, YD, X, XD, i - 8);
            }
        }
        __syncthreads();
    }
    if (warp >= 1) {
        changed |= flush_chunk(ring, drb, n, cBeg, tid - 32);
        helper_barrier();
        if (warp > HELPERS / 32) finish_chunk(cd, sg, N, n, cBeg, ch, tid - 32 - HELPERS);
    }
    if (changed) s_changed = 1;
    __syncthreads();
    if (threadIdx.x == 0) segChanged[b * KMAX + j] = s_changed;
}

// Before the first pass: every base = 9 n_b (the all-literal estimate in d-space,
// and the exact base of the last segment), every segment dirty.
__global__ void k_seg_init(const uint32_t *__restrict__ blkN, int B, int32_t *__restrict__ segBase,
                           uint32_t *__restrict__ segDirty, uint32_t *__restrict__ segChanged)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= B * KMAX) return;
    segBase[i] = 9 * (int32_t)(blkN[i / KMAX] - 8);
    segDirty[i] = 1;
    segChanged[i] = 0;
}

// After a pass: bases from the d values at the segment starts; a segment is
// dirty when something to its right changed.  counters[8] += dirty segments.
__global__ void k_seg_scan(const uint32_t *__restrict__ blkN, int B, int logS, int segLog,
                           const int32_t *__restrict__ dr, int32_t *__restrict__ segBase,
                           uint32_t *__restrict__ segDirty, uint32_t *__restrict__ segChanged,
                           uint32_t *__restrict__ counters)
{
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    const uint32_t N = blkN[b];
    const int Kb = (int)((N + (1u << segLog) - 1) >> segLog);
    const int32_t *drb = dr + ((size_t)b << logS);
    int32_t base = 9 * (int32_t)(N - 8);
    uint32_t anyRight = 0, dirty = 0;
    for (int j = Kb - 1; j >= 0; j--) {
        segBase[b * KMAX + j] = base;
        segDirty[b * KMAX + j] = anyRight;
        dirty += anyRight;
        anyRight |= segChanged[b * KMAX + j];
        segChanged[b * KMAX + j] = 0;
        base += drb[(size_t)j << segLog];  // d at the first position of segment j = base of segment j-1
    }
    if (dirty) atomicAdd(&counters[8], dirty);
}

// jump[p] = first position >= the end of p's PARSE_CHUNK that the chosen parse
// reaches from p (pointer jumping inside the chunk; used by the token walk in
// emit.cu).  One CTA per chunk.
__global__ void __launch_bounds__(256) k_jump(const uint32_t *__restrict__ blkN, const int2 *__restrict__ choice,
                                              int logS, uint32_t *__restrict__ jump)
{
    __shared__ uint32_t nxt[PARSE_CHUNK];
    const uint32_t chunksPerBlock = 1u << (logS - PARSE_CHUNK_LOG);
    const uint32_t b = blockIdx.x / chunksPerBlock, c = blockIdx.x % chunksPerBlock;
    const uint32_t N = blkN[b];
    const uint32_t base = c << PARSE_CHUNK_LOG;
    if (base >= N) return;
    const uint32_t end = min(N, base + PARSE_CHUNK);
    const size_t gb = (size_t)b << logS;
    constexpr int PER = PARSE_CHUNK / 256;
    uint32_t v[PER];
#pragma unroll
    for (int i = 0; i < PER; i++) {
        uint32_t p = base + threadIdx.x + i * 256;
        v[i] = (p < end) ? p + (uint32_t)choice[gb + p].y : end;
        nxt[threadIdx.x + i * 256] = v[i];
    }
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
        uint32_t p = base + threadIdx.x + i * 256;
        if (p < end) jump[gb + p] = v[i];
    }
}

}  // namespace

int parse_build(Workspace &ws, const BatchDesc &bd, cudaStream_t st, LaunchCounter &lc)
{
    static bool attrSet = false;
    if (!attrSet) {
        cudaFuncSetAttribute(k_parse, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)PARSE_SMEM);
        attrSet = true;
    }
    // Segments per block: enough chains to fill the CTA slots (4 per SM), but a cut
    // only pays when it is at least 4-fold (two passes + the confirming one).
    int segLog = bd.logS, K = 1;
    {
        const int slots = ws.numSMs * SALZ_PARSE_CTAS;
        int want = ws.parseSegments ? ws.parseSegments : slots / std::max(1, bd.B);
        if (!ws.parseSegments && want < 4) want = 1;
        while (2 * K <= want && K < KMAX && (bd.S >> 1) >> (bd.logS - segLog) >= 16u * PC) { K *= 2; segLog--; }
    }
    // cost holds d relative to the segment bases: cost[p] = dr[p] + segBase[p >> segLog] - 9p
    // (salz_cuda_debug_stages converts; nothing downstream needs the absolute costs)
    int32_t *dr = ws.cost;
    ws.parseSegLog = segLog;
    k_seg_init<<<(bd.B * KMAX + 255) / 256, 256, 0, st>>>(ws.blkN, bd.B, ws.segBase, ws.segDirty, ws.segChanged);
    lc.launches += 1;
    if (K > 1) {
        if (cudaMemsetAsync(dr, 0, (size_t)bd.B << bd.logS << 2, st) != cudaSuccess) return -1;
    }
    for (int pass = 0;; pass++) {
        k_parse<<<dim3(K, bd.B), PARSE_THREADS, PARSE_SMEM, st>>>(ws.cand, ws.blkN, bd.logS, segLog, ws.numSMs, dr, ws.segBase,
                                                                  ws.segDirty, ws.segChanged, ws.choice);
        lc.launches += 1;
        if (K == 1) break;
        if (cudaMemsetAsync(ws.counters + 8, 0, 4, st) != cudaSuccess) return -1;
        k_seg_scan<<<(bd.B + 63) / 64, 64, 0, st>>>(ws.blkN, bd.B, bd.logS, segLog, dr, ws.segBase, ws.segDirty,
                                                    ws.segChanged, ws.counters);
        lc.launches += 1;
        if (cudaMemcpyAsync(ws.h_counters + 8, ws.counters + 8, 4, cudaMemcpyDeviceToHost, st) != cudaSuccess) return -1;
        if (cudaStreamSynchronize(st) != cudaSuccess) return -1;
        if (ws.h_counters[8] == 0) break;
        if (pass > KMAX + 2) {
            fprintf(stderr, "salz_cuda: segmented parse failed to settle\n");
            return -1;
        }
    }
    k_jump<<<bd.B << (bd.logS - PARSE_CHUNK_LOG), 256, 0, st>>>(ws.blkN, ws.choice, bd.logS, ws.jump);
    lc.launches += 1;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        fprintf(stderr, "salz_cuda: parse_build: %s\n", cudaGetErrorString(e));
        return -1;
    }
    return 0;
}
