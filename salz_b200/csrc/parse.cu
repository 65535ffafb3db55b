// parse.cu - the cost-optimal LZ parse.
//
//   k_parse: right-to-left shortest-path DP with the reference's cost model and
//            strict '<' tie-break order literal, PSV, NSV (ref lib/salz.c:562-662).
//   k_jump : chunk-exit pointers of the chosen parse for the token walk in emit.cu.
#include "salz_dev.cuh"

namespace {

// ---------------------------------------------------------------------------
// Optimal parse.  One CTA per block.  The cost chain
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
//   stage(c-1)  consumer records + candidate records of chunk c-1
//   finish(c+1) the choice per position, recomputed from the final costs with
//               the reference's comparison order (strict '<': literal, PSV,
//               NSV; ref lib/salz.c:622-661), and written out
//   flush(c)    ring (d-space) -> global cost array, by the helper warps
// ---------------------------------------------------------------------------
constexpr int PC = 2048;           // positions per staged chunk (long enough to hide the helpers' memory latency)
constexpr int HELPERS = 128;       // threads per helper group (stagers, finishers)
constexpr int PER = 8;             // positions a helper thread has in flight
constexpr int SUBS = PC / (HELPERS * PER);  // batches of PER positions per helper thread and chunk
constexpr int PARSE_WARPS = 1 + 2 * HELPERS / 32;
constexpr int PARSE_THREADS = 32 * PARSE_WARPS;  // consumer warp + stagers + finishers
constexpr int NOT_FAR = -1;
constexpr int RING = 2 * PC;       // ring entries; entry RING is a constant INF (absent edge)
constexpr size_t PARSE_SMEM = 2 * PC * (sizeof(int4) + sizeof(uint32_t)) + (RING + 4) * sizeof(int32_t);

// Consumer record: x = F in d-space, y / z = weight - 9 len of the two ring
// edges, w and sd[] = shared-memory addresses of their ring entries.
__device__ __forceinline__ void stage_chunk(const int4 *__restrict__ cd, const int32_t *cs, uint32_t n,
                                            uint32_t chunk, int4 *sc, uint32_t *sd, uint32_t ringAddr, int tid)
{
    const uint32_t farFrom = (chunk + 2) * PC;  // targets >= this are final and visible
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
    for (int k = 0; k < PER; k++) {  // all far-cost loads in flight together
        const uint32_t p = base + tid + k * HELPERS;
        const uint32_t ta = p + (uint32_t)c[k].y, tb = p + (uint32_t)c[k].w;
        t1[k] = (c[k].y >= 3 && ta >= farFrom) ? __ldcg(&cs[ta]) : NOT_FAR;
        t2[k] = (c[k].w >= 3 && tb >= farFrom) ? __ldcg(&cs[tb]) : NOT_FAR;
    }
#pragma unroll
    for (int k = 0; k < PER; k++) {
        const uint32_t p = base + tid + k * HELPERS;
        int32_t F = (p == n) ? (int32_t)(9u * n) : SALZ_INF;  // cost[n] = 0 (ref lib/salz.c:621)
        int32_t w1 = 0, w2 = 0;
        uint32_t o1 = RING * 4, o2 = RING * 4;
        if (c[k].y >= 3) {
            const int w = factor_cost_bits((uint32_t)c[k].x, (uint32_t)c[k].y);
            if (t1[k] != NOT_FAR) F = min(F, w + t1[k] + (int32_t)(9u * p));
            else { w1 = w - 9 * c[k].y; o1 = ((p + (uint32_t)c[k].y) & (RING - 1)) * 4; }
        }
        if (c[k].w >= 3) {
            const int w = factor_cost_bits((uint32_t)c[k].z, (uint32_t)c[k].w);
            if (t2[k] != NOT_FAR) F = min(F, w + t2[k] + (int32_t)(9u * p));
            else { w2 = w - 9 * c[k].w; o2 = ((p + (uint32_t)c[k].w) & (RING - 1)) * 4; }
        }
        sc[tid + k * HELPERS] = make_int4(F, w1, w2, (int)(ringAddr + o1));
        sd[tid + k * HELPERS] = ringAddr + o2;
    }
  }
}

// Choices of a chunk whose costs (and every cost to its right) are final and in
// global memory: the reference's comparisons, in its order (ref lib/salz.c:622-661).
__device__ __forceinline__ void finish_chunk(const int4 *__restrict__ cd, const int32_t *cs, uint32_t N, uint32_t n,
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
    for (int k = 0; k < PER; k++) {  // all cost loads in flight together
        const uint32_t p = base + tid + k * HELPERS;
        const bool in = p >= 1 && p < n;
        r0[k] = in ? __ldcg(&cs[p + 1]) : 0;
        r1[k] = (in && c[k].y >= 3) ? __ldcg(&cs[p + (uint32_t)c[k].y]) : 0;
        r2[k] = (in && c[k].w >= 3) ? __ldcg(&cs[p + (uint32_t)c[k].w]) : 0;
    }
#pragma unroll
    for (int k = 0; k < PER; k++) {
        const uint32_t p = base + tid + k * HELPERS;
        if (p >= N) break;
        uint32_t bo = 0, bl = 1;
        if (p >= 1 && p < n) {
            int32_t best = 9 + r0[k];
            if (c[k].y >= 3) {
                const int32_t c1 = factor_cost_bits((uint32_t)c[k].x, (uint32_t)c[k].y) + r1[k];
                if (c1 < best) { best = c1; bo = (uint32_t)c[k].x; bl = (uint32_t)c[k].y; }
            }
            if (c[k].w >= 3) {
                const int32_t c2 = factor_cost_bits((uint32_t)c[k].z, (uint32_t)c[k].w) + r2[k];
                if (c2 < best) { best = c2; bo = (uint32_t)c[k].z; bl = (uint32_t)c[k].w; }
            }
        }
        ch[p] = make_int2((int)bo, (int)bl);
    }
  }
}

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

__global__ void __launch_bounds__(PARSE_THREADS, 2) k_parse(const int4 *__restrict__ cand,
                                                         const uint32_t *__restrict__ blkN, int logS, int numSMs,
                                                         int32_t *cost, int2 *__restrict__ choice)
{
    extern __shared__ __align__(16) uint8_t psmem[];
    int4 *stC = (int4 *)psmem;            // [2][PC] consumer records
    uint32_t *stD = (uint32_t *)(stC + 2 * PC);  // [2][PC] second ring address
    int32_t *ring = (int32_t *)(stD + 2 * PC);  // [RING + 1], d-space
    const uint32_t ringAddr = (uint32_t)__cvta_generic_to_shared(ring);
    const uint32_t b = blockIdx.x;
    // co-resident CTAs (b, b + numSMs, ...) put their consumer on different warp slots
    const int cw = (int)((b / (uint32_t)numSMs) % 4u);
    const int warp = ((int)(threadIdx.x >> 5) + PARSE_WARPS - cw) % PARSE_WARPS;  // role: 0 consumer, 1.. helpers
    const int lane = threadIdx.x & 31;
    const int tid = warp * 32 + lane;
    const uint32_t N = blkN[b];
    const uint32_t n = N - 8;
    const size_t gb = (size_t)b << logS;
    const int4 *cd = cand + gb;
    int32_t *cs = cost + gb + b;  // n_b + 1 entries per block (stride S + 1)
    int2 *ch = choice + gb;

    const uint32_t nChunks = (N + PC - 1) / PC;
    if (warp >= 1 && warp <= HELPERS / 32) {
        const uint32_t c = nChunks - 1;
        stage_chunk(cd, cs, n, c, stC + (c & 1) * PC, stD + (c & 1) * PC, ringAddr, tid - 32);
    }
    for (int i = threadIdx.x; i <= RING; i += PARSE_THREADS) ring[i] = SALZ_INF;
    __syncthreads();

    int32_t d = SALZ_INF;  // d[p + 1]
    for (uint32_t c = nChunks; c-- > 0;) {
        if (warp > HELPERS / 32) {
            if (c + 1 < nChunks)
                finish_chunk(cd, cs, N, n, c + 1, ch, tid - 32 - HELPERS);
        } else if (warp >= 1) {
            if (c > 0)
                stage_chunk(cd, cs, n, c - 1, stC + ((c - 1) & 1) * PC, stD + ((c - 1) & 1) * PC, ringAddr, tid - 32);
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
        // flush chunk c: d-space ring -> global costs (needed by stage(c-2) and finish(c))
        if (warp >= 1) {
            for (int i = tid - 32; i < PC; i += 2 * HELPERS) {
                const uint32_t p = c * PC + i;
                if (p <= n) cs[p] = ring[(c & 1u) * PC + i] - (int32_t)(9u * p);
            }
        }
        __syncthreads();
    }
    if (warp > HELPERS / 32) finish_chunk(cd, cs, N, n, 0, ch, tid - 32 - HELPERS);
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
    k_parse<<<bd.B, PARSE_THREADS, PARSE_SMEM, st>>>(ws.cand, ws.blkN, bd.logS, ws.numSMs, ws.cost, ws.choice);
    k_jump<<<bd.B << (bd.logS - PARSE_CHUNK_LOG), 256, 0, st>>>(ws.blkN, ws.choice, bd.logS, ws.jump);
    lc.launches += 2;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        fprintf(stderr, "salz_cuda: parse_build: %s\n", cudaGetErrorString(e));
        return -1;
    }
    return 0;
}
