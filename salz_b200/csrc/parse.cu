// parse.cu - the cost-optimal LZ parse.
//
//   k_parse: right-to-left shortest-path DP with the reference's cost model and
//            strict '<' tie-break order literal, PSV, NSV (ref lib/salz.c:562-662).
//   k_jump : chunk-exit pointers of the chosen parse for the token walk in emit.cu.
#include "salz_dev.cuh"

namespace {

// ---------------------------------------------------------------------------
// Optimal parse.  One CTA per block.  Warp 0 is the consumer: it walks the
// serial cost chain in 32-position windows right to left and does nothing
// else.  Warps 1..3 work one chunk (PC positions) ahead and one behind it:
//   stage(c-1)  candidates, edge weights and the final costs of targets that
//               lie at least one whole chunk to the right of chunk c-1, into
//               shared memory, so the chain never waits on global memory;
//   finish(c+1) the choice per position, recomputed from the now final costs
//               with the reference's comparison order (strict '<': literal,
//               PSV, NSV; ref lib/salz.c:622-661), and written out.
//
// Inside a window lane l owns position q - l (q = right end).  Targets
//   - staged as "far": weight + final cost, ready,
//   - beyond the window but inside the last two chunks: cost ring in smem,
//   - inside the window: relaxed to a fixpoint with shuffles.
// The literal chain inside the window is a prefix-min scan in (min,+).
// ---------------------------------------------------------------------------
constexpr int PC = 512;            // positions per staged chunk
constexpr int HELPERS = 128;       // threads per helper group (stagers, finishers)
constexpr int PER = PC / HELPERS;  // positions per helper thread and chunk
constexpr int PARSE_THREADS = 32 + 2 * HELPERS;  // consumer warp + stagers + finishers
constexpr int NOT_FAR = -1;
constexpr size_t PARSE_SMEM = 3 * PC * 2 * sizeof(int4) + 2 * PC * sizeof(int32_t);

// record: x = len (>= 3), y = weight, z = weight + final target cost or NOT_FAR, w = offs.
// An absent candidate is staged as far with cost INF.  `tid` in [0, HELPERS).
__device__ __forceinline__ void stage_chunk(const int4 *__restrict__ cd, const int32_t *cs, uint32_t n,
                                            uint32_t chunk, int4 *sa, int4 *sb, int tid)
{
    const uint32_t base = chunk * PC;
    const uint32_t farFrom = (chunk + 2) * PC;  // targets >= this are final and visible
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
        int4 a = make_int4(0, 0, SALZ_INF, 0), b = make_int4(0, 0, SALZ_INF, 0);
        if (c[k].y >= 3) {
            const int w = factor_cost_bits((uint32_t)c[k].x, (uint32_t)c[k].y);
            a = make_int4(c[k].y, w, (t1[k] != NOT_FAR) ? w + t1[k] : NOT_FAR, c[k].x);
        }
        if (c[k].w >= 3) {
            const int w = factor_cost_bits((uint32_t)c[k].z, (uint32_t)c[k].w);
            b = make_int4(c[k].w, w, (t2[k] != NOT_FAR) ? w + t2[k] : NOT_FAR, c[k].z);
        }
        sa[tid + k * HELPERS] = a;
        sb[tid + k * HELPERS] = b;
    }
}

// Choices of a chunk whose costs (and every cost to its right) are final.
__device__ __forceinline__ void finish_chunk(const int4 *sa, const int4 *sb, const int32_t *cs, uint32_t N, uint32_t n,
                                             uint32_t chunk, int2 *__restrict__ ch, int tid)
{
    const uint32_t base = chunk * PC;
    int4 a[PER], b[PER];
    int32_t r0[PER], r1[PER], r2[PER];
#pragma unroll
    for (int k = 0; k < PER; k++) {  // all cost loads in flight together
        const uint32_t p = base + tid + k * HELPERS;
        a[k] = sa[tid + k * HELPERS];
        b[k] = sb[tid + k * HELPERS];
        const bool in = p >= 1 && p < n;  // position 0 and the 8-byte tail are forced literals
        r0[k] = in ? __ldcg(&cs[p + 1]) : 0;
        r1[k] = (in && a[k].z == NOT_FAR) ? __ldcg(&cs[p + (uint32_t)a[k].x]) : 0;
        r2[k] = (in && b[k].z == NOT_FAR) ? __ldcg(&cs[p + (uint32_t)b[k].x]) : 0;
    }
#pragma unroll
    for (int k = 0; k < PER; k++) {
        const uint32_t p = base + tid + k * HELPERS;
        if (p >= N) break;
        uint32_t bo = 0, bl = 1;
        if (p >= 1 && p < n) {
            int32_t best = 9 + r0[k];
            if (a[k].z != SALZ_INF) {
                const int32_t c1 = (a[k].z != NOT_FAR) ? a[k].z : a[k].y + r1[k];
                if (c1 < best) { best = c1; bo = (uint32_t)a[k].w; bl = (uint32_t)a[k].x; }
            }
            if (b[k].z != SALZ_INF) {
                const int32_t c2 = (b[k].z != NOT_FAR) ? b[k].z : b[k].y + r2[k];
                if (c2 < best) { best = c2; bo = (uint32_t)b[k].w; bl = (uint32_t)b[k].x; }
            }
        }
        ch[p] = make_int2((int)bo, (int)bl);
    }
}

__global__ void __launch_bounds__(PARSE_THREADS) k_parse(const int4 *__restrict__ cand,
                                                         const uint32_t *__restrict__ blkN, int logS, int32_t *cost,
                                                         int2 *__restrict__ choice)
{
    extern __shared__ __align__(16) uint8_t psmem[];
    int4 *stA = (int4 *)psmem;            // [3][PC]
    int4 *stB = stA + 3 * PC;             // [3][PC]
    int32_t *ring = (int32_t *)(stB + 3 * PC);  // [2 * PC]
    const uint32_t b = blockIdx.x;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t N = blkN[b];
    const uint32_t n = N - 8;
    const size_t gb = (size_t)b << logS;
    const int4 *cd = cand + gb;
    int32_t *cs = cost + gb + b;  // n_b + 1 entries per block (stride S + 1)
    int2 *ch = choice + gb;

    const uint32_t nChunks = (N + PC - 1) / PC;
    if (warp >= 1 && warp <= HELPERS / 32) {
        const uint32_t c = nChunks - 1;
        stage_chunk(cd, cs, n, c, stA + (c % 3) * PC, stB + (c % 3) * PC, tid - 32);
    }
    __syncthreads();

    int32_t costRight = SALZ_INF;  // cost[q+1] of the window being processed
    for (uint32_t c = nChunks; c-- > 0;) {
        if (warp > HELPERS / 32) {
            if (c + 1 < nChunks)
                finish_chunk(stA + ((c + 1) % 3) * PC, stB + ((c + 1) % 3) * PC, cs, N, n, c + 1, ch,
                             tid - 32 - HELPERS);
        } else if (warp >= 1) {
            if (c > 0) stage_chunk(cd, cs, n, c - 1, stA + ((c - 1) % 3) * PC, stB + ((c - 1) % 3) * PC, tid - 32);
        } else {
            const int4 *sa = stA + (c % 3) * PC, *sb = stB + (c % 3) * PC;
            const uint32_t lastP = min(N - 1, c * PC + PC - 1);
            const uint32_t firstW = (c * PC) >> 5;
            int4 na = sa[((lastP | 31u) - lane) & (PC - 1)], nb = sb[((lastP | 31u) - lane) & (PC - 1)];
            for (uint32_t w = (lastP >> 5) + 1; w-- > firstW;) {
                const uint32_t q = (w << 5) + 31;
                const uint32_t p = q - lane;
                const int4 ra = na, rb = nb;
                if (w > firstW) {  // next window's records, off the critical path
                    na = sa[(p - 32) & (PC - 1)];
                    nb = sb[(p - 32) & (PC - 1)];
                }
                const int32_t rv1 = ring[(p + (uint32_t)ra.x) & (2 * PC - 1)];
                const int32_t rv2 = ring[(p + (uint32_t)rb.x) & (2 * PC - 1)];
                const bool near1 = ra.z == NOT_FAR && ra.x <= lane;
                const bool near2 = rb.z == NOT_FAR && rb.x <= lane;
                const int32_t f1 = (ra.z != NOT_FAR) ? ra.z : (near1 ? SALZ_INF : ra.y + rv1);
                const int32_t f2 = (rb.z != NOT_FAR) ? rb.z : (near2 ? SALZ_INF : rb.y + rv2);
                const int32_t x = (p == n) ? 0 : min(f1, f2);  // cost[n] = 0 (ref lib/salz.c:621)

                // --- literal chain: cost[l] = min( min_{k<=l} x[k] + 9(l-k), costRight + 9(l+1) ) ---
                int32_t y = x - 9 * lane;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    int32_t t = __shfl_up_sync(0xffffffffu, y, o);
                    if (lane >= o) y = min(y, t);
                }
                int32_t cst = min(y + 9 * lane, costRight + 9 * (lane + 1));
                if (p > n) cst = SALZ_INF;
                // --- in-window edges: relax every edge (factor edges, then one literal
                // step) until nothing changes; any relaxation order reaches the same
                // exact fixpoint, and runs of literals after a factor are short ---
                if (__any_sync(0xffffffffu, near1 || near2)) {
                    const int s1 = near1 ? (lane - ra.x) : lane, s2 = near2 ? (lane - rb.x) : lane;
                    const int32_t w1 = near1 ? ra.y : 0, w2 = near2 ? rb.y : 0;
                    for (;;) {
                        int32_t v = min(cst, min(w1 + __shfl_sync(0xffffffffu, cst, s1),
                                                 w2 + __shfl_sync(0xffffffffu, cst, s2)));
                        int32_t u = __shfl_up_sync(0xffffffffu, v, 1);
                        if (lane >= 1 && p <= n) v = min(v, u + 9);
                        const bool changed = v < cst;
                        cst = v;
                        if (!__any_sync(0xffffffffu, changed)) break;
                    }
                }
                ring[p & (2 * PC - 1)] = cst;
                if (p <= n) cs[p] = cst;
                costRight = __shfl_sync(0xffffffffu, cst, 31);
                __syncwarp();
            }
        }
        __syncthreads();
    }
    if (warp > HELPERS / 32) finish_chunk(stA, stB, cs, N, n, 0, ch, tid - 32 - HELPERS);
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
    k_parse<<<bd.B, PARSE_THREADS, PARSE_SMEM, st>>>(ws.cand, ws.blkN, bd.logS, ws.cost, ws.choice);
    k_jump<<<bd.B << (bd.logS - PARSE_CHUNK_LOG), 256, 0, st>>>(ws.blkN, ws.choice, bd.logS, ws.jump);
    lc.launches += 2;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        fprintf(stderr, "salz_cuda: parse_build: %s\n", cudaGetErrorString(e));
        return -1;
    }
    return 0;
}
