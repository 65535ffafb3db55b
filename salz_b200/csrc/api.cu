// api.cu - context, workspace and the extern "C" entry points of
// include/salz_cuda.h.  Host-side orchestration only; the kernels live in
// sa.cu / lcp.cu / parse.cu / emit.cu.
#include "../../include/salz_cuda.h"
#include "salz_dev.cuh"

#include <algorithm>
#include <cstdarg>
#include <cstdlib>
#include <cstring>
#include <vector>

struct salz_cuda_ctx {
    int device;
    cudaStream_t stream;
    cudaStream_t ownStream;
    Workspace ws;
    size_t maxBlockLen;
    int sortMode;
    salz_cuda_stats stats;
    char err[256];
    std::vector<void *> allocs;
    // pinned host mirrors
    uint32_t *h_blkN;
    uint32_t *h_outLen;
    uint32_t *h_salzLen;
    uint64_t *h_outOff;
    cudaEvent_t ev[8];
    // host-stream pipeline: second copy stream, two text and two output buffers
    cudaStream_t copyStream;
    uint8_t *Tbuf[2];
    uint8_t *outbuf[2];
    cudaEvent_t tReady[2], oDone[2], computed;
};

static void set_err(salz_cuda_ctx *c, const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(c->err, sizeof(c->err), fmt, ap);
    va_end(ap);
    fprintf(stderr, "salz_cuda: %s\n", c->err);
}

#define CKC(ctx, call)                                                                \
    do {                                                                              \
        cudaError_t e_ = (call);                                                      \
        if (e_ != cudaSuccess) {                                                      \
            set_err(ctx, "%s: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__,   \
                    __LINE__);                                                        \
            return -1;                                                                \
        }                                                                             \
    } while (0)

template <typename T>
static bool dalloc(salz_cuda_ctx *c, T **p, size_t count)
{
    void *q = nullptr;
    if (cudaMalloc(&q, std::max<size_t>(count, 1) * sizeof(T)) != cudaSuccess) {
        set_err(c, "cudaMalloc of %zu bytes failed", count * sizeof(T));
        return false;
    }
    c->allocs.push_back(q);
    *p = (T *)q;
    return true;
}

extern "C" int salz_cuda_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
    return n;
}

extern "C" salz_cuda_ctx *salz_cuda_ctx_create(int device, size_t max_block_len, size_t max_batch_blocks)
{
    if (max_block_len < 9 || max_batch_blocks < 1) return nullptr;
    int ndev = salz_cuda_device_count();
    if (device < 0 || device >= ndev) {
        fprintf(stderr, "salz_cuda: no CUDA device %d (found %d); there is no CPU fallback\n", device, ndev);
        return nullptr;
    }
    if (max_batch_blocks > 65535) {  // the block index is gridDim.y of three kernels
        fprintf(stderr, "salz_cuda: max_batch_blocks %zu exceeds 65535\n", max_batch_blocks);
        return nullptr;
    }
    if (cudaSetDevice(device) != cudaSuccess) return nullptr;
    if (sa_init_device() || parse_init_device()) return nullptr;
    salz_cuda_ctx *c = new salz_cuda_ctx();
    c->device = device;
    c->maxBlockLen = max_block_len;
    c->sortMode = 1;
    c->err[0] = 0;
    memset(&c->stats, 0, sizeof(c->stats));
    const char *sm = getenv("SALZ_SORT_MODE");
    if (sm) c->sortMode = atoi(sm) ? 1 : 0;
    const char *ps = getenv("SALZ_PARSE_SEGMENTS");

    Workspace &ws = c->ws;
    memset(&ws, 0, sizeof(ws));
    ws.parseSegments = ps ? atoi(ps) : 0;
    const char *scm = getenv("SALZ_SCATTER"), *sgr = getenv("SALZ_SORT_GROUP");
    ws.scatterMode = scm ? atoi(scm) : 6;   // measured, ms of sort per step: 0 10.89, 1 9.33, 2 10.58, 6 8.64
    const char *bro = getenv("SALZ_BAG_REORDER");
    ws.bagReorder = bro ? atoi(bro) : 3;
    const char *r0b = getenv("SALZ_R0_BITS");
    ws.round0Bits = r0b ? atoi(r0b) : 0;
    const char *lcr = getenv("SALZ_LOCAL_COUNT_ROUND");
    ws.localCountRound = lcr ? atoi(lcr) : 3;
    ws.sortGroup = sgr ? atoi(sgr) : 0;     // measured: one launch per pass 10.12 ms, groups of 8 blocks in one launch 10.80
    cudaDeviceGetAttribute(&ws.numSMs, cudaDevAttrMultiProcessorCount, device);
    if (ws.numSMs <= 0) ws.numSMs = 148;
    int logS = 12;
    while (((size_t)1 << logS) < max_block_len) logS++;
    ws.logS = logS;
    ws.S = 1u << logS;
    ws.Bmax = max_batch_blocks;
    ws.Mcap = ws.Bmax * (size_t)ws.S;
    // 32-bit slot indices everywhere; ~90 B of workspace per position: 2^30 + 2^27 positions
    // (the 1 GiB + 4096 B inputs of BASELINE.json in ONE batch) take ~105 GB of the 180 GB
    if (ws.Mcap > ((size_t)1 << 30) + ((size_t)1 << 27)) {
        fprintf(stderr, "salz_cuda: batch of %zu x %u exceeds 2^30 + 2^27 positions\n", ws.Bmax, ws.S);
        delete c;
        return nullptr;
    }
    const size_t M = ws.Mcap, B = ws.Bmax;
    bool ok = cudaStreamCreateWithFlags(&c->ownStream, cudaStreamNonBlocking) == cudaSuccess;
    c->stream = c->ownStream;
    for (int i = 0; i < 8 && ok; i++) ok = cudaEventCreate(&c->ev[i]) == cudaSuccess;
    ws.Tcap = M + 256;
    ok = ok && dalloc(c, &ws.T, ws.Tcap);
    c->Tbuf[0] = ws.T;
    ok = ok && dalloc(c, &c->Tbuf[1], ws.Tcap);
    ok = ok && dalloc(c, &ws.blkN, B);
    const bool packed = 3 * logS <= 64;
    for (int i = 0; i < 2 && ok; i++)
        ok = dalloc(c, &ws.keys[i], M + M / 4096 + 64) && dalloc(c, &ws.vals[i], packed ? 1 : M);  // keys[0] doubles as the ANSV far list
    for (int i = 0; i < 3 && ok; i++)
        ok = dalloc(c, &ws.listCnt[i], B) && dalloc(c, &ws.listBase[i], B + 1) && dalloc(c, &ws.tileBase[i], B + 1);
    ws.tileCap = M / sa_tile_elems() + B + 2;
    for (int i = 0; i < 3 && ok; i++) ok = dalloc(c, &ws.tileDesc[i], ws.tileCap);
    ok = ok && dalloc(c, &ws.rank, M + 64) && dalloc(c, &ws.SA, M + 64);
    ok = ok && dalloc(c, &ws.cntSnap, B) && dalloc(c, &ws.alphaMask, B * 8) && dalloc(c, &ws.codeTab, B * 256) &&
         dalloc(c, &ws.symBits, B) && dalloc(c, &ws.h0, B);
    ok = ok && dalloc(c, &ws.hist, B * 8 * 256) && dalloc(c, &ws.digitBase, B * 8 * 256);
    ok = ok && dalloc(c, &ws.tileState, (ws.tileCap + 1) * 256);
    ok = ok && dalloc(c, &ws.sched, ws.tileCap * 8) && dalloc(c, &ws.passDone, B * 8);
    ok = ok && dalloc(c, &ws.counters, 16);
    ok = ok && dalloc(c, &ws.plcp, M) && dalloc(c, &ws.lcp, M) && dalloc(c, &ws.tileMax, M / 1024 + 1);
    for (int l = 1; l < 6 && ok; l++) {
        size_t cnt = (5 * l <= logS) ? (M >> (5 * l)) : 0;
        ok = dalloc(c, &ws.pyrSA[l], cnt + 32) && dalloc(c, &ws.pyrLCP[l], cnt + 32);
    }
    ok = ok && dalloc(c, &ws.tileMinV, M / 4096 + 1) && dalloc(c, &ws.tileMinL, M / 4096 + 1) &&
         dalloc(c, &ws.tilePreV, M / 4096 + 1) && dalloc(c, &ws.tileSufV, M / 4096 + 1);
    ok = ok && dalloc(c, &ws.cand, M + 64) && dalloc(c, &ws.cost, M + B + 64) && dalloc(c, &ws.choice, M);
    ws.keys[2] = (uint64_t *)ws.cand;  // the candidates are written after the suffix array is complete
    ok = ok && dalloc(c, &ws.jump, M + 64);
    ok = ok && dalloc(c, &ws.segBase, B * 64) && dalloc(c, &ws.cost2, M + 64);
    const size_t chunks = M >> PARSE_CHUNK_LOG;
    ok = ok && dalloc(c, &ws.chunkEntry, chunks) && dalloc(c, &ws.chunkTok, chunks) &&
         dalloc(c, &ws.chunkBits, chunks);
    ok = ok && dalloc(c, &ws.tokPos, M) && dalloc(c, &ws.tokBit, M) && dalloc(c, &ws.firstTok, M / 4 + 1);
    ok = ok && dalloc(c, &ws.blkTokens, B) && dalloc(c, &ws.blkBits, B) && dalloc(c, &ws.blkSalzLen, B) &&
         dalloc(c, &ws.blkOutLen, B) && dalloc(c, &ws.blkOutOff, B + 1);
    ws.outCap = B * ((size_t)ws.S + ws.S / 8 + 16) + 64;
    ok = ok && dalloc(c, &ws.out, ws.outCap);
    c->outbuf[0] = ws.out;
    ok = ok && dalloc(c, &c->outbuf[1], ws.outCap);
    ok = ok && cudaStreamCreateWithFlags(&c->copyStream, cudaStreamNonBlocking) == cudaSuccess;
    for (int i = 0; i < 2 && ok; i++)
        ok = cudaEventCreateWithFlags(&c->tReady[i], cudaEventDisableTiming) == cudaSuccess &&
             cudaEventCreateWithFlags(&c->oDone[i], cudaEventDisableTiming) == cudaSuccess;
    ok = ok && cudaEventCreateWithFlags(&c->computed, cudaEventDisableTiming) == cudaSuccess;
    ok = ok && cudaMallocHost((void **)&ws.h_counters, 64) == cudaSuccess;
    ok = ok && cudaMallocHost((void **)&c->h_blkN, B * 4) == cudaSuccess;
    ws.h_blkN = c->h_blkN;
    ok = ok && cudaMallocHost((void **)&c->h_outLen, B * 4) == cudaSuccess;
    ok = ok && cudaMallocHost((void **)&c->h_salzLen, B * 4) == cudaSuccess;
    ok = ok && cudaMallocHost((void **)&c->h_outOff, (B + 1) * 8) == cudaSuccess;
    if (!ok) {
        fprintf(stderr, "salz_cuda: context creation failed: %s\n", cudaGetErrorString(cudaGetLastError()));
        salz_cuda_ctx_destroy(c);
        return nullptr;
    }
    return c;
}

extern "C" void salz_cuda_ctx_destroy(salz_cuda_ctx *c)
{
    if (!c) return;
    cudaSetDevice(c->device);
    cudaDeviceSynchronize();
    for (void *p : c->allocs) cudaFree(p);
    if (c->ws.h_counters) cudaFreeHost(c->ws.h_counters);
    if (c->h_blkN) cudaFreeHost(c->h_blkN);
    if (c->h_outLen) cudaFreeHost(c->h_outLen);
    if (c->h_salzLen) cudaFreeHost(c->h_salzLen);
    if (c->h_outOff) cudaFreeHost(c->h_outOff);
    for (int i = 0; i < 8; i++)
        if (c->ev[i]) cudaEventDestroy(c->ev[i]);
    if (c->ownStream) cudaStreamDestroy(c->ownStream);
    if (c->copyStream) cudaStreamDestroy(c->copyStream);
    for (int i = 0; i < 2; i++) {
        if (c->tReady[i]) cudaEventDestroy(c->tReady[i]);
        if (c->oDone[i]) cudaEventDestroy(c->oDone[i]);
    }
    if (c->computed) cudaEventDestroy(c->computed);
    delete c;
}

extern "C" void *salz_cuda_host_alloc(size_t bytes)
{
    void *p = nullptr;
    if (cudaMallocHost(&p, bytes) != cudaSuccess) return nullptr;
    return p;
}
extern "C" void salz_cuda_host_free(void *p) { if (p) cudaFreeHost(p); }

extern "C" const char *salz_cuda_last_error(const salz_cuda_ctx *c) { return c ? c->err : "no context"; }
extern "C" void salz_cuda_set_sort_mode(salz_cuda_ctx *c, int mode) { if (c) c->sortMode = mode ? 1 : 0; }
extern "C" void salz_cuda_set_parse_segments(salz_cuda_ctx *c, int k) { if (c) c->ws.parseSegments = k; }
extern "C" void salz_cuda_set_stream(salz_cuda_ctx *c, void *stream)
{
    if (!c) return;
    cudaStreamSynchronize(c->stream);
    c->stream = stream ? (cudaStream_t)stream : c->ownStream;
}
extern "C" int salz_cuda_get_stats(const salz_cuda_ctx *c, salz_cuda_stats *out)
{
    if (!c || !out) return -1;
    *out = c->stats;
    return 0;
}
extern "C" void salz_cuda_reset_stats(salz_cuda_ctx *c) { if (c) memset(&c->stats, 0, sizeof(c->stats)); }

// Runs the device pipeline on the B blocks whose text already sits in ws.T at
// `Tstride` and whose lengths are in c->h_blkN.  upto: 1 = SA only, 5 = all.
static int run_batch(salz_cuda_ctx *c, int B, size_t Tstride, int framing, int upto)
{
    Workspace &ws = c->ws;
    cudaStream_t st = c->stream;
    BatchDesc bd;
    bd.B = B;
    bd.logS = ws.logS;
    bd.S = ws.S;
    bd.Tstride = Tstride;
    uint32_t M = 0, maxN = 0;
    for (int b = 0; b < B; b++) {
        M += c->h_blkN[b] - 8;
        maxN = std::max(maxN, c->h_blkN[b]);
    }
    bd.M = M;
    bd.maxN = maxN;
    CKC(c, cudaMemcpyAsync(ws.blkN, c->h_blkN, (size_t)B * 4, cudaMemcpyHostToDevice, st));

    LaunchCounter lc = {0, 0, 0, 0, 0};
    uint64_t algSa = 0;
    int rounds = 0;
    float msSort = 0.f;
    CKC(c, cudaEventRecord(c->ev[0], st));
    if (sa_build(ws, bd, c->sortMode, st, lc, algSa, rounds, &msSort)) { set_err(c, "sa_build failed"); return -1; }
    CKC(c, cudaEventRecord(c->ev[1], st));
    if (upto >= 2 && lcp_build(ws, bd, st, lc)) { set_err(c, "lcp_build failed"); return -1; }
    CKC(c, cudaEventRecord(c->ev[2], st));
    if (upto >= 3 && ansv_build(ws, bd, st, lc)) { set_err(c, "ansv_build failed"); return -1; }
    CKC(c, cudaEventRecord(c->ev[3], st));
    if (upto >= 4 && parse_build(ws, bd, st, lc)) { set_err(c, "parse_build failed"); return -1; }
    CKC(c, cudaEventRecord(c->ev[4], st));
    if (upto >= 5) {
        if (emit_build(ws, bd, framing, st, lc)) { set_err(c, "emit_build failed"); return -1; }
        CKC(c, cudaMemcpyAsync(c->h_outLen, ws.blkOutLen, (size_t)B * 4, cudaMemcpyDeviceToHost, st));
        CKC(c, cudaMemcpyAsync(c->h_salzLen, ws.blkSalzLen, (size_t)B * 4, cudaMemcpyDeviceToHost, st));
        CKC(c, cudaMemcpyAsync(c->h_outOff, ws.blkOutOff, (size_t)(B + 1) * 8, cudaMemcpyDeviceToHost, st));
    }
    CKC(c, cudaEventRecord(c->ev[5], st));
    CKC(c, cudaStreamSynchronize(st));
    float ms[5];
    for (int i = 0; i < 5; i++) CKC(c, cudaEventElapsedTime(&ms[i], c->ev[i], c->ev[i + 1]));
    salz_cuda_stats &s = c->stats;
    s.ms_sa += ms[0]; s.ms_sort += msSort; s.ms_lcp += ms[1]; s.ms_ansv += ms[2]; s.ms_parse += ms[3];
    s.ms_emit += ms[4];
    s.ms_total += ms[0] + ms[1] + ms[2] + ms[3] + ms[4];
    s.launches += lc.launches;
    s.sort_launches += lc.sort_launches;
    s.sort_elems += lc.sort_elems;
    s.sort_bytes += lc.sort_bytes;
    s.doubling_rounds += rounds;
    s.alg_bytes_sa += algSa;
    if (upto >= 5) {
        s.alg_bytes_rest += lc.alg_bytes;   // counted at the launch sites (salz_dev.cuh)
        s.blocks += B;
        for (int b = 0; b < B; b++) {
            s.input_bytes += c->h_blkN[b];
            s.output_bytes += c->h_outLen[b];
            s.alg_bytes_rest += c->h_outLen[b];
        }
    }
    return 0;
}

static int stream_common(salz_cuda_ctx *c, const uint8_t *src, bool srcOnDevice, size_t src_len, size_t block_len,
                         uint8_t *dst, bool dstOnDevice, size_t dst_cap, size_t *dst_len, uint32_t *block_sizes,
                         size_t *nblocks_out)
{
    if (!c || !src || !dst || !dst_len || block_len < 9 || src_len == 0) return -1;
    if (cudaSetDevice(c->device) != cudaSuccess) return -1;
    Workspace &ws = c->ws;
    if (block_len > ws.S) { set_err(c, "block_len %zu exceeds context block capacity %u", block_len, ws.S); return -1; }
    const size_t nblocks = (src_len + block_len - 1) / block_len;
    const size_t tail = src_len - (nblocks - 1) * block_len;
    if (tail <= 8) { set_err(c, "last block has %zu bytes (<= 8): not encodable", tail); return -1; }
    cudaStream_t st = c->stream;
    size_t outPos = 0;
    for (size_t b0 = 0; b0 < nblocks; b0 += ws.Bmax) {
        const int B = (int)std::min(ws.Bmax, nblocks - b0);
        const size_t off = b0 * block_len;
        const size_t bytes = std::min(src_len - off, (size_t)B * block_len);
        for (int b = 0; b < B; b++)
            c->h_blkN[b] = (uint32_t)std::min(block_len, src_len - (b0 + b) * block_len);
        CKC(c, cudaEventRecord(c->ev[6], st));
        CKC(c, cudaMemcpyAsync(ws.T, src + off, bytes, srcOnDevice ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, st));
        CKC(c, cudaEventRecord(c->ev[7], st));
        if (run_batch(c, B, block_len, 1, 5)) return -1;
        float ms;
        CKC(c, cudaEventElapsedTime(&ms, c->ev[6], c->ev[7]));
        c->stats.ms_h2d += ms;
        // The reference writes the SALZ form into a buffer of
        // salz_encoded_len_max(N) bytes before deciding on PLAIN; a nearly
        // incompressible block can overflow it by < 8 bytes and the reference
        // CLI then fails the whole file (ref lib/salz.c:260,274).  Same here.
        for (int b = 0; b < B; b++) {
            // capacity the reference CLI gives every block, the short tail block included
            // (ref programs/salzcli.c:129: salz_encoded_len_max(plain_len))
            if (c->h_salzLen[b] > 4 + block_len + (block_len + 63) / 64 * 8) {
                set_err(c, "block %zu: intermediate stream exceeds salz_encoded_len_max (reference fails too)", b0 + b);
                return -1;
            }
        }
        const size_t produced = (size_t)c->h_outOff[B];
        if (outPos + produced > dst_cap) { set_err(c, "output buffer too small"); return -1; }
        CKC(c, cudaEventRecord(c->ev[6], st));
        CKC(c, cudaMemcpyAsync(dst + outPos, ws.out, produced, dstOnDevice ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, st));
        CKC(c, cudaEventRecord(c->ev[7], st));
        CKC(c, cudaStreamSynchronize(st));
        CKC(c, cudaEventElapsedTime(&ms, c->ev[6], c->ev[7]));
        c->stats.ms_d2h += ms;
        if (block_sizes)
            for (int b = 0; b < B; b++) block_sizes[b0 + b] = c->h_outLen[b];
        outPos += produced;
    }
    *dst_len = outPos;
    if (nblocks_out) *nblocks_out = nblocks;
    return 0;
}

// Host buffers: the same loop with the copies on a second stream and two text / output
// buffers, so that the upload of batch i+1 and the download of batch i-1 run under the
// kernels of batch i.  A stream that would fit one batch is cut in two for the same reason.
// (Pageable host memory makes the copies synchronous; the result is the same.)
constexpr int SPLIT_PCT = 25;  // share of a one-batch stream that goes into its first piece

static int stream_host(salz_cuda_ctx *c, const uint8_t *src, size_t src_len, size_t block_len, uint8_t *dst,
                       size_t dst_cap, size_t *dst_len, uint32_t *block_sizes, size_t *nblocks_out)
{
    if (!c || !src || !dst || !dst_len || block_len < 9 || src_len == 0) return -1;
    if (cudaSetDevice(c->device) != cudaSuccess) return -1;
    Workspace &ws = c->ws;
    if (block_len > ws.S) { set_err(c, "block_len %zu exceeds context block capacity %u", block_len, ws.S); return -1; }
    const size_t nblocks = (src_len + block_len - 1) / block_len;
    const size_t tail = src_len - (nblocks - 1) * block_len;
    if (tail <= 8) { set_err(c, "last block has %zu bytes (<= 8): not encodable", tail); return -1; }
    cudaStream_t st = c->stream, cs = c->copyStream;
    // Batches of Bmax blocks; a stream that fits one batch is cut in two so that most of its
    // upload and download run under kernels.  The first piece is the smaller one: its upload is
    // the copy nothing can hide.
    std::vector<size_t> cut;  // cut[i] = first block of batch i, cut[nbatch] = nblocks
    for (size_t b = 0; b < nblocks; b += ws.Bmax) cut.push_back(b);
    cut.push_back(nblocks);
    if (cut.size() == 2 && nblocks >= 32) {
        static const int pct = [] { const char *e = getenv("SALZ_SPLIT_PCT"); int v = e ? atoi(e) : 0; return (v > 0 && v < 100) ? v : SPLIT_PCT; }();
        cut.insert(cut.begin() + 1, std::max<size_t>(16, nblocks * pct / 100));
    }
    const size_t nbatch = cut.size() - 1;
    auto batch_off = [&](size_t i) { return cut[i] * block_len; };
    auto batch_bytes = [&](size_t i) { return std::min(src_len, cut[i + 1] * block_len) - cut[i] * block_len; };
    // on every return (errors included): no copy still reads the caller's src or writes its dst
    struct Restore {
        salz_cuda_ctx *c;
        ~Restore()
        {
            cudaStreamSynchronize(c->copyStream);
            cudaStreamSynchronize(c->stream);
            c->ws.T = c->Tbuf[0];
            c->ws.out = c->outbuf[0];
        }
    } restore{c};
    // everything queued earlier on the compute stream is done before the buffers are reused
    CKC(c, cudaEventRecord(c->computed, st));
    CKC(c, cudaStreamWaitEvent(cs, c->computed, 0));
    CKC(c, cudaMemcpyAsync(c->Tbuf[0], src, batch_bytes(0), cudaMemcpyHostToDevice, cs));
    CKC(c, cudaEventRecord(c->tReady[0], cs));
    size_t outPos = 0;
    bool outBusy[2] = {false, false};
    for (size_t i = 0; i < nbatch; i++) {
        const int k = (int)(i & 1);
        const size_t b0 = cut[i];
        const int B = (int)(cut[i + 1] - b0);
        if (i + 1 < nbatch) {  // Tbuf[k^1] was batch i-1's text: its kernels are complete (run_batch synchronises)
            CKC(c, cudaMemcpyAsync(c->Tbuf[k ^ 1], src + batch_off(i + 1), batch_bytes(i + 1),
                                   cudaMemcpyHostToDevice, cs));
            CKC(c, cudaEventRecord(c->tReady[k ^ 1], cs));
        }
        for (int b = 0; b < B; b++)
            c->h_blkN[b] = (uint32_t)std::min(block_len, src_len - (b0 + b) * block_len);
        CKC(c, cudaStreamWaitEvent(st, c->tReady[k], 0));
        if (outBusy[k]) CKC(c, cudaStreamWaitEvent(st, c->oDone[k], 0));  // batch i-2's records have left outbuf[k]
        ws.T = c->Tbuf[k];
        ws.out = c->outbuf[k];
        if (run_batch(c, B, block_len, 1, 5)) return -1;
        for (int b = 0; b < B; b++) {
            // capacity the reference CLI gives every block, the short tail block included
            // (ref programs/salzcli.c:129: salz_encoded_len_max(plain_len))
            if (c->h_salzLen[b] > 4 + block_len + (block_len + 63) / 64 * 8) {
                set_err(c, "block %zu: intermediate stream exceeds salz_encoded_len_max (reference fails too)", b0 + b);
                return -1;
            }
        }
        const size_t produced = (size_t)c->h_outOff[B];
        if (outPos + produced > dst_cap) { set_err(c, "output buffer too small"); return -1; }
        CKC(c, cudaMemcpyAsync(dst + outPos, c->outbuf[k], produced, cudaMemcpyDeviceToHost, cs));  // kernels are complete
        CKC(c, cudaEventRecord(c->oDone[k], cs));
        outBusy[k] = true;
        if (block_sizes)
            for (int b = 0; b < B; b++) block_sizes[b0 + b] = c->h_outLen[b];
        outPos += produced;
    }
    CKC(c, cudaStreamSynchronize(cs));
    *dst_len = outPos;
    if (nblocks_out) *nblocks_out = nblocks;
    return 0;
}

extern "C" int salz_cuda_encode_stream(salz_cuda_ctx *c, const uint8_t *src, size_t src_len, size_t block_len,
                                       uint8_t *dst, size_t dst_cap, size_t *dst_len, uint32_t *block_sizes,
                                       size_t *nblocks)
{
    return stream_host(c, src, src_len, block_len, dst, dst_cap, dst_len, block_sizes, nblocks);
}

extern "C" int salz_cuda_encode_stream_device(salz_cuda_ctx *c, const uint8_t *d_src, size_t src_len,
                                              size_t block_len, uint8_t *d_dst, size_t dst_cap, size_t *dst_len,
                                              uint32_t *block_sizes, size_t *nblocks)
{
    return stream_common(c, d_src, true, src_len, block_len, d_dst, true, dst_cap, dst_len, block_sizes, nblocks);
}

extern "C" int salz_cuda_encode_blocks(salz_cuda_ctx *c, const uint8_t *const *srcs, const size_t *src_lens,
                                       uint8_t *const *dsts, size_t *dst_lens, int *rcs, size_t nblocks)
{
    if (!c || !srcs || !src_lens || !dsts || !dst_lens || !rcs) return -1;
    if (cudaSetDevice(c->device) != cudaSuccess) return -1;
    Workspace &ws = c->ws;
    cudaStream_t st = c->stream;
    std::vector<size_t> idx;
    for (size_t i0 = 0; i0 < nblocks;) {
        idx.clear();
        size_t i = i0;
        for (; i < nblocks && idx.size() < ws.Bmax; i++) {
            rcs[i] = -1;
            // ref lib/salz.c:783-786 (NULL buffers), SURVEY F2 (N <= 8 cannot be encoded)
            if (!srcs[i] || !dsts[i] || src_lens[i] <= 8 || src_lens[i] > ws.S || dst_lens[i] < 4) continue;
            idx.push_back(i);
        }
        i0 = i;
        const int B = (int)idx.size();
        if (B == 0) continue;
        for (int b = 0; b < B; b++) {
            c->h_blkN[b] = (uint32_t)src_lens[idx[b]];
            CKC(c, cudaMemcpyAsync(ws.T + (size_t)b * ws.S, srcs[idx[b]], src_lens[idx[b]], cudaMemcpyHostToDevice, st));
        }
        if (run_batch(c, B, ws.S, 0, 5)) return -1;
        for (int b = 0; b < B; b++) {
            const size_t k = idx[b];
            // the reference writes the SALZ form first, so it must fit even when
            // PLAIN is chosen in the end (ref lib/salz.c:260,274,755-767)
            if (c->h_salzLen[b] > dst_lens[k]) continue;
            CKC(c, cudaMemcpyAsync(dsts[k], ws.out + c->h_outOff[b], c->h_outLen[b], cudaMemcpyDeviceToHost, st));
            dst_lens[k] = c->h_outLen[b];
            rcs[k] = 0;
        }
        CKC(c, cudaStreamSynchronize(st));
    }
    return 0;
}

extern "C" int salz_cuda_debug_stages(salz_cuda_ctx *c, const uint8_t *src, size_t N, int32_t *sa, int32_t *plcp,
                                      int32_t *lcp, int32_t *psvnsv, int32_t *cand, int32_t *parse, uint8_t *dst,
                                      size_t *dst_len)
{
    if (!c || !src || N <= 8 || N > c->ws.S) return -1;
    if (cudaSetDevice(c->device) != cudaSuccess) return -1;
    Workspace &ws = c->ws;
    cudaStream_t st = c->stream;
    const size_t n = N - 8;
    c->h_blkN[0] = (uint32_t)N;
    CKC(c, cudaMemcpyAsync(ws.T, src, N, cudaMemcpyHostToDevice, st));
    if (run_batch(c, 1, ws.S, 0, 5)) return -1;
    if (sa) CKC(c, cudaMemcpy(sa, ws.SA, n * 4, cudaMemcpyDeviceToHost));
    if (plcp) CKC(c, cudaMemcpy(plcp, ws.plcp, n * 4, cudaMemcpyDeviceToHost));
    if (lcp) CKC(c, cudaMemcpy(lcp, ws.lcp, n * 4, cudaMemcpyDeviceToHost));
    if (cand) CKC(c, cudaMemcpy(cand, ws.cand, n * 16, cudaMemcpyDeviceToHost));
    if (psvnsv) {
        // the ANSV pass leaves distances (p - PSV, p - NSV; p + 1 when there is none); the
        // reference's aux[4p+0/1] after build_psvnsv_array are the positions themselves
        std::vector<int32_t> cd(4 * n);
        CKC(c, cudaMemcpy(cd.data(), ws.cand, n * 16, cudaMemcpyDeviceToHost));
        psvnsv[0] = psvnsv[1] = -1;  // nothing is smaller than position 0
        for (size_t p = 1; p < n; p++) {
            psvnsv[2 * p] = (int32_t)p - cd[4 * p];
            psvnsv[2 * p + 1] = (int32_t)p - cd[4 * p + 2];
        }
    }
    if (parse) {
        // cost[p] (the reference's aux[4p+2]) = d[p] - 9p, d = relative value + segment base
        std::vector<int32_t> ch(2 * n), cs(n + 1), sb(64);
        CKC(c, cudaMemcpy(ch.data(), ws.choice, n * 8, cudaMemcpyDeviceToHost));
        CKC(c, cudaMemcpy(cs.data(), ws.parseDr, (n + 1) * 4, cudaMemcpyDeviceToHost));
        CKC(c, cudaMemcpy(sb.data(), ws.segBase, 64 * 4, cudaMemcpyDeviceToHost));
        for (size_t p = 0; p < n; p++) {
            parse[3 * p] = ch[2 * p];
            parse[3 * p + 1] = ch[2 * p + 1];
            parse[3 * p + 2] = cs[p] + sb[p >> ws.parseSegLog] - 9 * (int32_t)p;
        }
    }
    if (dst && dst_len) {
        if (c->h_salzLen[0] > *dst_len) return -1;
        CKC(c, cudaMemcpy(dst, ws.out, c->h_outLen[0], cudaMemcpyDeviceToHost));
        *dst_len = c->h_outLen[0];
    }
    return 0;
}

extern "C" int salz_cuda_suffix_arrays(salz_cuda_ctx *c, const uint8_t *texts, const uint32_t *lens, size_t stride,
                                       size_t ntexts, int32_t *sa_out)
{
    if (!c || !texts || !lens) return -1;
    if (cudaSetDevice(c->device) != cudaSuccess) return -1;
    Workspace &ws = c->ws;
    cudaStream_t st = c->stream;
    if (stride > ws.S) return -1;
    for (size_t b0 = 0; b0 < ntexts; b0 += ws.Bmax) {
        const int B = (int)std::min(ws.Bmax, ntexts - b0);
        for (int b = 0; b < B; b++) {
            if (lens[b0 + b] == 0 || lens[b0 + b] > stride || lens[b0 + b] + 8 > ws.S) return -1;
            c->h_blkN[b] = lens[b0 + b] + 8;  // sort the whole text: n = len
        }
        CKC(c, cudaMemcpyAsync(ws.T, texts + b0 * stride, (size_t)B * stride, cudaMemcpyHostToDevice, st));
        if (run_batch(c, B, stride, 0, 1)) return -1;
        if (sa_out)
            for (int b = 0; b < B; b++)
                CKC(c, cudaMemcpyAsync(sa_out + (b0 + b) * stride, ws.SA + (size_t)b * ws.S,
                                       (size_t)lens[b0 + b] * 4, cudaMemcpyDeviceToHost, st));
        CKC(c, cudaStreamSynchronize(st));
    }
    return 0;
}
