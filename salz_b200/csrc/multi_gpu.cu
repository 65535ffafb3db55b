// multi_gpu.cu - blocks of one stream sharded over the GPUs of a box.
//
// Blocks are independent (the reference creates its context per call, ref
// lib/salz.c:788,821, and the CLI frames them as a flat concatenation, ref
// programs/salzcli.c:143-179), so rank r of `world` encodes the contiguous block
// range salz_cuda_block_range(nblocks, r, world) with the single-GPU pipeline and
// nothing crosses between GPUs on the data path.  The one exchange is an NCCL
// all-gather of the per-block encoded sizes; an exclusive scan over them gives
// every record's byte offset in the assembled stream
// (off_b = sum_{j<b} (4 + len_j), ref programs/salzcli.c:162-169).
//
// Two ways in:
//   salz_cuda_comm_*     one rank per caller (a process per GPU under torchrun/MPI, or a
//                        thread per GPU): the caller ships the 128-byte NCCL id itself
//   salz_cuda_multi_*    one process, all devices: a host thread per GPU, records staged
//                        in pinned memory and copied to their offsets after the all-gather
//
// NCCL is bound at run time (dlopen of libnccl.so.2, the soname both the system package
// and the copy bundled with PyTorch carry), so libsalz_cuda.so loads and the single-GPU
// entry points work on a machine without NCCL.
#include "../../include/salz_cuda.h"

#include <cuda_runtime.h>
#include <dlfcn.h>
#include <nccl.h>

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <thread>
#include <vector>

namespace {

struct NcclApi {
    void *handle = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
    ncclResult_t (*GetVersion)(int *) = nullptr;
    bool ok = false;
};

NcclApi &nccl()
{
    static NcclApi api;
    static std::once_flag once;
    std::call_once(once, [] {
        const char *names[] = {"libnccl.so.2", "libnccl.so"};
        for (const char *n : names) {
            api.handle = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
            if (api.handle) break;
        }
        if (!api.handle) {
            fprintf(stderr, "salz_cuda: NCCL not found (%s)\n", dlerror());
            return;
        }
#define BIND(field, sym) *(void **)&api.field = dlsym(api.handle, sym)
        BIND(GetUniqueId, "ncclGetUniqueId");
        BIND(CommInitRank, "ncclCommInitRank");
        BIND(CommDestroy, "ncclCommDestroy");
        BIND(AllGather, "ncclAllGather");
        BIND(GetErrorString, "ncclGetErrorString");
        BIND(GetVersion, "ncclGetVersion");
#undef BIND
        api.ok = api.GetUniqueId && api.CommInitRank && api.CommDestroy && api.AllGather && api.GetErrorString;
        if (!api.ok) fprintf(stderr, "salz_cuda: libnccl lacks a required symbol\n");
    });
    return api;
}

#define NCK(call)                                                                                   \
    do {                                                                                            \
        ncclResult_t r_ = (call);                                                                   \
        if (r_ != ncclSuccess) {                                                                    \
            fprintf(stderr, "salz_cuda: %s: %s (%s:%d)\n", #call, nccl().GetErrorString(r_), __FILE__, __LINE__); \
            return -1;                                                                              \
        }                                                                                           \
    } while (0)
#define CUK(call)                                                                                   \
    do {                                                                                            \
        cudaError_t e_ = (call);                                                                    \
        if (e_ != cudaSuccess) {                                                                    \
            fprintf(stderr, "salz_cuda: %s: %s (%s:%d)\n", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
            return -1;                                                                              \
        }                                                                                           \
    } while (0)

}  // namespace

struct salz_cuda_comm {
    int device, rank, world;
    ncclComm_t comm;
    cudaStream_t stream;
    uint32_t *d_send, *d_recv;   // [per], [world * per]
    uint32_t *h_recv;            // pinned mirror of d_recv
    uint32_t *h_send;            // pinned
    size_t per;                  // capacity in sizes per rank
    uint64_t gathers;            // all-gathers executed (bench: proof the exchange ran)
};

extern "C" void salz_cuda_block_range(size_t nblocks, int rank, int world, size_t *lo, size_t *hi)
{
    const size_t base = nblocks / (size_t)world, rem = nblocks % (size_t)world;
    const size_t r = (size_t)rank;
    const size_t l = r * base + std::min(r, rem);
    if (lo) *lo = l;
    if (hi) *hi = l + base + (r < rem ? 1 : 0);
}

extern "C" int salz_cuda_comm_unique_id(uint8_t id[SALZ_CUDA_UNIQUE_ID_BYTES])
{
    static_assert(sizeof(ncclUniqueId) == SALZ_CUDA_UNIQUE_ID_BYTES, "NCCL unique id size");
    if (!id || !nccl().ok) return -1;
    ncclUniqueId u;
    NCK(nccl().GetUniqueId(&u));
    memcpy(id, &u, sizeof(u));
    return 0;
}

static int comm_reserve(salz_cuda_comm *c, size_t per)
{
    if (per <= c->per) return 0;
    per = std::max<size_t>(per, 1024);
    if (c->d_send) cudaFree(c->d_send);
    if (c->d_recv) cudaFree(c->d_recv);
    if (c->h_recv) cudaFreeHost(c->h_recv);
    if (c->h_send) cudaFreeHost(c->h_send);
    c->d_send = c->d_recv = c->h_recv = c->h_send = nullptr;
    c->per = 0;
    CUK(cudaMalloc((void **)&c->d_send, per * 4));
    CUK(cudaMalloc((void **)&c->d_recv, per * 4 * (size_t)c->world));
    CUK(cudaMallocHost((void **)&c->h_recv, per * 4 * (size_t)c->world));
    CUK(cudaMallocHost((void **)&c->h_send, per * 4));
    c->per = per;
    return 0;
}

extern "C" salz_cuda_comm *salz_cuda_comm_create(int device, const uint8_t id[SALZ_CUDA_UNIQUE_ID_BYTES], int rank,
                                                 int world)
{
    if (!id || world < 1 || rank < 0 || rank >= world || !nccl().ok) return nullptr;
    if (cudaSetDevice(device) != cudaSuccess) return nullptr;
    salz_cuda_comm *c = new salz_cuda_comm();
    memset(c, 0, sizeof(*c));
    c->device = device;
    c->rank = rank;
    c->world = world;
    ncclUniqueId u;
    memcpy(&u, id, sizeof(u));
    ncclResult_t r = nccl().CommInitRank(&c->comm, world, u, rank);
    if (r != ncclSuccess) {
        fprintf(stderr, "salz_cuda: ncclCommInitRank(rank %d of %d, device %d): %s\n", rank, world, device,
                nccl().GetErrorString(r));
        delete c;
        return nullptr;
    }
    if (cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess || comm_reserve(c, 1024)) {
        salz_cuda_comm_destroy(c);
        return nullptr;
    }
    return c;
}

extern "C" void salz_cuda_comm_destroy(salz_cuda_comm *c)
{
    if (!c) return;
    cudaSetDevice(c->device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    if (c->comm) nccl().CommDestroy(c->comm);
    if (c->d_send) cudaFree(c->d_send);
    if (c->d_recv) cudaFree(c->d_recv);
    if (c->h_recv) cudaFreeHost(c->h_recv);
    if (c->h_send) cudaFreeHost(c->h_send);
    if (c->stream) cudaStreamDestroy(c->stream);
    delete c;
}

extern "C" int salz_cuda_comm_rank(const salz_cuda_comm *c) { return c ? c->rank : -1; }
extern "C" int salz_cuda_comm_world(const salz_cuda_comm *c) { return c ? c->world : -1; }
extern "C" uint64_t salz_cuda_comm_gathers(const salz_cuda_comm *c) { return c ? c->gathers : 0; }

// Every rank passes the sizes of its own block range (salz_cuda_block_range of
// nblocks_total); all ranks get all sizes in stream order and, optionally, the record
// offsets (nblocks_total + 1 entries, the last one = length of the assembled stream).
extern "C" int salz_cuda_comm_allgather_sizes(salz_cuda_comm *c, const uint32_t *local_sizes, size_t nlocal,
                                              size_t nblocks_total, uint32_t *all_sizes, uint64_t *record_offsets)
{
    if (!c || (!local_sizes && nlocal) || !all_sizes || nblocks_total == 0) return -1;
    size_t lo, hi;
    salz_cuda_block_range(nblocks_total, c->rank, c->world, &lo, &hi);
    if (hi - lo != nlocal) {
        fprintf(stderr, "salz_cuda: rank %d passes %zu sizes, its range of %zu blocks holds %zu\n", c->rank, nlocal,
                nblocks_total, hi - lo);
        return -1;
    }
    CUK(cudaSetDevice(c->device));
    const size_t per = (nblocks_total + (size_t)c->world - 1) / (size_t)c->world;
    if (comm_reserve(c, per)) return -1;
    memset(c->h_send, 0, per * 4);
    if (nlocal) memcpy(c->h_send, local_sizes, nlocal * 4);
    CUK(cudaMemcpyAsync(c->d_send, c->h_send, per * 4, cudaMemcpyHostToDevice, c->stream));
    NCK(nccl().AllGather(c->d_send, c->d_recv, per, ncclUint32, c->comm, c->stream));
    CUK(cudaMemcpyAsync(c->h_recv, c->d_recv, per * 4 * (size_t)c->world, cudaMemcpyDeviceToHost, c->stream));
    CUK(cudaStreamSynchronize(c->stream));
    c->gathers++;
    for (int r = 0; r < c->world; r++) {
        size_t l, h;
        salz_cuda_block_range(nblocks_total, r, c->world, &l, &h);
        memcpy(all_sizes + l, c->h_recv + (size_t)r * per, (h - l) * 4);
    }
    if (record_offsets) {
        uint64_t off = 0;
        for (size_t b = 0; b < nblocks_total; b++) {
            record_offsets[b] = off;
            off += 4ull + all_sizes[b];
        }
        record_offsets[nblocks_total] = off;
    }
    return 0;
}

// ---- one process, all devices -------------------------------------------------------------

struct salz_cuda_multi {
    int ndev;
    std::vector<int> devices;
    std::vector<salz_cuda_ctx *> ctx;
    std::vector<salz_cuda_comm *> comm;
    std::vector<uint8_t *> stage;      // pinned staging of a rank's records
    std::vector<size_t> stageCap;
    size_t maxBlockLen, maxBatchBlocks;
    char err[256];
};

extern "C" salz_cuda_multi *salz_cuda_multi_create(const int *devices, int ndev, size_t max_block_len,
                                                   size_t max_batch_blocks)
{
    const int have = salz_cuda_device_count();
    if (ndev <= 0) ndev = have;            // all visible devices
    if (ndev <= 0 || ndev > have) {
        fprintf(stderr, "salz_cuda: %d device(s) requested, %d visible; there is no CPU fallback\n", ndev, have);
        return nullptr;
    }
    salz_cuda_multi *m = new salz_cuda_multi();
    m->ndev = ndev;
    m->maxBlockLen = max_block_len;
    m->maxBatchBlocks = max_batch_blocks;
    m->err[0] = 0;
    m->ctx.assign(ndev, nullptr);
    m->comm.assign(ndev, nullptr);
    m->stage.assign(ndev, nullptr);
    m->stageCap.assign(ndev, 0);
    for (int i = 0; i < ndev; i++) m->devices.push_back(devices ? devices[i] : i);
    // One device: nothing to exchange, no communicator (and no dependence on NCCL being
    // installed); the stream goes straight through that device's context.
    uint8_t id[SALZ_CUDA_UNIQUE_ID_BYTES];
    bool ok = ndev == 1 || salz_cuda_comm_unique_id(id) == 0;
    // ncclCommInitRank blocks until every rank has joined: one thread per device
    std::vector<std::thread> th;
    for (int i = 0; i < ndev && ok; i++)
        th.emplace_back([m, i, &id, max_block_len, max_batch_blocks] {
            m->ctx[i] = salz_cuda_ctx_create(m->devices[i], max_block_len, max_batch_blocks);
            if (m->ndev > 1) m->comm[i] = salz_cuda_comm_create(m->devices[i], id, i, m->ndev);
        });
    for (auto &t : th) t.join();
    for (int i = 0; i < ndev; i++) ok = ok && m->ctx[i] && (ndev == 1 || m->comm[i]);
    if (!ok) {
        fprintf(stderr, "salz_cuda: multi-GPU context creation failed\n");
        salz_cuda_multi_destroy(m);
        return nullptr;
    }
    return m;
}

extern "C" void salz_cuda_multi_destroy(salz_cuda_multi *m)
{
    if (!m) return;
    // ncclCommDestroy of the ranks of one communicator may wait for one another
    std::vector<std::thread> th;
    for (int i = 0; i < m->ndev; i++)
        th.emplace_back([m, i] {
            if (m->comm[i]) salz_cuda_comm_destroy(m->comm[i]);
            if (m->ctx[i]) salz_cuda_ctx_destroy(m->ctx[i]);
            if (m->stage[i]) { cudaSetDevice(m->devices[i]); cudaFreeHost(m->stage[i]); }
        });
    for (auto &t : th) t.join();
    delete m;
}

extern "C" int salz_cuda_multi_devices(const salz_cuda_multi *m) { return m ? m->ndev : 0; }
extern "C" salz_cuda_ctx *salz_cuda_multi_ctx(salz_cuda_multi *m, int i) { return (m && i >= 0 && i < m->ndev) ? m->ctx[i] : nullptr; }
extern "C" const char *salz_cuda_multi_last_error(const salz_cuda_multi *m) { return m ? m->err : "no context"; }

extern "C" int salz_cuda_multi_encode_stream(salz_cuda_multi *m, const uint8_t *src, size_t src_len, size_t block_len,
                                             uint8_t *dst, size_t dst_cap, size_t *dst_len, uint32_t *block_sizes,
                                             size_t *nblocks_out)
{
    if (!m || !src || !dst || !dst_len || block_len < 9 || src_len == 0) return -1;
    const size_t nblocks = (src_len + block_len - 1) / block_len;
    const size_t tail = src_len - (nblocks - 1) * block_len;
    if (tail <= 8) {
        snprintf(m->err, sizeof(m->err), "last block has %zu bytes (<= 8): not encodable", tail);
        return -1;
    }
    const int G = m->ndev;
    if (G == 1) {
        const int r = salz_cuda_encode_stream(m->ctx[0], src, src_len, block_len, dst, dst_cap, dst_len, block_sizes, nblocks_out);
        if (r != 0) snprintf(m->err, sizeof(m->err), "device %d failed: %s", m->devices[0], salz_cuda_last_error(m->ctx[0]));
        return r;
    }
    std::vector<int> rc(G, 0);
    std::vector<size_t> produced(G, 0);
    std::vector<std::vector<uint32_t>> all(G);       // every rank's view of all sizes (must agree)
    std::vector<std::vector<uint64_t>> offs(G);
    std::vector<std::thread> th;
    for (int r = 0; r < G; r++)
        th.emplace_back([&, r] {
            size_t lo, hi;
            salz_cuda_block_range(nblocks, r, G, &lo, &hi);
            std::vector<uint32_t> mine(hi - lo);
            all[r].assign(nblocks, 0);
            offs[r].assign(nblocks + 1, 0);
            if (hi > lo) {
                const size_t off = lo * block_len;
                const size_t bytes = std::min(src_len, hi * block_len) - off;
                const size_t need = bytes + (hi - lo) * (block_len / 8 + 24) + 64;
                if (need > m->stageCap[r]) {
                    cudaSetDevice(m->devices[r]);
                    if (m->stage[r]) cudaFreeHost(m->stage[r]);
                    m->stage[r] = nullptr;
                    m->stageCap[r] = 0;
                    if (cudaMallocHost((void **)&m->stage[r], need) != cudaSuccess) rc[r] = -1;
                    else m->stageCap[r] = need;
                }
                if (rc[r] == 0 && salz_cuda_encode_stream(m->ctx[r], src + off, bytes, block_len, m->stage[r],
                                                          m->stageCap[r], &produced[r], mine.data(), nullptr) != 0)
                    rc[r] = -1;
            }
            // a failed rank still joins the collective (zero sizes) so that nobody hangs
            if (rc[r] != 0) std::fill(mine.begin(), mine.end(), 0u);
            if (salz_cuda_comm_allgather_sizes(m->comm[r], mine.data(), mine.size(), nblocks, all[r].data(),
                                               offs[r].data()) != 0)
                rc[r] = -1;
        });
    for (auto &t : th) t.join();
    for (int r = 0; r < G; r++)
        if (rc[r] != 0) {
            snprintf(m->err, sizeof(m->err), "rank %d (device %d) failed: %s", r, m->devices[r],
                     salz_cuda_last_error(m->ctx[r]));
            return -1;
        }
    for (size_t b = 0; b < nblocks; b++)
        if (all[0][b] == 0) {
            snprintf(m->err, sizeof(m->err), "block %zu has no size after the all-gather", b);
            return -1;
        }
    const uint64_t total = offs[0][nblocks];
    if (total > dst_cap) {
        snprintf(m->err, sizeof(m->err), "output buffer too small (%llu needed)", (unsigned long long)total);
        return -1;
    }
    // every rank copies its staged records to the offset the scan over the gathered sizes gave it
    th.clear();
    for (int r = 0; r < G; r++)
        th.emplace_back([&, r] {
            size_t lo, hi;
            salz_cuda_block_range(nblocks, r, G, &lo, &hi);
            if (hi <= lo) return;
            if (offs[r][hi] - offs[r][lo] != produced[r]) { rc[r] = -1; return; }
            memcpy(dst + offs[r][lo], m->stage[r], produced[r]);
        });
    for (auto &t : th) t.join();
    for (int r = 0; r < G; r++)
        if (rc[r] != 0) {
            snprintf(m->err, sizeof(m->err), "rank %d: staged bytes disagree with the gathered sizes", r);
            return -1;
        }
    *dst_len = (size_t)total;
    if (block_sizes) memcpy(block_sizes, all[0].data(), nblocks * 4);
    if (nblocks_out) *nblocks_out = nblocks;
    return 0;
}
