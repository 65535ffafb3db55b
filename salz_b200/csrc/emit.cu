// emit.cu - token selection and bit/byte encoding with prefix-sum offsets.
//
// Replaces emit_encoding + finalize_encoding and the sequential writer
// (ref lib/salz.c:258-457, 664-775) by the closed form of the stream layout:
//
//   token t (t-th token of the parse, the 8 tail literals included) emits
//   pre-bits, one direct byte, post-bits:
//     literal        pre = '0'                     byte = T[p]          post = -
//     factor (o,l)   pre = '1' + vnibble((o-1)>>8) byte = (o-1)&0xff    post = GR3(l-3)
//   B_t = bits emitted before token t.  Bit words are 64-bit, MSB first, stored
//   little-endian, and each word's 8-byte slot is reserved when its first bit
//   is produced, so
//     word k  sits at  4 + 8k + y_k,  y_k = t   if bit 64k is a pre-bit of token t
//                                           t+1 if it is a post-bit of token t
//     byte of token t sits at  4 + t + 8*ceil((B_t + |pre_t|)/64)
//   total = 4 + m + 8*ceil(B_m/64); PLAIN fallback when total > N + 4.
//
// Every output byte is written exactly once (bytes by their token, words by
// one thread per word), so no zero-fill and no atomics are needed.
#include "salz_dev.cuh"

namespace {

constexpr uint32_t NONE = 0xffffffffu;

struct TokBits { int pre; int post; };

__device__ __forceinline__ TokBits token_bits(int2 c)
{
    TokBits tb;
    if (c.y == 1) { tb.pre = 1; tb.post = 0; }
    else { tb.pre = factor_pre_bits((uint32_t)c.x); tb.post = factor_post_bits((uint32_t)c.y); }
    return tb;
}

// Chunk entries: the positions e_0 = 0, e_{i+1} = jump[e_i] and the chunks they fall in -
// a chain of dependent loads per block (512 for a 1 MiB block), cut short by exact
// speculation.  One warp per block; lane j walks the chunks of range j from the range's
// first position (lane 0: position 0, the true chain).  The true chain then goes through
// the ranges in order: where it steps on a position the range's lane has visited, the rest
// of that lane's walk IS the true chain (the walk is a function of the position) and its
// exit is taken over; until then it walks by itself and erases what the lane had left.
// Paths merge within a chunk or two, so the chain is ~16 + 3 x 31 loads long.
// chunkEntry is NONE everywhere before the launch.  grid = B, 32 threads.
__global__ void __launch_bounds__(32) k_chunk_entries(const uint32_t *__restrict__ blkN,
                                                      const uint32_t *__restrict__ jump, int logS,
                                                      uint32_t *chunkEntry)
{
    __shared__ uint32_t ex[32];
    const uint32_t b = blockIdx.x, lane = threadIdx.x;
    const uint32_t N = blkN[b];
    const size_t gb = (size_t)b << logS;
    uint32_t *ce = chunkEntry + (gb >> PARSE_CHUNK_LOG);
    const uint32_t chunks = (N + PARSE_CHUNK - 1) >> PARSE_CHUNK_LOG;
    const uint32_t perRange = (chunks + 31) / 32;
    {
        uint32_t e = (lane * perRange) << PARSE_CHUNK_LOG;
        const uint32_t lim = min(N, ((lane + 1) * perRange) << PARSE_CHUNK_LOG);
        while (e < lim) {
            ce[e >> PARSE_CHUNK_LOG] = e;
            e = jump[gb + e];
        }
        ex[lane] = e;
    }
    __threadfence_block();
    __syncwarp();
    if (lane != 0) return;
    uint32_t x = ex[0];
    uint32_t clean = min(chunks, perRange);  // chunks below this hold their final value
    while (x < N) {
        uint32_t c = x >> PARSE_CHUNK_LOG;
        const uint32_t j = c / perRange, rEnd = min(chunks, (j + 1) * perRange);
        for (uint32_t k = clean; k < c; k++) ce[k] = NONE;  // chunks the chain has jumped over
        for (;;) {
            if (__ldcg(&ce[c]) == x) {  // on the lane's path: the rest of range j is final
                x = ex[j];
                break;
            }
            const uint32_t nx = jump[gb + x];
            const uint32_t c2 = (nx < N) ? nx >> PARSE_CHUNK_LOG : chunks;  // N itself may lie inside the last chunk
            ce[c] = x;
            for (uint32_t k = c + 1; k < min(c2, rEnd); k++) ce[k] = NONE;
            x = nx;
            c = c2;
            if (c >= rEnd) break;
        }
        clean = rEnd;
    }
    for (uint32_t k = clean; k < chunks; k++) ce[k] = NONE;
}

// One thread per chunk: count the tokens that start in the chunk and their bits.
__global__ void __launch_bounds__(128) k_walk_count(const uint32_t *__restrict__ blkN,
                                                    const int2 *__restrict__ choice,
                                                    const uint32_t *__restrict__ chunkEntry, int logS,
                                                    uint32_t totalChunks, uint32_t *__restrict__ chunkTok,
                                                    uint32_t *__restrict__ chunkBits)
{
    uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= totalChunks) return;
    uint32_t e = chunkEntry[c];
    uint32_t ntok = 0, nbits = 0;
    if (e != NONE) {
        const uint32_t b = c >> (logS - PARSE_CHUNK_LOG);
        const uint32_t N = blkN[b];
        const size_t gb = (size_t)b << logS;
        const uint32_t end = min(N, ((e >> PARSE_CHUNK_LOG) + 1) << PARSE_CHUNK_LOG);
        uint32_t p = e;
        while (p < end) {
            int2 ch = choice[gb + p];
            TokBits tb = token_bits(ch);
            ntok++;
            nbits += tb.pre + tb.post;
            p += (uint32_t)ch.y;
        }
    }
    chunkTok[c] = ntok;
    chunkBits[c] = nbits;
}

// One CTA per block: exclusive scan over its chunks; block totals and sizes.
__global__ void __launch_bounds__(256) k_block_scan(const uint32_t *__restrict__ blkN, uint32_t chunksPerBlock,
                                                    uint32_t *__restrict__ chunkTok, uint32_t *__restrict__ chunkBits,
                                                    uint32_t *__restrict__ blkTokens, uint32_t *__restrict__ blkBits,
                                                    uint32_t *__restrict__ blkSalzLen, uint32_t *__restrict__ blkOutLen)
{
    __shared__ uint32_t pt[256], pb[256];
    const uint32_t b = blockIdx.x;
    uint32_t *ct = chunkTok + (size_t)b * chunksPerBlock;
    uint32_t *cb = chunkBits + (size_t)b * chunksPerBlock;
    uint32_t per = (chunksPerBlock + 255) / 256;
    uint32_t lo = threadIdx.x * per, hi = min(chunksPerBlock, lo + per);
    uint32_t st = 0, sb = 0;
    for (uint32_t i = lo; i < hi; i++) { st += ct[i]; sb += cb[i]; }
    pt[threadIdx.x] = st;
    pb[threadIdx.x] = sb;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t rt = 0, rb = 0;
        for (int i = 0; i < 256; i++) {
            uint32_t a = pt[i], c = pb[i];
            pt[i] = rt; pb[i] = rb;
            rt += a; rb += c;
        }
        const uint32_t N = blkN[b];
        uint32_t salz = 4 + rt + 8 * ((rb + 63) / 64);
        blkTokens[b] = rt;
        blkBits[b] = rb;
        blkSalzLen[b] = salz;
        blkOutLen[b] = (salz > N + 4) ? (N + 4) : salz;
    }
    __syncthreads();
    st = pt[threadIdx.x];
    sb = pb[threadIdx.x];
    for (uint32_t i = lo; i < hi; i++) {
        uint32_t a = ct[i], c = cb[i];
        ct[i] = st; cb[i] = sb;
        st += a; sb += c;
    }
}

// Offsets of the block records in the packed output stream.
__global__ void k_out_offsets(const uint32_t *__restrict__ blkOutLen, int B, int framing,
                              uint64_t *__restrict__ blkOutOff)
{
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        uint64_t off = 0;
        for (int b = 0; b < B; b++) {
            blkOutOff[b] = off;
            off += (uint64_t)blkOutLen[b] + (framing ? 4 : 0);
        }
        blkOutOff[B] = off;
    }
}

// One WARP per chunk.  A thread-per-chunk walk (the first form of this kernel) is a chain of ~400 dependent 8-byte loads per thread,
// 131 K chains each pulling single sectors out of a 2 GB array (1.6 TB/s of DRAM traffic at
// 5 % issue utilisation), and its per-token stores are scattered across the threads.  Here
// the warp loads the chunk's chosen lengths coalesced into shared memory (clamped to 16
// bits: anything longer leaves the chunk anyway), lane 0 walks them there and leaves the
// list of token positions, and then all lanes handle 32 tokens at a time: bit offsets by a
// warp scan, tokPos / tokBit written as coalesced runs.  Same stores, same values.
constexpr int EW = 4;   // warps (chunks) per CTA
__global__ void __launch_bounds__(EW * 32) k_emit_tokens_warp(
    const uint8_t *__restrict__ T, size_t Tstride, const uint32_t *__restrict__ blkN,
    const int2 *__restrict__ choice, const uint32_t *__restrict__ chunkEntry, const uint32_t *__restrict__ chunkTok,
    const uint32_t *__restrict__ chunkBits, const uint32_t *__restrict__ blkSalzLen,
    const uint64_t *__restrict__ blkOutOff, int framing, int logS, uint32_t totalChunks,
    uint32_t *__restrict__ tokPos, uint32_t *__restrict__ tokBit, uint32_t *__restrict__ firstTok,
    uint8_t *__restrict__ out)
{
    __shared__ uint16_t slen[EW][PARSE_CHUNK];
    __shared__ uint16_t stok[EW][PARSE_CHUNK];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const uint32_t c = blockIdx.x * EW + w;
    if (c >= totalChunks) return;
    const uint32_t e = chunkEntry[c];
    if (e == NONE) return;
    const uint32_t b = c >> (logS - PARSE_CHUNK_LOG);
    const uint32_t N = blkN[b];
    if (blkSalzLen[b] > N + 4) return;  // PLAIN block
    const size_t gb = (size_t)b << logS;
    const uint8_t *Tb = T + (size_t)b * Tstride;
    uint8_t *ob = out + blkOutOff[b] + (framing ? 4 : 0);
    uint32_t *tp = tokPos + gb, *tbit = tokBit + gb;
    uint32_t *ft = firstTok + (gb >> 2);
    const uint32_t cs = (e >> PARSE_CHUNK_LOG) << PARSE_CHUNK_LOG;   // first position of the chunk
    const uint32_t cnt = min(N, cs + PARSE_CHUNK) - cs;              // positions of the chunk inside the block
    const int2 *chb = choice + gb + cs;
    for (uint32_t j = lane; j < cnt; j += 32) slen[w][j] = (uint16_t)min((uint32_t)chb[j].y, 0xffffu);
    __syncwarp();
    uint32_t nt = 0;
    if (lane == 0) {
        uint32_t q = e - cs;
        while (q < cnt) {
            stok[w][nt++] = (uint16_t)q;
            q += slen[w][q];
        }
    }
    nt = __shfl_sync(0xffffffffu, nt, 0);
    __syncwarp();
    const uint32_t t0 = chunkTok[c];
    uint32_t Bbase = chunkBits[c];
    for (uint32_t i0 = 0; i0 < nt; i0 += 32) {
        const uint32_t i = i0 + lane;
        const bool live = i < nt;
        uint32_t p = 0, bits = 0;
        int2 ch = make_int2(0, 1);
        TokBits tb = {0, 0};
        if (live) {
            p = cs + stok[w][i];
            ch = choice[gb + p];
            tb = token_bits(ch);
            bits = (uint32_t)(tb.pre + tb.post);
        }
        uint32_t inc = bits;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t v = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += v;
        }
        const uint32_t Bt = Bbase + inc - bits;
        if (live) {
            const uint32_t t = t0 + i;
            tp[t] = p;
            tbit[t] = Bt;
            const uint32_t bytePos = 4 + t + 8 * ((Bt + tb.pre + 63) >> 6);
            ob[bytePos] = (ch.y == 1) ? Tb[p] : (uint8_t)(((uint32_t)ch.x - 1u) & 0xffu);
            const uint32_t Bn = Bt + bits;
            for (uint32_t k = (Bt + 63) >> 6; k < ((Bn + 63) >> 6); k++) ft[k] = t;
        }
        Bbase += __shfl_sync(0xffffffffu, inc, 31);
    }
}


// ORs the bits of a field into word `w`.  The field covers global bits
// [s, s+nb) with its least significant bit last; the word covers [w0, w0+64).
__device__ __forceinline__ void place_field(uint64_t &w, int64_t w0, int64_t s, int nb, uint64_t val)
{
    int64_t rel = s + nb - 1 - w0;  // index of the field's last bit inside the word
    if (rel < 0) return;
    int64_t pos = 63 - rel;
    if (pos >= 0) {
        w |= val << pos;
    } else if (-pos < nb) {
        w |= val >> (-pos);
    }
}

// One thread per 64-bit word of every SALZ block.
__global__ void __launch_bounds__(256) k_emit_words(
    const uint32_t *__restrict__ blkN, const int2 *__restrict__ choice, const uint32_t *__restrict__ blkTokens,
    const uint32_t *__restrict__ blkBits, const uint32_t *__restrict__ blkSalzLen,
    const uint64_t *__restrict__ blkOutOff, int framing, int logS, const uint32_t *__restrict__ tokPos,
    const uint32_t *__restrict__ tokBit, const uint32_t *__restrict__ firstTok, uint8_t *__restrict__ out)
{
    const uint32_t wi = blockIdx.x * blockDim.x + threadIdx.x;   // b * (S/4) + k
    const uint32_t b = wi >> (logS - 2);
    const uint32_t k = wi & ((1u << (logS - 2)) - 1);
    const uint32_t N = blkN[b];
    if (blkSalzLen[b] > N + 4) return;
    const uint32_t Bm = blkBits[b];
    if (k >= ((Bm + 63) >> 6)) return;
    const uint32_t m = blkTokens[b];
    const size_t gb = (size_t)b << logS;
    const uint32_t *tp = tokPos + gb, *tbit = tokBit + gb;
    const int64_t w0 = (int64_t)k << 6;
    const uint32_t t0 = firstTok[(gb >> 2) + k];
    uint64_t w = 0;
    uint32_t y = t0;
    for (uint32_t t = t0; t < m; t++) {
        const uint32_t Bt = tbit[t];
        if ((int64_t)Bt >= w0 + 64) break;
        int2 ch = choice[gb + tp[t]];
        if (ch.y == 1) continue;  // literal: a single '0' bit
        const uint32_t offs = (uint32_t)ch.x, len = (uint32_t)ch.y;
        // pre-bits: '1' then the vnibbles of (offs-1)>>8, first-written nibble on top
        int32_t v = (int32_t)((offs - 1u) >> 8);
        uint64_t pv = (uint64_t)((v & 7) | 8);
        int nn = 1;
        v = (v >> 3) - 1;
        while (v >= 0) {
            pv |= (uint64_t)(v & 7) << (4 * nn);
            nn++;
            v = (v >> 3) - 1;
        }
        const int pre = 1 + 4 * nn;
        pv |= 1ull << (4 * nn);
        place_field(w, w0, (int64_t)Bt, pre, pv);
        if (t == t0 && w0 >= (int64_t)Bt + pre) y = t0 + 1;
        // post-bits: q zeros, '1', three low bits of len-3
        const uint32_t q = (len - 3u) >> 3;
        place_field(w, w0, (int64_t)Bt + pre + q, 4, 8u | ((len - 3u) & 7u));
    }
    uint8_t *dst = out + blkOutOff[b] + (framing ? 4 : 0) + 4 + 8 * (size_t)k + y;
#pragma unroll
    for (int i = 0; i < 8; i++) dst[i] = (uint8_t)(w >> (8 * i));
}

// Block header, optional u32 framing, and the PLAIN copy.
// grid = (S / 4096, B), 256 threads.
__global__ void __launch_bounds__(256) k_finalize(const uint8_t *__restrict__ T, size_t Tstride,
                                                  const uint32_t *__restrict__ blkN,
                                                  const uint32_t *__restrict__ blkSalzLen,
                                                  const uint32_t *__restrict__ blkOutLen,
                                                  const uint64_t *__restrict__ blkOutOff, int framing,
                                                  uint8_t *__restrict__ out)
{
    const uint32_t b = blockIdx.y;
    const uint32_t N = blkN[b];
    const uint32_t salz = blkSalzLen[b];
    const bool plain = salz > N + 4;
    uint8_t *rec = out + blkOutOff[b];
    if (blockIdx.x == 0 && threadIdx.x < 8) {
        // ref lib/salz.c:755-772: type in the top byte, 24-bit length (overflow and all)
        uint32_t hdr = plain ? (N & 0xffffffu) : ((1u << 24) | ((salz - 4) & 0xffffffu));
        uint32_t fr = blkOutLen[b];
        int i = threadIdx.x;
        if (framing) {
            rec[i] = (i < 4) ? (uint8_t)(fr >> (8 * i)) : (uint8_t)(hdr >> (8 * (i - 4)));
        } else if (i < 4) {
            rec[i] = (uint8_t)(hdr >> (8 * i));
        }
    }
    if (plain) {
        uint8_t *dst = rec + (framing ? 4 : 0) + 4;
        const uint8_t *src = T + (size_t)b * Tstride;
        uint32_t lo = blockIdx.x * 4096u;
        for (uint32_t i = lo + threadIdx.x; i < min(N, lo + 4096u); i += 256) dst[i] = src[i];
    }
}

}  // namespace

int emit_build(Workspace &ws, const BatchDesc &bd, int framing, cudaStream_t st, LaunchCounter &lc)
{
    const size_t total = (size_t)bd.B * bd.S;
    const uint32_t totalChunks = (uint32_t)(total >> PARSE_CHUNK_LOG);
    const uint32_t chunksPerBlock = bd.S >> PARSE_CHUNK_LOG;
    cudaMemsetAsync(ws.chunkEntry, 0xff, (size_t)totalChunks * 4, st);
    k_chunk_entries<<<bd.B, 32, 0, st>>>(ws.blkN, ws.jump, bd.logS, ws.chunkEntry);
    k_walk_count<<<(totalChunks + 127) / 128, 128, 0, st>>>(ws.blkN, ws.choice, ws.chunkEntry, bd.logS,
                                                            totalChunks, ws.chunkTok, ws.chunkBits);
    k_block_scan<<<bd.B, 256, 0, st>>>(ws.blkN, chunksPerBlock, ws.chunkTok, ws.chunkBits, ws.blkTokens,
                                       ws.blkBits, ws.blkSalzLen, ws.blkOutLen);
    k_out_offsets<<<1, 32, 0, st>>>(ws.blkOutLen, bd.B, framing, ws.blkOutOff);
    k_emit_tokens_warp<<<(totalChunks + EW - 1) / EW, EW * 32, 0, st>>>(
        ws.T, bd.Tstride, ws.blkN, ws.choice, ws.chunkEntry, ws.chunkTok, ws.chunkBits, ws.blkSalzLen,
        ws.blkOutOff, framing, bd.logS, totalChunks, ws.tokPos, ws.tokBit, ws.firstTok, ws.out);
    k_emit_words<<<(uint32_t)((total >> 2) / 256), 256, 0, st>>>(ws.blkN, ws.choice, ws.blkTokens, ws.blkBits,
                                                                 ws.blkSalzLen, ws.blkOutOff, framing, bd.logS,
                                                                 ws.tokPos, ws.tokBit, ws.firstTok, ws.out);
    dim3 fg(bd.S / 4096, bd.B);
    k_finalize<<<fg, 256, 0, st>>>(ws.T, bd.Tstride, ws.blkN, ws.blkSalzLen, ws.blkOutLen, ws.blkOutOff, framing,
                                   ws.out);
    lc.launches += 7;
    lc.alg_bytes += (uint64_t)bd.M * 29;   // + the output bytes, added by the caller once the sizes are known
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        fprintf(stderr, "salz_cuda: emit_build: %s\n", cudaGetErrorString(e));
        return -1;
    }
    return 0;
}
