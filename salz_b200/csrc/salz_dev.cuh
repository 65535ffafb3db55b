// salz_dev.cuh - internal device-side declarations shared by the .cu files.
//
// Index spaces used throughout (see DESIGN.md "Data layout in HBM"):
//   b      block index inside the current batch, 0 <= b < B
//   p      text position inside a block, 0 <= p < N_b   (n_b = N_b - 8 parsed)
//   g      padded global index  g = b*S + p, S = power of two >= max block len
//   slot   padded global suffix-array index b*S + r, r = rank inside block
//   e,k    dense indices into the (compacted) sort list
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#define SALZ_INF 0x3fffffff

struct BatchDesc {
    int B;             // blocks in this batch
    int logS;          // log2(S)
    uint32_t S;        // index stride per block (power of two)
    size_t Tstride;    // byte stride between blocks in d_T
    uint32_t M;        // sum of n_b
    uint32_t maxN;     // largest N_b
};

struct Workspace {
    int numSMs;
    // capacity
    size_t Bmax;
    uint32_t S;
    int logS;
    size_t Mcap;       // Bmax * S

    uint8_t *T;        // text, Bmax*Tstride + pad
    size_t Tcap;
    uint32_t *blkN;    // [Bmax]   N_b

    // suffix sorting (sa.cu)
    uint64_t *keys[3];   // [Mcap] list elements (packed) or keys (pairs); keys[2] aliases cand (packed only)
    uint32_t *vals[2];   // [Mcap] values of the pairs format (block capacity > 2^21 only)
    uint64_t *rank;      // [Mcap] by g: (stamp:8 | old:24 | new:24) block-local slot of the group head; new = final ISA
    uint32_t *SA;        // [Mcap] by slot: local text position p
    uint32_t *listCnt[3];   // [Bmax]   list elements per block (views 0/1: the bag / pairs list, 2: global path)
    uint32_t *listBase[3];  // [Bmax+1] list offset per block (padded: b << logS)
    uint32_t *tileBase[3];  // [Bmax+1] first tile per block
    uint4 *tileDesc[3];     // [tileCap] (block, tile in block, first element, elements) of every tile
    size_t tileCap;
    uint32_t *cntSnap;   // [Bmax] snapshot of a list count
    uint32_t *alphaMask; // [Bmax][8] byte values present
    uint16_t *codeTab;   // [Bmax][256] dense codes, from 1
    uint32_t *symBits;   // [Bmax] bits per code
    uint32_t *h0;        // [Bmax] characters compared by round 0
    uint32_t *hist;      // [Bmax][8][256] digit histograms for the coming sort
    uint32_t *digitBase; // [Bmax][8][256] exclusive scans of hist
    uint32_t *tileState; // [1 + tileCap][256]: ticket, then look-back states
    uint32_t *sched;     // [8 * tileCap] ticket order of a multi-pass scatter launch (tile | pass << 26)
    uint32_t *passDone;  // [Bmax][8] tiles that have finished (block, pass) in that launch
    int scatterMode;     // 0 = register-staged k_scatter, 1 = TMA-staged, 2 = TMA-staged + two keys per ranking step
    int localCountRound; // first doubling round whose k_local may rank small groups by counting (SALZ_LOCAL_COUNT_ROUND, 3)
    int bagReorder;      // first doubling round before which a large bag is reordered by text position (SALZ_BAG_REORDER, 3; 0 = never)
    int round0Bits;      // > 0: cap on the round-0 key width (SALZ_R0_BITS; fewer radix passes, more left for the doubling rounds)
    int sortGroup;       // > 0: all passes of a sort in one launch, block groups of this many blocks (L2-resident)
    uint32_t *counters; // [16] device scalars: [0]=list length, [1]=tiles
    uint32_t *h_counters; // pinned mirror
    uint32_t *h_blkN;     // pinned host copy of blkN (owned by the context)

    // LCP
    int32_t *plcp;     // [Mcap] by g (first V = PLCP+i for irreducible, then PLCP)
    int32_t *lcp;      // [Mcap] by slot
    int32_t *tileMax;  // [Mcap/1024]
    // min pyramids, levels 1.. (level l has Mcap >> (5l) entries)
    int32_t *pyrSA[6];
    int32_t *pyrLCP[6];
    int nlevels;       // number of levels including level 0

    // ANSV tile summaries, [Mcap / 4096] each
    int32_t *tileMinV, *tileMinL, *tilePreV, *tileSufV;
    // candidates / parse
    int4 *cand;        // [Mcap] by g
    int32_t *cost;     // [Mcap + 64] by g: the parse's d values relative to the segment bases (parse.cu)
    int32_t *cost2;    // [Mcap + 64] the other array of the parse's two (a pass reads the previous pass's values)
    int32_t *parseDr;  // whichever of cost / cost2 holds the final values of the last parse
    int parseSegLog;   // log2 of the segment length of the last parse
    int parseSegments; // requested segments per block, 0 = automatic
    int2 *choice;      // [Mcap] by g: (offs, len)
    uint32_t *jump;    // [Mcap] by g: first position >= chunk end reached from p 
    int32_t *segBase;  // [Bmax][64] parse segments: d at the segment's right end
    // emit
    uint32_t *chunkEntry; // [Mcap / CHUNK]
    uint32_t *chunkTok;   // [Mcap / CHUNK] tokens starting in chunk -> exclusive prefix
    uint32_t *chunkBits;  // [Mcap / CHUNK] bits of those tokens -> exclusive prefix
    uint32_t *tokPos;     // [Mcap] by b*S + t
    uint32_t *tokBit;     // [Mcap] by b*S + t : B_t
    uint32_t *firstTok;   // [Mcap/4] by b*(S/4) + k
    uint32_t *blkTokens;  // [Bmax] m
    uint32_t *blkBits;    // [Bmax] B_m
    uint32_t *blkSalzLen; // [Bmax] 4 + m + 8*ceil(B_m/64)
    uint32_t *blkOutLen;  // [Bmax] final encoded length
    uint64_t *blkOutOff;  // [Bmax+1] offset of block b's record in the packed stream
    uint8_t *out;         // packed output stream
    size_t outCap;
};

#define PARSE_CHUNK 2048      // positions per token-walk chunk
#define PARSE_CHUNK_LOG 11

// ---- small device helpers -------------------------------------------------

// 8 bytes starting at an arbitrary byte address (little-endian), built from two
// aligned 8-byte loads.  The text buffer is padded so addr+15 is readable.
__device__ __forceinline__ uint64_t load_u64_unaligned(const uint8_t *addr)
{
    uintptr_t a = (uintptr_t)addr;
    const uint64_t *w = (const uint64_t *)(a & ~(uintptr_t)7);
    unsigned sh = (unsigned)(a & 7) * 8;
    uint64_t lo = __ldg(w);
    if (sh == 0)
        return lo;
    uint64_t hi = __ldg(w + 1);
    return (lo >> sh) | (hi << (64 - sh));
}

// Bit cost of a factor (ref lib/salz.c:562-608): token bit + 8 direct offset
// bits + 4 bits per vnibble of (offs-1)>>8 + GR3 code of len-3.
__device__ __forceinline__ int vnibble_count(uint32_t v)
{
    int k = 1;
    if (v >= 8) k = 2;
    if (v >= 72) k = 3;
    if (v >= 584) k = 4;
    if (v >= 4680) k = 5;
    if (v >= 37448) k = 6;
    if (v >= 299592) k = 7;
    if (v >= 2396744) k = 8;
    return k;
}
__device__ __forceinline__ int factor_pre_bits(uint32_t offs)
{
    return 1 + 4 * vnibble_count((offs - 1u) >> 8);
}
__device__ __forceinline__ int factor_post_bits(uint32_t len)
{
    return (int)((len - 3u) >> 3) + 4;
}
__device__ __forceinline__ int factor_cost_bits(uint32_t offs, uint32_t len)
{
    return factor_pre_bits(offs) + 8 + factor_post_bits(len);
}

// ---- L2 residency hints ------------------------------------------------------
// Scatters by text position (candidates, ranks, SA slots) write a fraction of a 32-byte
// sector at a time; the other fractions arrive later from other CTAs.  Stores carry an
// evict_last policy so that the partially written sectors wait in the 126 MB L2 for their
// neighbours instead of going to DRAM as read-modify-write pairs; the streaming inputs of
// the same kernels are read evict_first.
#ifndef SALZ_L2_HINTS
#define SALZ_L2_HINTS 1
#endif
__device__ __forceinline__ uint64_t l2_evict_last()
{
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ uint64_t l2_evict_first()
{
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ void st_hint(int4 *a, int4 v, uint64_t pol)
{
#if SALZ_L2_HINTS
    asm volatile("st.global.L2::cache_hint.v4.s32 [%0], {%1, %2, %3, %4}, %5;" ::"l"(a), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w), "l"(pol) : "memory");
#else
    *a = v;
#endif
}
__device__ __forceinline__ void st_hint(int2 *a, int2 v, uint64_t pol)
{
#if SALZ_L2_HINTS
    asm volatile("st.global.L2::cache_hint.v2.s32 [%0], {%1, %2}, %3;" ::"l"(a), "r"(v.x), "r"(v.y), "l"(pol) : "memory");
#else
    *a = v;
#endif
}
__device__ __forceinline__ void st_hint(uint64_t *a, uint64_t v, uint64_t pol)
{
#if SALZ_L2_HINTS
    asm volatile("st.global.L2::cache_hint.u64 [%0], %1, %2;" ::"l"(a), "l"(v), "l"(pol) : "memory");
#else
    *a = v;
#endif
}
__device__ __forceinline__ void st_hint(uint32_t *a, uint32_t v, uint64_t pol)
{
#if SALZ_L2_HINTS
    asm volatile("st.global.L2::cache_hint.u32 [%0], %1, %2;" ::"l"(a), "r"(v), "l"(pol) : "memory");
#else
    *a = v;
#endif
}
__device__ __forceinline__ void st_hint(int32_t *a, int32_t v, uint64_t pol)
{
#if SALZ_L2_HINTS
    asm volatile("st.global.L2::cache_hint.s32 [%0], %1, %2;" ::"l"(a), "r"(v), "l"(pol) : "memory");
#else
    *a = v;
#endif
}
__device__ __forceinline__ uint32_t ld_stream(const uint32_t *a, uint64_t pol)
{
#if SALZ_L2_HINTS
    uint32_t v;
    asm volatile("ld.global.L1::no_allocate.L2::cache_hint.u32 %0, [%1], %2;" : "=r"(v) : "l"(a), "l"(pol));
    return v;
#else
    return *a;
#endif
}
__device__ __forceinline__ int32_t ld_stream(const int32_t *a, uint64_t pol)
{
    return (int32_t)ld_stream((const uint32_t *)a, pol);
}

// ---- host-side launchers (one per pipeline stage) --------------------------
// alg_bytes: algorithmic HBM bytes of the stages after the suffix array, counted where the
// kernels are launched (per position of the batch, M = sum of n_b):
//   k_plcp_irreducible  rank 8 + SA 4 + text ~1 + V 4 = 17      k_plcp_apply  V 4 + rank 8 + PLCP 4 + LCP 4 = 20
//   k_ansv_tile         SA 4 + LCP 4 + candidates 16 = 24        (far items, pyramids: < 1)
//   k_parse             candidates 16 + d 4 per position walked  k_choose  candidates 16 + d 4 + choice 8 + jump 4 = 32
//   emit                choice 8 (walk) + choice 8, text 1, tokPos/tokBit 8 (tokens) + ~4 (words) + output
struct LaunchCounter { uint64_t launches, sort_launches, sort_elems, sort_bytes, alg_bytes; };

// sa.cu
int sa_tile_elems();  // list elements per sort tile
int sa_init_device();    // kernel attributes of the current device (per context)
int sa_build(Workspace &ws, const BatchDesc &bd, int sortMode, cudaStream_t st,
             LaunchCounter &lc, uint64_t &algBytes, int &rounds, float *msSort);
// lcp.cu
int lcp_build(Workspace &ws, const BatchDesc &bd, cudaStream_t st, LaunchCounter &lc);
int pyramid_upper_build(Workspace &ws, const BatchDesc &bd, cudaStream_t st, LaunchCounter &lc);
// ansv.cu
int ansv_build(Workspace &ws, const BatchDesc &bd, cudaStream_t st, LaunchCounter &lc);
// parse.cu
int parse_init_device();
int parse_build(Workspace &ws, const BatchDesc &bd, cudaStream_t st, LaunchCounter &lc);
// emit.cu
int emit_build(Workspace &ws, const BatchDesc &bd, int framing, cudaStream_t st,
               LaunchCounter &lc);
