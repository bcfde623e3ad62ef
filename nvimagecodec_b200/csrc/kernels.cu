// Hand-written sm_100a kernels of the b200 baseline-JPEG decode path.
//
//   K0  scan_count / scan_prefix / scan_scatter : marker + restart-interval scan, byte un-stuffing
//   K1  huff_sync   : self-synchronising Huffman decode (state only), intra-sequence fixed point
//   K2  verify_scan : sequence-boundary verification + segmented prefix pass (block index, DC predictors)
//   K3  huff_write  : re-decode from verified states, scatter int16 coefficients
//   K4a idct        : de-zigzag + dequantise + ISLOW 8x8 integer IDCT -> component planes
//   K4b color       : chroma upsampling (fancy / box) + YCbCr->RGB + store into the caller's buffer
//
// They replace what the reference gets from libjpeg-turbo behind
// extensions/libjpeg_turbo/jpeg_mem.cpp:102-463 (jdhuff / jidctint / jdsample / jdcolor);
// the arithmetic is the one written out in SURVEY.md 8(c) and checked by oracle/jpeg_oracle.c.
// No tensor cores: the work is bit-serial entropy decode and 8x8 integer butterflies.
#include "kernels.cuh"

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <mutex>

#include "b200jpeg.h"

#ifndef B200_K1_CHAIN
#define B200_K1_CHAIN 3 // symbols per pair-table entry at most (measured 2: 0.714 ms, 3: 0.694 ms on C2)
#endif

namespace b200jpeg {

namespace {

// ------------------------------------------------------------------------------------------------
// K0: marker scan + un-stuffing
// ------------------------------------------------------------------------------------------------
// Byte classes inside the entropy-coded segment (T.81 B.1.1.5, F.1.2.3):
//   FF 00        -> data byte FF (the 00 is dropped)
//   FF FF        -> the first FF is a fill byte (dropped)
//   FF D0..D7    -> RSTm: both bytes dropped, a restart boundary sits here
//   FF xx other  -> a real marker (EOI, ...): end of the entropy-coded segment
struct ByteMasks {
    uint32_t kept, rst, end;
};

__device__ __noinline__ ByteMasks classify16_tail(const uint8_t* __restrict__ raw, uint32_t len, uint32_t i0)
{
    ByteMasks m{0u, 0u, 0u};
    if (i0 >= len)
        return m;
    const uint4 v = __ldg(reinterpret_cast<const uint4*>(raw + i0));
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
    uint32_t prv = i0 ? __ldg(raw + i0 - 1) : 0u;
    const uint32_t after = (i0 + 16 < len) ? __ldg(raw + i0 + 16) : 0xD9u;
#pragma unroll
    for (int j = 0; j < 16; ++j) {
        const uint32_t cur = (w[j >> 2] >> (8 * (j & 3))) & 0xFFu;
        uint32_t nxt = (j < 15) ? ((w[(j + 1) >> 2] >> (8 * ((j + 1) & 3))) & 0xFFu) : after;
        if (i0 + j + 1 >= len)
            nxt = 0xD9u;
        const bool in = (i0 + j) < len;
        const bool after_ff = (prv == 0xFFu);
        const bool keep = (cur == 0xFFu) ? (nxt == 0u) : !after_ff;
        const bool is_rst = after_ff && ((cur & 0xF8u) == 0xD0u);
        const bool is_end = after_ff && cur != 0u && cur != 0xFFu && !is_rst;
        if (in) {
            m.kept |= uint32_t(keep) << j;
            m.rst |= uint32_t(is_rst) << j;
            m.end |= uint32_t(is_end) << j;
        }
        prv = cur;
    }
    return m;
}

// The same classification four bytes at a time (all but the one chunk of a segment that holds its last byte).
// flag words carry one bit per byte, in bit 7 of the byte.
__device__ __forceinline__ uint32_t zero_bytes(uint32_t v) // 0x80 in every byte of v that is 0 (exact, no carries across bytes)
{
    const uint32_t t = (v & 0x7F7F7F7Fu) + 0x7F7F7F7Fu;
    return ~(t | v | 0x7F7F7F7Fu);
}
__device__ __forceinline__ uint32_t gather_bit7(uint32_t m) // bit 7 of bytes 0..3 -> bits 0..3
{
    return (m * 0x00204081u) >> 28;
}
__device__ __forceinline__ ByteMasks classify16(const uint8_t* __restrict__ raw, uint32_t len, uint32_t i0)
{
    if (i0 + 16u >= len) // (also i0 >= len)
        return classify16_tail(raw, len, i0);
    const uint4 v = __ldg(reinterpret_cast<const uint4*>(raw + i0));
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
    const uint32_t prv = i0 ? __ldg(raw + i0 - 1) : 0u;
    const uint32_t after = __ldg(raw + i0 + 16);
    uint32_t ff[4], zz[4], rs[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        ff[k] = zero_bytes(~w[k]);
        zz[k] = zero_bytes(w[k]);
        rs[k] = zero_bytes((w[k] & 0xF8F8F8F8u) ^ 0xD0D0D0D0u);
    }
    const uint32_t ff_before = (prv == 0xFFu) ? 0x80000000u : 0u; // flag of the byte before the chunk, in its top byte
    const uint32_t zz_after = (after == 0u) ? 0x00000080u : 0u;   // flag of the byte after the chunk, in its low byte
    ByteMasks m{0u, 0u, 0u};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const uint32_t pf = __funnelshift_l(k ? ff[k - 1] : ff_before, ff[k], 8);  // byte j: was byte j-1 an FF?
        const uint32_t nz = __funnelshift_r(zz[k], k < 3 ? zz[k + 1] : zz_after, 8); // byte j: is byte j+1 a 00?
        const uint32_t keep = (ff[k] & nz) | (~ff[k] & ~pf & 0x80808080u);
        const uint32_t rst = pf & rs[k];
        const uint32_t end = pf & ~zz[k] & ~ff[k] & ~rs[k];
        m.kept |= gather_bit7(keep) << (4 * k);
        m.rst |= gather_bit7(rst) << (4 * k);
        m.end |= gather_bit7(end) << (4 * k);
    }
    return m;
}

__device__ __forceinline__ uint32_t block_min_u32(uint32_t v, uint32_t* s_tmp)
{
    v = __reduce_min_sync(0xffffffffu, v);
    if (threadIdx.x == 0)
        *s_tmp = 0xFFFFFFFFu;
    __syncthreads();
    if ((threadIdx.x & 31) == 0)
        atomicMin(s_tmp, v);
    __syncthreads();
    return *s_tmp;
}

// Staging-buffer swizzle for K0c: a thread compacts 64 consecutive input bytes, so without it the
// 32 lanes of a byte store would hit two shared-memory banks (16-way conflict).  XOR-ing the
// word-in-row index with the 128-byte row number spreads them over all 32 banks; the sequential
// read-out stays conflict-free because lanes of one row see the same XOR constant.
__device__ __forceinline__ uint32_t stage_swz(uint32_t byte_index) { return byte_index ^ (((byte_index >> 7) & 15u) << 2); }

__device__ __forceinline__ uint32_t valid_mask16(uint32_t e, uint32_t i0)
{
    if (e == 0xFFFFFFFFu)
        return 0xFFFFu;
    return (e <= i0) ? 0u : ((e - i0 >= 16) ? 0xFFFFu : ((1u << (e - i0)) - 1u));
}

// K0a: per-tile counts.  A thread owns kScanChunks consecutive 16-byte chunks (64 contiguous bytes),
// so a CTA streams 16 KiB with four independent 128-bit loads in flight per thread.
__global__ void __launch_bounds__(kScanThreads) scan_count_kernel(BatchDev bd)
{
    __shared__ uint32_t s_min, s_kept, s_rst;
    const uint32_t tile = blockIdx.x;
    const ImageDev& im = bd.img[bd.tile_img[tile]];
    const uint8_t* raw = bd.raw + im.raw_off;
    const uint32_t i0 = (tile - im.tile_off) * kScanTileBytes + threadIdx.x * (16 * kScanChunks);
    ByteMasks m[kScanChunks];
    uint32_t my_end = 0xFFFFFFFFu;
#pragma unroll
    for (int c = 0; c < kScanChunks; ++c) {
        m[c] = classify16(raw, im.raw_len, i0 + 16 * c);
        if (m[c].end && my_end == 0xFFFFFFFFu)
            my_end = i0 + 16 * c + __ffs(m[c].end) - 1;
    }
    if (threadIdx.x == 0) {
        s_kept = 0;
        s_rst = 0;
    }
    const uint32_t e = block_min_u32(my_end, &s_min); // includes the barriers that publish s_kept/s_rst
    uint32_t k = 0, r = 0;
#pragma unroll
    for (int c = 0; c < kScanChunks; ++c) {
        const uint32_t valid = valid_mask16(e, i0 + 16 * c);
        k += __popc(m[c].kept & valid);
        r += __popc(m[c].rst & valid);
    }
    k = __reduce_add_sync(0xffffffffu, k);
    r = __reduce_add_sync(0xffffffffu, r);
    if ((threadIdx.x & 31) == 0) {
        atomicAdd(&s_kept, k);
        atomicAdd(&s_rst, r);
    }
    __syncthreads();
    if (threadIdx.x == 0)
        bd.tile_info[tile] = make_uint2(s_kept | (e != 0xFFFFFFFFu ? 0x80000000u : 0u), s_rst);
}

// One warp per image: exclusive prefix over its tiles, stops at the tile holding the end marker.
__global__ void __launch_bounds__(128) scan_prefix_kernel(BatchDev bd)
{
    const int img = blockIdx.x * 4 + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (img >= bd.n_images)
        return;
    const ImageDev& im = bd.img[img];
    uint32_t carry_k = 0, carry_r = 0;
    bool ended = false;
    for (uint32_t c = 0; c < im.n_tiles; c += 32) {
        const uint32_t t = c + lane;
        uint2 info = make_uint2(0u, 0u);
        if (t < im.n_tiles)
            info = bd.tile_info[im.tile_off + t];
        uint32_t k = info.x & 0x7FFFFFFFu, r = info.y;
        const uint32_t endmask = __ballot_sync(0xffffffffu, (info.x >> 31) != 0u);
        const int first_end = endmask ? (__ffs(endmask) - 1) : 32;
        const bool live = !ended && lane <= first_end;
        if (!live) {
            k = 0;
            r = 0;
        }
        uint32_t ik = k, ir = r;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t a = __shfl_up_sync(0xffffffffu, ik, d);
            const uint32_t b = __shfl_up_sync(0xffffffffu, ir, d);
            if (lane >= d) {
                ik += a;
                ir += b;
            }
        }
        if (t < im.n_tiles)
            bd.tile_base[im.tile_off + t] = live ? make_uint2(carry_k + ik - k, carry_r + ir - r) : make_uint2(0xFFFFFFFFu, 0u);
        carry_k += __shfl_sync(0xffffffffu, ik, 31);
        carry_r += __shfl_sync(0xffffffffu, ir, 31);
        if (endmask)
            ended = true;
    }
    if (lane == 0) {
        ImageState st;
        st.cmp_len = carry_k;
        st.n_rst = carry_r;
        st.flags = (carry_r != im.n_rst_expected) ? kFlagBadMarker : 0u;
        // (images without subsequences of their own -- the scans of a progressive stream -- stay out of K1-K3)
        st.n_sub = (st.flags || im.n_sub_max == 0u) ? 0u : min(im.n_sub_max, (carry_k * 8u + bd.sub_bits - 1u) / bd.sub_bits);
        bd.state[img] = st;
        const uint32_t slot = carry_r < im.n_bnd_cap ? carry_r : im.n_bnd_cap - 1;
        bd.bnd[im.bnd_off + slot] = carry_k; // end of stream closes the last restart interval
    }
}

__global__ void __launch_bounds__(kScanThreads) scan_scatter_kernel(BatchDev bd)
{
    __shared__ uint32_t s_min;
    __shared__ uint32_t s_warp[kScanThreads / 32];
    __shared__ __align__(16) uint8_t s_stage[kScanTileBytes];
    const uint32_t tile = blockIdx.x;
    const uint32_t img = bd.tile_img[tile];
    const ImageDev& im = bd.img[img];
    const uint2 base = bd.tile_base[tile];
    if (base.x == 0xFFFFFFFFu)
        return; // past the end marker
    const uint8_t* raw = bd.raw + im.raw_off;
    const uint32_t i0 = (tile - im.tile_off) * kScanTileBytes + threadIdx.x * (16 * kScanChunks);
    ByteMasks m[kScanChunks];
    uint32_t my_end = 0xFFFFFFFFu;
#pragma unroll
    for (int c = 0; c < kScanChunks; ++c) {
        m[c] = classify16(raw, im.raw_len, i0 + 16 * c);
        if (m[c].end && my_end == 0xFFFFFFFFu)
            my_end = i0 + 16 * c + __ffs(m[c].end) - 1;
    }
    const uint32_t e = block_min_u32(my_end, &s_min);
    uint32_t mine = 0;
#pragma unroll
    for (int c = 0; c < kScanChunks; ++c) {
        const uint32_t valid = valid_mask16(e, i0 + 16 * c);
        m[c].kept &= valid;
        m[c].rst &= valid;
        mine += __popc(m[c].kept) | (__popc(m[c].rst) << 16);
    }
    // block-wide exclusive scan of (kept count | rst count << 16)
    uint32_t inc = mine;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t a = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= d)
            inc += a;
    }
    if (lane == 31)
        s_warp[warp] = inc;
    __syncthreads();
    uint32_t woff = 0, total = 0;
#pragma unroll
    for (int i = 0; i < kScanThreads / 32; ++i) {
        const uint32_t v = s_warp[i];
        if (i < warp)
            woff += v;
        total += v;
    }
    const uint32_t excl = woff + inc - mine;
    uint32_t lk = excl & 0xFFFFu, lr = excl >> 16;
    const uint32_t tile_kept = total & 0xFFFFu;
    // stage kept bytes, record restart boundaries
#pragma unroll
    for (int c = 0; c < kScanChunks; ++c) {
        const uint32_t kept = m[c].kept, rst = m[c].rst;
        if (kept | rst) {
            const uint4 v = __ldg(reinterpret_cast<const uint4*>(raw + i0 + 16 * c));
            const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int j = 0; j < 16; ++j) {
                const uint32_t cur = (w[j >> 2] >> (8 * (j & 3))) & 0xFFu;
                if ((kept >> j) & 1u)
                    s_stage[stage_swz(lk++)] = uint8_t(cur);
                if ((rst >> j) & 1u) {
                    const uint32_t idx = base.y + lr++;
                    if (idx + 1 < im.n_bnd_cap)
                        bd.bnd[im.bnd_off + idx] = base.x + lk;
                    if (cur != 0xD0u + (idx & 7u))
                        atomicOr(&bd.state[img].flags, kFlagBadMarker);
                }
            }
        }
    }
    __syncthreads();
    // write out as big-endian 32-bit words (a little-endian word load then yields MSB-first bits)
    uint32_t* dst32 = bd.cmp + im.cmp_off / 4;
    uint8_t* dst8 = reinterpret_cast<uint8_t*>(dst32);
    const uint32_t g0 = base.x, g1 = base.x + tile_kept;
    const uint32_t w0 = (g0 + 3) / 4, w1 = g1 / 4;
    if (w1 > w0) {
        for (uint32_t wi = w0 + threadIdx.x; wi < w1; wi += kScanThreads) {
            const uint32_t li = wi * 4 - g0;
            dst32[wi] = (uint32_t(s_stage[stage_swz(li)]) << 24) | (uint32_t(s_stage[stage_swz(li + 1)]) << 16) |
                        (uint32_t(s_stage[stage_swz(li + 2)]) << 8) | uint32_t(s_stage[stage_swz(li + 3)]);
        }
        const uint32_t head = w0 * 4 - g0, tail = g1 - w1 * 4; // < 4 each
        if (threadIdx.x < head)
            dst8[(g0 + threadIdx.x) ^ 3u] = s_stage[stage_swz(threadIdx.x)];
        if (threadIdx.x >= 32 && threadIdx.x - 32 < tail)
            dst8[(w1 * 4 + threadIdx.x - 32) ^ 3u] = s_stage[stage_swz(w1 * 4 - g0 + threadIdx.x - 32)];
    } else {
        for (uint32_t i = threadIdx.x; i < tile_kept; i += kScanThreads)
            dst8[(g0 + i) ^ 3u] = s_stage[stage_swz(i)];
    }
}

// ------------------------------------------------------------------------------------------------
// K1 / K3: Huffman decode of one subsequence
// ------------------------------------------------------------------------------------------------
// Decoder state at a subsequence boundary, packed in 32 bits:
//   [0,6)  bit offset of the first code word past the boundary   [6,12) zig-zag index z
//   [12,16) block-in-MCU index b                                  bit 16: end of stream reached
constexpr uint32_t kStateDone = 1u << 16;

struct DecodeResult {
    uint32_t state;     // exit state
    int32_t blk;        // block index at exit (relative to entry unless restarted, then absolute)
    int32_t restarted;  // crossed at least one restart boundary
    uint32_t err;
};

struct StreamView {
    const uint32_t* words;   // unstuffed stream (big-endian words)
    const uint32_t* bnd;     // restart boundaries as byte positions, then end of stream
    uint32_t nb;             // entries in bnd (>= 1)
    uint32_t total_bits;
    uint32_t bpm, blk_comp, blk_tab;
    uint32_t ri_blocks;      // restart_interval * bpm
    uint32_t pair_off;       // K1: bytes from an AC table of the set in shared memory to its pair table (pair_ac_tables); 0 in K3
    uint32_t total_blocks;
    // K3 counts blocks in the batch's coefficient arena (blk_base = first block of this image; 0 in K1):
    uint32_t blk_base;
    uint32_t write_lo;           // AC coefficients are wanted from here (first MCU row a ROI touches) ...
    uint32_t dc_lo, dc_span;     // ... DC differences in [dc_lo, dc_lo + dc_span): the chain K3b integrates, up to the last wanted row
    uint32_t bit_base;           // subtracted from bit positions stored in blkpos (0: positions are relative to the image's stream)
};

// Coefficient block assembly for K3: the 64 int16 of the block being decoded are emitted as
// 16-byte groups (8 coefficients in zig-zag order) straight from registers; groups that are all zero
// -- most of them at photographic qualities -- are not written at all, a one-byte mask per block says
// which groups exist.  No clearing pass, no partial-sector read-modify-write in L2.
struct BlockWriter {
    uint32_t r0, r1, r2, r3; // the eight coefficients of the current group, zig-zag order, two per register
    uint32_t g;              // current group, 0..7
    uint32_t mask;           // groups of this block that hold a non-zero coefficient (and were stored)
    __device__ __forceinline__ void reset()
    {
        r0 = r1 = r2 = r3 = 0;
        g = 0;
        mask = 0;
    }
    // next block: the group registers are zero already (store_group clears what it stores, and a block
    // that is not wanted never put anything)
    __device__ __forceinline__ void next_block()
    {
        g = 0;
        mask = 0;
    }
    // coef: the batch's coefficient arena, blk: block index in it
    __device__ __forceinline__ void store_group(int16_t* __restrict__ coef, uint32_t blk)
    {
        if ((r0 | r1) | (r2 | r3)) {
            reinterpret_cast<uint4*>(coef)[blk * 8u + g] = make_uint4(r0, r1, r2, r3);
            mask |= 1u << g;
            r0 = r1 = r2 = r3 = 0;
        }
    }
    __device__ __forceinline__ void put(int16_t* __restrict__ coef, uint32_t blk, uint32_t pos, int32_t val)
    {
        const uint32_t ng = pos >> 3;
        if (ng != g) { // positions only grow inside a block: the previous group is complete
            store_group(coef, blk);
            g = ng;
        }
        // 16-bit value to halfword (pos & 7) of r0..r3: a 64-bit shift by the position inside the half, then
        // predicated ORs into the lower or the upper pair
        const uint64_t v = uint64_t(uint32_t(val) & 0xFFFFu) << (16u * (pos & 3u));
        asm("{\n\t.reg .pred q;\n\t"
            "setp.ne.u32 q, %4, 0;\n\t"
            "@!q or.b32 %0, %0, %5;\n\t"
            "@!q or.b32 %1, %1, %6;\n\t"
            "@q or.b32 %2, %2, %5;\n\t"
            "@q or.b32 %3, %3, %6;\n\t}"
            : "+r"(r0), "+r"(r1), "+r"(r2), "+r"(r3)
            : "r"(pos & 4u), "r"(uint32_t(v)), "r"(uint32_t(v >> 32)));
    }
    // end of block: last group out, then the block record = presence mask (all-zero groups are never
    // written; K4a substitutes zeros for them) | DC difference << 16 (K3b turns it into the DC value)
    __device__ __forceinline__ void finish(int16_t* __restrict__ coef, uint32_t* __restrict__ rec, uint32_t blk, int32_t dc_diff)
    {
        store_group(coef, blk);
        rec[blk] = mask | (uint32_t(dc_diff) << 16);
    }
};

// Per-block-in-MCU constants, one 16-byte entry per block index, built once per CTA in shared memory.
// The addresses are 32-bit shared-window addresses, so the decode loop forms every table access with
// one add:
//   .x = address of the block's DC table      .y = address of its AC table
//   .z = address of the NEXT block's entry    .w = component | block index << 8 | component consumed << 24
struct BlockTable {
    uint4 e[kMaxBlocksPerMcu];
};
__device__ __forceinline__ uint32_t smem_addr(const void* p) { return uint32_t(__cvta_generic_to_shared(p)); }
__device__ __forceinline__ uint32_t lds_u16(uint32_t a)
{
    uint16_t v;
    asm("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(a));
    return v;
}
__device__ __forceinline__ uint4 lds_v4(uint32_t a)
{
    uint4 v;
    asm("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a));
    return v;
}
__device__ __forceinline__ void build_block_table(BlockTable* bt, const HuffSetDev* s_hs, const ImageDev& im)
{
    if (threadIdx.x < im.bpm) {
        const uint32_t b = threadIdx.x;
        const uint32_t comp = (im.blk_comp >> (2 * b)) & 3u;
        const uint32_t dc = (im.blk_tab >> b) & 1u;
        const uint32_t ac = 2u + ((im.blk_tab >> (16 + b)) & 1u);
        const uint32_t next = (b + 1u == im.bpm) ? 0u : b + 1u;
        const uint32_t used = (comp == 0u || im.ncomp_needed >= 3) ? 1u : 0u; // Y output consumes luma only
        bt->e[b] = make_uint4(smem_addr(&s_hs->t[dc]), smem_addr(&s_hs->t[ac]), smem_addr(&bt->e[next]), comp | (b << 8) | (used << 24));
    }
}

// Bit reader of the decode loop: the 64 bits at the read position in (whi, wlo), the following word
// prefetched in wnx, wp -> the word held in whi.  A symbol advances the position by at most 31 bits,
// i.e. across at most one word boundary; the refill is predicated, not branched: a branch here is taken
// by some lane of the warp in nearly every iteration and then costs every lane its issue slots.
// (Measured and dropped: an L2 prefetch ahead of the refill, and a subsequence-interleaved stream layout
// that makes a warp's refills share cache lines -- neither shortens the kernels, profiles/README.md.)
struct BitReader {
    uint32_t whi, wlo, wnx;
    const uint32_t* wp;
    __device__ __forceinline__ void seek(const uint32_t* words, uint32_t p)
    {
        wp = words + (p >> 5);
        whi = __ldg(wp);
        wlo = __ldg(wp + 1);
        wnx = __ldg(wp + 2);
    }
    __device__ __forceinline__ uint32_t window(uint32_t p) const { return __funnelshift_l(wlo, whi, p & 31u); }
    // p_old -> p_new crossed a word boundary?  then shift the words up and fetch the next one
    __device__ __forceinline__ void advance(uint32_t p_old, uint32_t p_new)
    {
        asm("{\n\t.reg .pred q;\n\t"
            "setp.gt.u32 q, %4, 31;\n\t"
            "@q mov.b32 %0, %1;\n\t"
            "@q mov.b32 %1, %2;\n\t"
            "@q add.u64 %3, %3, 4;\n\t"
            "@q ld.global.nc.u32 %2, [%3+8];\n\t}"
            : "+r"(whi), "+r"(wlo), "+r"(wnx), "+l"(wp)
            : "r"(p_old ^ p_new));
    }
};

// Decodes the code words that start in [start bit given by in_state, end_bit).
//   kWrite == false (K1): only the decoder state and the block count are tracked -- no symbol values.
//   kWrite == true  (K3): the thread OWNS every block that starts inside its range: a block already
//     under way at entry is skipped (its owner, the previous thread, finishes it), and the block
//     under way at end_bit is decoded to its end.  AC coefficients go out as 16-byte groups, the DC
//     DIFFERENCE goes into the block record; K3b integrates the differences afterwards, so neither
//     kernel carries DC predictors.
// The per-symbol path is straight-line: table address kept in a register, word refill and (K1) the
// switch to the next block's tables predicated.  Only the rare events branch (restart boundary, code
// longer than the first-level table, K3's group / block stores).
template <bool kWrite>
__device__ __forceinline__ void decode_subsequence(const StreamView& sv, const HuffSetDev* __restrict__ hs, const BlockTable* __restrict__ bt,
    uint32_t base_bit, uint32_t end_bit, uint32_t in_state, int32_t blk, int16_t* __restrict__ coef, uint32_t* __restrict__ blkrec,
    DecodeResult& out)
{
    out.err = 0;
    out.restarted = 0;
    if (in_state & kStateDone) {
        out.state = kStateDone;
        out.blk = blk;
        return;
    }
    uint32_t p = base_bit + (in_state & 63u);
    uint32_t z = (in_state >> 6) & 63u;
    // first boundary strictly after p
    uint32_t r = 0;
    if (sv.nb > 1) {
        uint32_t lo = 0, hi = sv.nb - 1; // bnd[nb-1]*8 == total_bits > p for every live thread
        while (lo < hi) {
            const uint32_t mid = (lo + hi) >> 1;
            if (__ldg(sv.bnd + mid) * 8u > p)
                hi = mid;
            else
                lo = mid + 1;
        }
        r = lo;
    }
    uint32_t next_bnd = (r == sv.nb - 1) ? sv.total_bits : __ldg(sv.bnd + r) * 8u;
    // one comparison per symbol covers both "subsequence finished" and "restart boundary close"
    uint32_t limit = min(end_bit, max(next_bnd, 7u) - 7u);
    const uint32_t s_hs = smem_addr(hs), s_bt = smem_addr(bt);
    BitReader br;
    br.seek(sv.words, p);
    uint4 bi = lds_v4(s_bt + ((in_state >> 12) & 15u) * 16u); // tables / component of the block under way
    uint32_t taddr = (z == 0u) ? bi.x : bi.y;                 // table of the next symbol
    bool owned = (z == 0u); // kWrite: is the block under way mine (it started inside my range)?
    // ... and is it wanted at all: its DC difference (component needed, not past the ROI's MCU rows), its AC
    // coefficients (inside those rows)?  Both imply blk < total_blocks.
    auto want_dc = [&](int32_t k, uint32_t w) { return uint32_t(k) - sv.dc_lo < sv.dc_span && (w & (1u << 24)); };
    bool w_dc = kWrite && owned && want_dc(blk, bi.w);
    bool w_ac = w_dc && uint32_t(blk) >= sv.write_lo;
    int32_t dc_diff = 0;
    BlockWriter bw;
    bw.reset();
    bool done = false;
    // The 32 bits at the read position are carried around the loop (computed at the end of an iteration for
    // the next one): the word refill issued in iteration i is then first needed by the register moves of
    // iteration i+1, a whole iteration later, instead of at the loop head.
    uint32_t win = br.window(p);
    for (;;) {
        if (p >= limit) {
            // End of a restart interval (or of the stream): either the decoder already ran across the
            // boundary (only an unsynchronised one can), or fewer than 8 bits remain and they are all
            // 1-bits, i.e. the byte-alignment padding T.81 F.1.2.3 prescribes -- no complete code word
            // fits there, because the all-ones code is never assigned.
            bool at_boundary = false;
            if (p + 7u >= next_bnd) {
                const uint32_t rem = next_bnd - p; // meaningful only when p < next_bnd
                at_boundary = p >= next_bnd || (win >> (32u - rem)) == ((1u << rem) - 1u);
            }
            if (at_boundary) {
                if (r == sv.nb - 1) {
                    done = true;
                    break;
                }
                p = next_bnd;
                ++r;
                next_bnd = (r == sv.nb - 1) ? sv.total_bits : __ldg(sv.bnd + r) * 8u;
                limit = min(end_bit, max(next_bnd, 7u) - 7u);
                if (kWrite && (z != 0u || ((bi.w >> 8) & 15u) != 0u || uint32_t(blk) != sv.blk_base + r * sv.ri_blocks))
                    out.err |= kFlagOverrun;
                blk = int32_t(sv.blk_base + r * sv.ri_blocks);
                out.restarted = 1;
                z = 0;
                owned = true;
                bw.reset();
                bi = lds_v4(s_bt);
                taddr = bi.x;
                w_dc = kWrite && want_dc(blk, bi.w);
                w_ac = w_dc && uint32_t(blk) >= sv.write_lo;
                br.seek(sv.words, p);
                win = br.window(p);
                continue;
            }
            if (p >= end_bit && (!kWrite || z == 0u || !owned))
                break;
            // otherwise: real code words within 7 bits of a boundary, or K3 finishing the block it
            // owns past its range -- keep taking this path until one of the exits above fires
        }
        const bool is_dc = (z == 0u);
        // K1 has a second first-level table per AC table whose entries cover up to B200_K1_CHAIN symbols where they fit the
        // lookup window (pair_ac_tables; `pair_off` bytes behind the table set).  Such an entry may be taken only if its last
        // code word still starts before `limit` and none but the last symbol can end the block (a run is <= 16).
        const bool pair = !kWrite && !is_dc && p + uint32_t(kLutBits) <= limit && z <= 63u - 16u * (B200_K1_CHAIN - 1);
        uint32_t e = lds_u16(taddr + (pair ? sv.pair_off : 0u) + ((win >> (31 - kLutBits)) & ((2u << kLutBits) - 2u)));
        if ((e & 31u) == 0u) {
            // code longer than the first-level table: second-level lookup on the 16-bit window
            const HuffTableDev* tab = reinterpret_cast<const HuffTableDev*>(reinterpret_cast<const uint8_t*>(hs) + (taddr - s_hs));
            const uint32_t i2 = (win >> 16) - tab->first16; // wraps to a huge value below first16
            e = (i2 < tab->n2) ? tab->lut2[i2] : 0u;
            if (tab->n2 == 0u && (win >> 16) >= tab->first16) {
                // table with too many long codes for the second level: canonical search (T.81 F.2.2.3)
#pragma unroll 1
                for (int l = kLutBits + 1; l <= 16; ++l) {
                    const int32_t code = int32_t(win >> (32 - l));
                    if (code <= tab->maxcode[l]) {
                        e = tab->sym[(tab->valoff[l] + code) & 255];
                        break;
                    }
                }
            }
            if ((e & 31u) == 0u) {
                e = 16u | (1u << 9); // unassigned code: consumes 16 bits (only an unsynchronised decoder gets here on a valid stream)
                out.err |= kFlagBadHuffman;
            }
        }
        const uint32_t tot = e & 31u, size = (e >> 5) & 15u, skip = e >> 9;
        br.advance(p, p + tot);
        p += tot;
        if (kWrite) {
            // the value the symbol carries (T.81 F.2.2.1 EXTEND)
            const uint32_t rest = win << (tot - size);
            const uint32_t mag = __funnelshift_l(rest, 0u, size); // top `size` bits, 0 when size == 0
            const int32_t val = (rest >> 31) ? int32_t(mag) : int32_t(mag) - int32_t((1u << size) - 1u);
            dc_diff = is_dc ? val : dc_diff;
            if (w_ac && !is_dc && size != 0u) {
                const uint32_t pos = z + skip - 1u; // zig-zag position of the coefficient this symbol carries
                out.err |= (pos > 63u) ? uint32_t(kFlagBadHuffman) : 0u;
                bw.put(coef, uint32_t(blk), pos & 63u, val);
            }
            z += skip;
            if (z > 63u) {
                if (w_dc)
                    bw.finish(coef, blkrec, uint32_t(blk), dc_diff);
                bw.next_block();
                owned = true;
                z = 0;
                ++blk;
                bi = lds_v4(bi.z);
                w_dc = want_dc(blk, bi.w);
                w_ac = w_dc && uint32_t(blk) >= sv.write_lo;
            }
        } else {
            z += skip;
            // end of block: next block's entry, predicated (some lane ends a block in most iterations)
            asm("{\n\t.reg .pred q;\n\t"
                "setp.gt.u32 q, %4, 63;\n\t"
                "@q ld.shared.v4.u32 {%0,%1,%2,%3}, [%2];\n\t}"
                : "+r"(bi.x), "+r"(bi.y), "+r"(bi.z), "+r"(bi.w)
                : "r"(z));
            blk += (z > 63u) ? 1 : 0;
            z = (z > 63u) ? 0u : z;
        }
        taddr = (z == 0u) ? bi.x : bi.y;
        win = br.window(p);
    }
    if (done) {
        out.state = kStateDone;
        if (kWrite && (z != 0u || uint32_t(blk) != sv.blk_base + sv.total_blocks))
            out.err |= kFlagOverrun;
    } else {
        out.state = ((p - end_bit) & 63u) | (z << 6) | (((bi.w >> 8) & 15u) << 12);
    }
    out.blk = blk;
}

__device__ __forceinline__ void load_huff_set(HuffSetDev* s_hs, const HuffSetDev* g_hs)
{
    const uint4* src = reinterpret_cast<const uint4*>(g_hs);
    uint4* dst = reinterpret_cast<uint4*>(s_hs);
    for (uint32_t i = threadIdx.x; i < sizeof(HuffSetDev) / 16; i += blockDim.x)
        dst[i] = __ldg(src + i);
}

// K1 only: a second first-level table per AC table whose entries cover up to B200_K1_CHAIN symbols wherever the code words
// after the first lie entirely inside the lookup window: advance = all their bits, zig-zag advance = all their runs (the
// last may be an EOB).
// K1 tracks neither values nor magnitude sizes, so a pair is just a longer symbol to it and the loop runs fewer iterations.
// pair table k sits k * sizeof(HuffTableDev) bytes into `s_pair`, i.e. at one constant offset from AC table k of the set.
__device__ __forceinline__ void pair_ac_tables(const HuffSetDev* s_hs, uint8_t* s_pair)
{
    constexpr uint32_t N = 1u << kLutBits;
    for (uint32_t idx = threadIdx.x; idx < 2u * N; idx += blockDim.x) {
        const HuffTableDev& tab = s_hs->t[2 + idx / N];
        const uint32_t i = idx % N;
        const uint32_t e1 = tab.lut[i];
        uint32_t out = e1;
        uint32_t tot = e1 & 31u, skip = e1 >> 9, last = skip;
        if (tot != 0u) {
            for (int n = 1; n < B200_K1_CHAIN && tot < uint32_t(kLutBits) && last != 64u; ++n) {
                const uint32_t e2 = tab.lut[(i << tot) & (N - 1u)]; // the window behind the symbols so far, unknown bits as zeros
                const uint32_t tot2 = e2 & 31u, len2 = tot2 - ((e2 >> 5) & 15u);
                if (tot2 == 0u || len2 > uint32_t(kLutBits) - tot || tot + tot2 > 31u)
                    break;
                tot += tot2;
                last = e2 >> 9;
                skip += last;
                out = tot | (skip << 9);
            }
        }
        reinterpret_cast<uint16_t*>(s_pair + (idx / N) * sizeof(HuffTableDev))[i] = uint16_t(out);
    }
}

__device__ __forceinline__ StreamView make_view(const BatchDev& bd, const ImageDev& im, const ImageState& st)
{
    StreamView sv;
    sv.words = bd.cmp + im.cmp_off / 4;
    sv.bnd = bd.bnd + im.bnd_off;
    sv.nb = st.n_rst + 1;
    sv.total_bits = st.cmp_len * 8u;
    sv.bpm = im.bpm;
    sv.blk_comp = im.blk_comp;
    sv.blk_tab = im.blk_tab;
    sv.ri_blocks = im.restart_interval * im.bpm;
    sv.pair_off = 0;
    sv.total_blocks = im.total_blocks;
    sv.blk_base = 0;
    sv.write_lo = im.write_blk_lo;
    sv.dc_lo = im.dc_blk_lo;
    sv.dc_span = im.write_blk_hi - im.dc_blk_lo;
    sv.bit_base = 0;
    return sv;
}

// K1.  One CTA = kSeqLogical - seq_warm consecutive subsequences of one image (+ seq_warm leading threads whose
// only job is to make the CTA's entry state right with high probability).
//   round 0  every thread decodes the last quarter subsequence BEFORE its own from the bare guess
//            "a code word starts here, at a block start": JPEG Huffman streams self-synchronise
//            within a few dozen symbols, so this usually lands on the true entry state;
//   round 1  every thread decodes its subsequence from that entry state (exit state, block count, DC sums);
//   round 2+ "my entry state := my predecessor's exit state; re-decode if it changed", to a fixed
//            point; the few threads that changed are compacted into the leading warps so the sparse
//            rounds do not burn issue slots of 8 mostly idle warps.
// fix mode (images K2 found dirty): the first thread takes the stored exit state of the previous
// sequence as entry state and only the fixed-point rounds run.
struct SyncShared {
    uint32_t in[kSeqLogical];
    uint32_t out[kSeqLogical];
    uint32_t info[kSeqLogical];     // block count (or absolute index | 1 << 31 after a restart)
    uint32_t queue[2][kSeqLogical]; // work lists of the current and the next fixed-point round
    uint8_t fresh[kSeqLogical];     // (re)computed by this CTA -> has to be written back
    uint32_t count[2];
};
constexpr size_t kSyncPairOffset = (sizeof(HuffSetDev) + sizeof(BlockTable) + sizeof(SyncShared) + 15) / 16 * 16;

__device__ __forceinline__ uint32_t pack_info(const DecodeResult& r) { return uint32_t(r.blk) | (r.restarted ? 0x80000000u : 0u); }

// A CTA of kSeqThreads threads owns kSeqLogical = kSeqMult * kSeqThreads consecutive subsequences
// ("logical threads" l = m * kSeqThreads + t): the full pass runs kSeqMult times back to back, and the
// repair rounds batch the work of the whole range, so they run with (nearly) full warps instead of
// idling most of the CTA at the round barriers.
template <int kWarm>
__global__ void __launch_bounds__(kSeqThreads) huff_sync_kernel(BatchDev bd, int mode)
{
    extern __shared__ uint4 s_raw[];
    HuffSetDev* s_hs = reinterpret_cast<HuffSetDev*>(s_raw);
    BlockTable* s_bt = reinterpret_cast<BlockTable*>(reinterpret_cast<uint8_t*>(s_raw) + sizeof(HuffSetDev));
    SyncShared& sh = *reinterpret_cast<SyncShared*>(reinterpret_cast<uint8_t*>(s_raw) + sizeof(HuffSetDev) + sizeof(BlockTable));
    uint8_t* s_pair = reinterpret_cast<uint8_t*>(s_raw) + kSyncPairOffset;
    const bool fix_mode = (mode & 1) != 0;
    const bool no_warm = (mode & 2) != 0; // test hook: bare guesses, no warm-up threads, no probe
    const uint32_t S = bd.sub_bits;
    const uint32_t seq = blockIdx.x;
    const uint32_t img = bd.seq_img[seq];
    const ImageDev& im = bd.img[img];
    const ImageState st = bd.state[img];
    if (st.n_sub == 0u || (fix_mode && !(st.flags & kFlagDirty)))
        return;
    constexpr int warm = kWarm; // leading subsequences of the CTA that only warm up its entry state (BatchDev::seq_warm)
    const uint32_t first_k = (seq - im.seq_off) * uint32_t(kSeqLogical - warm);
    if (first_k >= st.n_sub)
        return;
    load_huff_set(s_hs, bd.huff + im.huff_set);
    build_block_table(s_bt, s_hs, im);
    StreamView sv = make_view(bd, im, st);
    sv.pair_off = uint32_t(kSyncPairOffset - 2 * sizeof(HuffTableDev)); // AC table k of the set is t[2 + k]
    const int t = threadIdx.x;
    const int k0 = int(first_k) - warm; // subsequence of logical thread 0 (may be negative)
    const bool skip_warm = fix_mode || no_warm;
    auto is_active = [&](int l) { return k0 + l >= 0 && k0 + l < int(st.n_sub) && (!skip_warm || l >= warm); };
    auto has_pred = [&](int l) { return is_active(l) && l > 0 && k0 + l > 0 && (!skip_warm || l > warm); };
    if (t == 0)
        sh.count[0] = sh.count[1] = 0;
    for (int l = t; l < kSeqLogical; l += kSeqThreads)
        sh.fresh[l] = 0;
    __syncthreads(); // tables loaded, counters cleared
    pair_ac_tables(s_hs, s_pair);
    __syncthreads();
    DecodeResult res;
    if (fix_mode) {
        for (int l = t; l < kSeqLogical; l += kSeqThreads) {
            const bool active = is_active(l);
            const uint32_t g = im.sub_off + uint32_t(k0 + l);
            uint32_t in = 0;
            sh.out[l] = active ? bd.sub_out[g] : kStateDone;
            if (active) {
                in = bd.sub_in[g];
                if (l == warm) {
                    const uint32_t want = (k0 + l > 0) ? bd.sub_out[g - 1] : 0u;
                    if (want != in) {
                        in = want;
                        sh.fresh[l] = 1;
                        sh.queue[0][atomicAdd(&sh.count[0], 1u)] = uint32_t(l);
                    }
                }
            }
            sh.in[l] = in;
        }
    } else {
        for (int l = t; l < kSeqLogical; l += kSeqThreads) {
            const bool active = is_active(l);
            const int k = k0 + l;
            uint32_t in = 0; // k == 0: the true start state; otherwise the bare guess
            if (active && k > 0 && !no_warm) {
                // round 0: probe the tail of the previous subsequence
                const uint32_t pb = uint32_t(k) * S - S / 4;
                decode_subsequence<false>(sv, s_hs, s_bt, pb, uint32_t(k) * S, 0u, 0, nullptr, nullptr, res);
                in = res.state;
            }
            // round 1: everybody
            res.state = kStateDone;
            if (active) {
                decode_subsequence<false>(sv, s_hs, s_bt, uint32_t(k) * S, uint32_t(k + 1) * S, in, 0, nullptr, nullptr, res);
                sh.info[l] = pack_info(res);
                sh.fresh[l] = 1;
            }
            sh.in[l] = in;
            sh.out[l] = res.state;
        }
        __syncthreads();
        for (int l = t; l < kSeqLogical; l += kSeqThreads) {
            if (has_pred(l) && sh.out[l - 1] != sh.in[l]) {
                sh.in[l] = sh.out[l - 1];
                sh.queue[0][atomicAdd(&sh.count[0], 1u)] = uint32_t(l);
            }
        }
    }
    // fixed-point rounds over the compacted work list.  Invariant: in[l] == out[l-1] for every l,
    // and out[l] == decode(in[l]) for every l that is not queued.
    uint32_t dbg_rounds = 0, dbg_items = 0;
    for (uint32_t cur = 0;; cur ^= 1u) {
        __syncthreads();
        const uint32_t n = sh.count[cur];
        if (n == 0u)
            break;
        ++dbg_rounds;
        dbg_items += n;
        for (uint32_t i = uint32_t(t); i < n; i += kSeqThreads) {
            const uint32_t item = sh.queue[cur][i];
            const uint32_t kk = uint32_t(k0 + int(item));
            decode_subsequence<false>(sv, s_hs, s_bt, kk * S, (kk + 1) * S, sh.in[item], 0, nullptr, nullptr, res);
            sh.out[item] = res.state; // nobody reads out[] before the barrier below
            sh.info[item] = pack_info(res);
        }
        if (t == 0)
            sh.count[cur ^ 1u] = 0; // the other list was drained in the previous round
        __syncthreads();
        for (uint32_t i = uint32_t(t); i < n; i += kSeqThreads) {
            const uint32_t item = sh.queue[cur][i];
            const uint32_t succ = item + 1;
            const uint32_t state = sh.out[item];
            if (succ < uint32_t(kSeqLogical) && k0 + int(succ) < int(st.n_sub) && sh.in[succ] != state) {
                sh.in[succ] = state;
                sh.fresh[succ] = 1;
                sh.queue[cur ^ 1u][atomicAdd(&sh.count[cur ^ 1u], 1u)] = succ;
            }
        }
        if (t == 0)
            sh.count[cur] = 0;
    }
    if (t == 0 && bd.sync_stats) { // [0] CTAs, [1] repair rounds, [2] repaired subsequences, [3] max rounds in one CTA
        atomicAdd(bd.sync_stats + 0, 1u);
        atomicAdd(bd.sync_stats + 1, dbg_rounds);
        atomicAdd(bd.sync_stats + 2, dbg_items);
        atomicMax(bd.sync_stats + 3, dbg_rounds);
    }
    // results of the real (non warm-up) subsequences
    for (int l = t; l < kSeqLogical; l += kSeqThreads) {
        if (is_active(l) && l >= warm && sh.fresh[l]) {
            const uint32_t g = im.sub_off + uint32_t(k0 + l);
            bd.sub_in[g] = sh.in[l];
            bd.sub_out[g] = sh.out[l];
            bd.sub_info[g] = sh.info[l];
        }
    }
}

// K2.  One CTA per image: (a) every subsequence's entry state must equal its predecessor's exit
// state (else the image is dirty and goes through the fix-up path), (b) segmented exclusive scan
// of the block counts where a restart boundary resets the running value to the absolute index.
struct ScanElem {
    int32_t f, blk;
};
__device__ __forceinline__ ScanElem scan_combine(const ScanElem& a, const ScanElem& b)
{
    ScanElem r;
    r.f = a.f | b.f;
    r.blk = b.f ? b.blk : a.blk + b.blk;
    return r;
}
__device__ __forceinline__ ScanElem scan_shfl_up(const ScanElem& a, int d)
{
    ScanElem r;
    r.f = __shfl_up_sync(0xffffffffu, a.f, d);
    r.blk = __shfl_up_sync(0xffffffffu, a.blk, d);
    return r;
}

__global__ void __launch_bounds__(256) verify_scan_kernel(BatchDev bd, int fix_mode)
{
    __shared__ ScanElem s_warp[8];
    __shared__ ScanElem s_carry;
    __shared__ uint32_t s_dirty;
    const uint32_t img = blockIdx.x;
    const ImageDev& im = bd.img[img];
    const ImageState st = bd.state[img];
    if (st.n_sub == 0u || (fix_mode && !(st.flags & kFlagDirty)))
        return;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) {
        s_carry = ScanElem{0, 0};
        s_dirty = 0;
    }
    __syncthreads();
    for (uint32_t c = 0; c < st.n_sub; c += 256) {
        const uint32_t k = c + threadIdx.x;
        ScanElem e{0, 0};
        if (k < st.n_sub) {
            const uint32_t g = im.sub_off + k;
            const uint32_t in = bd.sub_in[g];
            const uint32_t want = k ? bd.sub_out[g - 1] : 0u;
            if (in != want)
                atomicOr(&s_dirty, 1u);
            const uint32_t info = bd.sub_info[g];
            e.f = int32_t(info >> 31);
            e.blk = int32_t(info & 0x7FFFFFFFu);
        }
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const ScanElem o = scan_shfl_up(e, d);
            if (lane >= d)
                e = scan_combine(o, e);
        }
        if (lane == 31)
            s_warp[warp] = e;
        __syncthreads();
        ScanElem pre = s_carry;
        for (int i = 0; i < warp; ++i)
            pre = scan_combine(pre, s_warp[i]);
        // exclusive value = pre (+) inclusive-of-previous-lane
        ScanElem prev = scan_shfl_up(e, 1);
        ScanElem ex = (lane == 0) ? pre : scan_combine(pre, prev);
        if (k < st.n_sub)
            bd.sub_start[im.sub_off + k] = uint32_t(ex.blk);
        __syncthreads();
        if (threadIdx.x == 255)
            s_carry = scan_combine(pre, e);
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        uint32_t f = st.flags & ~kFlagDirty;
        if (s_dirty)
            f |= kFlagDirty;
        if (fix_mode)
            f |= kFlagRedo; // its coefficients and pixels have to be produced again once it is clean
        bd.state[img].flags = f;
    }
}

// K3.  Same CTA shape as K1; every thread re-decodes its subsequence from the verified entry
// state, now knowing its absolute block index, and writes the coefficient blocks it owns (zig-zag
// order, MCU decode order) as whole 16-byte groups plus one record per block (group mask, DC difference).
__global__ void __launch_bounds__(kSeqThreads, 5) huff_write_kernel(BatchDev bd, int redo_only)
{
    extern __shared__ uint4 s_raw[];
    HuffSetDev* s_hs = reinterpret_cast<HuffSetDev*>(s_raw);
    BlockTable* s_bt = reinterpret_cast<BlockTable*>(reinterpret_cast<uint8_t*>(s_raw) + sizeof(HuffSetDev));
    const uint32_t cta = blockIdx.x;
    const uint32_t img = bd.wseq_img[cta];
    const ImageDev& im = bd.img[img];
    const ImageState st = bd.state[img];
    if (st.n_sub == 0u || (st.flags & kFlagDirty) || (redo_only && !(st.flags & kFlagRedo)))
        return;
    const uint32_t first_k = (cta - im.wseq_off) * kSeqThreads;
    if (first_k >= st.n_sub)
        return;
    load_huff_set(s_hs, bd.huff + im.huff_set);
    build_block_table(s_bt, s_hs, im);
    __syncthreads();
    const uint32_t k = first_k + threadIdx.x;
    if (k >= st.n_sub)
        return;
    StreamView sv = make_view(bd, im, st);
    sv.blk_base = uint32_t(im.coef_off / 64); // blocks are counted in the batch's arena from here on
    sv.write_lo += sv.blk_base;
    sv.dc_lo += sv.blk_base;
    const uint32_t g = im.sub_off + k;
    const uint32_t in = bd.sub_in[g];
    const uint32_t start = bd.sub_start[g];
    DecodeResult res;
    uint32_t err = 0;
    // ROI: nothing to write if every block this thread can own lies past the wanted MCU rows or before the
    // restart interval they start in (DC differences are needed from there on).  (The end-of-stream
    // consistency check then falls to whoever decodes the tail; a ROI that excludes the last MCU row skips
    // it, like libjpeg's jpeg_skip_scanlines does.)
    const uint32_t next_start = (k + 1 < st.n_sub) ? bd.sub_start[g + 1] : im.total_blocks;
    if (start >= im.write_blk_hi || next_start < im.dc_blk_lo)
        return;
    if (!(in & kStateDone) && (start % im.bpm) != ((in >> 12) & 15u)) {
        err |= kFlagPhase;
    } else {
        decode_subsequence<true>(sv, s_hs, s_bt, k * bd.sub_bits, (k + 1) * bd.sub_bits, in, int32_t(sv.blk_base + start), bd.coef, bd.blkrec, res);
        err |= res.err;
    }
    if (err)
        atomicOr(&bd.state[img].flags, err);
}

// K3, entry-stream form.  Every thread decodes exactly the code words that start in its subsequence (as K1 does: no block
// ownership, nothing is decoded twice) and writes one 32-bit entry per loop iteration, whatever the iteration met; the
// four entries of a chunk are four registers (the chunk loop is unrolled), and the lanes of a warp store their chunks side
// by side -- 512 contiguous bytes per warp and chunk row.  Nothing depends on the symbol but the entry's value: no group
// stores, no per-block flush, the end of a block is a predicated table switch as in K1.  The one scattered store left is
// the block record at a DC symbol (entry slot + DC difference, and the bit position in blkpos).
// An entry is made for its consumer (sym_fetch_block): value in the low half; in the high half the byte offset at which
// the IDCT kernel's scatter column holds that zig-zag position, plus two flags.  An iteration that carries no coefficient
// (DC symbol, sixteen zeros, restart hop) leaves 0, which the consumer may store like any other entry: offset 0 is the
// unused DC position.
constexpr uint32_t kEntNop = 0u, kEntLast = 0x40000000u /* the block ends with this entry */, kEntEnd = 0x80000000u /* the subsequence's entries end */;
__device__ __forceinline__ uint32_t ent_offset(uint32_t pos) { return ((pos >> 1) * kScatterStride) | ((pos & 1u) << 1); }

__device__ __forceinline__ uint32_t emit_subsequence(const StreamView& sv, const HuffSetDev* __restrict__ hs, const BlockTable* __restrict__ bt,
    uint32_t base_bit, uint32_t end_bit, uint32_t in_state, int32_t blk, uint32_t* __restrict__ blkrec, uint32_t* __restrict__ blkpos,
    uint4* __restrict__ out, uint32_t max_chunks)
{
    uint32_t err = 0;
    if (in_state & kStateDone) { // past the end of the stream: an empty entry list
        out[0] = make_uint4(kEntEnd, kEntNop, kEntNop, kEntNop);
        return 0;
    }
    uint32_t p = base_bit + (in_state & 63u);
    uint32_t z = (in_state >> 6) & 63u;
    uint32_t r = 0; // first boundary strictly after p
    if (sv.nb > 1) {
        uint32_t lo = 0, hi = sv.nb - 1;
        while (lo < hi) {
            const uint32_t mid = (lo + hi) >> 1;
            if (__ldg(sv.bnd + mid) * 8u > p)
                hi = mid;
            else
                lo = mid + 1;
        }
        r = lo;
    }
    uint32_t next_bnd = (r == sv.nb - 1) ? sv.total_bits : __ldg(sv.bnd + r) * 8u;
    uint32_t limit = min(end_bit, max(next_bnd, 7u) - 7u);
    const uint32_t s_hs = smem_addr(hs), s_bt = smem_addr(bt);
    BitReader br;
    br.seek(sv.words, p);
    uint4 bi = lds_v4(s_bt + ((in_state >> 12) & 15u) * 16u);
    uint32_t taddr = (z == 0u) ? bi.x : bi.y;
    uint32_t win = br.window(p);
    bool running = true;
#pragma unroll 1
    for (uint32_t c = 0;; ++c) {
        uint32_t q[4] = {kEntNop, kEntNop, kEntNop, kEntNop};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            if (!running)
                continue;
            uint32_t ent = kEntNop;
            bool symbol = true;
            if (p >= limit) {
                // End of a restart interval (or of the stream): the decoder ran across the boundary, or fewer than 8 bits
                // remain and they are all 1-bits (the byte-alignment padding of T.81 F.1.2.3), see decode_subsequence.
                bool at_boundary = false;
                if (p + 7u >= next_bnd) {
                    const uint32_t rem = next_bnd - p;
                    at_boundary = p >= next_bnd || (win >> (32u - rem)) == ((1u << rem) - 1u);
                }
                if (at_boundary) {
                    symbol = false;
                    if (r == sv.nb - 1) { // end of the stream
                        if (z != 0u || uint32_t(blk) != sv.blk_base + sv.total_blocks)
                            err |= kFlagOverrun;
                        ent = kEntEnd;
                        running = false;
                    } else {
                        p = next_bnd;
                        ++r;
                        next_bnd = (r == sv.nb - 1) ? sv.total_bits : __ldg(sv.bnd + r) * 8u;
                        limit = min(end_bit, max(next_bnd, 7u) - 7u);
                        if (z != 0u || ((bi.w >> 8) & 15u) != 0u || uint32_t(blk) != sv.blk_base + r * sv.ri_blocks)
                            err |= kFlagOverrun;
                        blk = int32_t(sv.blk_base + r * sv.ri_blocks);
                        z = 0;
                        bi = lds_v4(s_bt);
                        taddr = bi.x;
                        br.seek(sv.words, p);
                        win = br.window(p);
                        if (p >= end_bit) { // the interval that starts here belongs to the next subsequence
                            ent = kEntEnd;
                            running = false;
                        }
                    }
                } else if (p >= end_bit) {
                    symbol = false;
                    ent = kEntEnd;
                    running = false;
                }
            }
            if (symbol) {
                const bool is_dc = (z == 0u);
                uint32_t e = lds_u16(taddr + ((win >> (31 - kLutBits)) & ((2u << kLutBits) - 2u)));
                if ((e & 31u) == 0u) {
                    const HuffTableDev* tab = reinterpret_cast<const HuffTableDev*>(reinterpret_cast<const uint8_t*>(hs) + (taddr - s_hs));
                    const uint32_t i2 = (win >> 16) - tab->first16;
                    e = (i2 < tab->n2) ? tab->lut2[i2] : 0u;
                    if (tab->n2 == 0u && (win >> 16) >= tab->first16) {
#pragma unroll 1
                        for (int l = kLutBits + 1; l <= 16; ++l) {
                            const int32_t code = int32_t(win >> (32 - l));
                            if (code <= tab->maxcode[l]) {
                                e = tab->sym[(tab->valoff[l] + code) & 255];
                                break;
                            }
                        }
                    }
                    if ((e & 31u) == 0u) {
                        e = 16u | (1u << 9);
                        err |= kFlagBadHuffman;
                    }
                }
                const uint32_t tot = e & 31u, size = (e >> 5) & 15u, skip = e >> 9;
                const uint32_t p_sym = p;
                br.advance(p, p + tot);
                p += tot;
                const uint32_t rest = win << (tot - size);
                const uint32_t mag = __funnelshift_l(rest, 0u, size);
                const int32_t val = (rest >> 31) ? int32_t(mag) : int32_t(mag) - int32_t((1u << size) - 1u);
                z += skip;
                if (is_dc) {
                    // block record: where the block's entries start (this slot, which itself carries nothing) and its DC difference
                    if (uint32_t(blk) - sv.dc_lo < sv.dc_span && (bi.w & (1u << 24))) {
                        blkrec[blk] = (c * 4u + uint32_t(j)) | (uint32_t(val) << 16);
                        blkpos[blk] = p_sym - sv.bit_base;
                    }
                } else {
                    const uint32_t pos = z - 1u; // zig-zag position of the coefficient a run/size symbol carries
                    err |= (size != 0u && pos > 63u) ? uint32_t(kFlagBadHuffman) : 0u;
                    ent = (size != 0u) ? ((ent_offset(pos & 63u) << 16) | (uint32_t(val) & 0xFFFFu)) : kEntNop;
                    ent |= (z > 63u) ? kEntLast : 0u; // an end-of-block symbol, or the coefficient at position 63
                }
                asm("{\n\t.reg .pred q;\n\t"
                    "setp.gt.u32 q, %4, 63;\n\t"
                    "@q ld.shared.v4.u32 {%0,%1,%2,%3}, [%2];\n\t}"
                    : "+r"(bi.x), "+r"(bi.y), "+r"(bi.z), "+r"(bi.w)
                    : "r"(z));
                blk += (z > 63u) ? 1 : 0;
                z = (z > 63u) ? 0u : z;
                taddr = (z == 0u) ? bi.x : bi.y;
                win = br.window(p);
            }
            q[j] = ent;
        }
        if (c < max_chunks) {
            out[size_t(c) * 32] = make_uint4(q[0], q[1], q[2], q[3]);
        } else { // more entries than a subsequence can legitimately have: damaged data
            err |= kFlagOverrun;
            break;
        }
        if (!running)
            break;
    }
    return err;
}

__global__ void __launch_bounds__(kSeqThreads, 5) huff_emit_kernel(BatchDev bd, int redo_only)
{
    extern __shared__ uint4 s_raw[];
    HuffSetDev* s_hs = reinterpret_cast<HuffSetDev*>(s_raw);
    BlockTable* s_bt = reinterpret_cast<BlockTable*>(reinterpret_cast<uint8_t*>(s_raw) + sizeof(HuffSetDev));
    const uint32_t cta = blockIdx.x;
    const uint32_t img = bd.wseq_img[cta];
    const ImageDev& im = bd.img[img];
    const ImageState st = bd.state[img];
    if (st.n_sub == 0u || (st.flags & kFlagDirty) || (redo_only && !(st.flags & kFlagRedo)))
        return;
    const uint32_t first_k = (cta - im.wseq_off) * kSeqThreads;
    if (first_k >= st.n_sub)
        return;
    load_huff_set(s_hs, bd.huff + im.huff_set);
    build_block_table(s_bt, s_hs, im);
    __syncthreads();
    const uint32_t k = first_k + threadIdx.x;
    if (k >= st.n_sub)
        return;
    StreamView sv = make_view(bd, im, st);
    sv.blk_base = uint32_t(im.coef_off / 64); // blocks are counted in the batch's arena from here on
    sv.dc_lo += sv.blk_base;
    const uint32_t g = im.sub_off + k;
    const uint32_t in = bd.sub_in[g];
    const uint32_t start = bd.sub_start[g];
    // ROI: threads whose blocks all lie past the wanted MCU rows, or before the restart interval they start in, have nothing to
    // contribute (no wanted block continues into them either: the block under way at their entry is `start`)
    const uint32_t next_start = (k + 1 < st.n_sub) ? bd.sub_start[g + 1] : im.total_blocks;
    if (start >= im.write_blk_hi || next_start < im.dc_blk_lo)
        return;
    uint32_t err = 0;
    if (!(in & kStateDone) && (start % im.bpm) != ((in >> 12) & 15u)) {
        err |= kFlagPhase;
    } else {
        uint4* out = bd.sym + im.sym_off + size_t(k >> 5) * bd.sym_chunks * 32 + (k & 31u);
        err |= emit_subsequence(sv, s_hs, s_bt, k * bd.sub_bits, (k + 1) * bd.sub_bits, in, int32_t(sv.blk_base + start), bd.blkrec, bd.blkpos, out,
            bd.sym_chunks);
    }
    if (err)
        atomicOr(&bd.state[img].flags, err);
}

// K3b.  DC differences -> DC values, in place in the block records K3 wrote (T.81 F.2.1.3.1: every DC is coded
// as the difference from the previous block of the same component; restart intervals reset the predictors).
// A CTA takes one chunk of kDcChunk consecutive blocks (MCU decode order) of one image, a thread kDcPer of
// them.  Two launches: (a) per-chunk sums, (b) every chunk combines the sums of the chunks before it in its
// image (one warp, a few loads) and integrates its own blocks.  Arithmetic is int32 with the low 16 bits
// kept, exactly what storing into a 16-bit coefficient does.
struct DcElem {
    int32_t f, d0, d1, d2; // f: a predictor reset lies inside -> the sums are absolute
};
__device__ __forceinline__ DcElem dc_combine(const DcElem& a, const DcElem& b)
{
    DcElem r;
    r.f = a.f | b.f;
    r.d0 = b.f ? b.d0 : a.d0 + b.d0;
    r.d1 = b.f ? b.d1 : a.d1 + b.d1;
    r.d2 = b.f ? b.d2 : a.d2 + b.d2;
    return r;
}
__device__ __forceinline__ DcElem dc_shfl_up(const DcElem& a, int d)
{
    DcElem r;
    r.f = __shfl_up_sync(0xffffffffu, a.f, d);
    r.d0 = __shfl_up_sync(0xffffffffu, a.d0, d);
    r.d1 = __shfl_up_sync(0xffffffffu, a.d1, d);
    r.d2 = __shfl_up_sync(0xffffffffu, a.d2, d);
    return r;
}
__device__ __forceinline__ DcElem dc_warp_scan(DcElem e, int lane)
{
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const DcElem o = dc_shfl_up(e, d);
        if (lane >= d)
            e = dc_combine(o, e);
    }
    return e;
}

// The thread's kDcPer blocks starting at j0: records r[], running per-component sums v[] (inclusive, restarted at
// a predictor reset), the thread's element, and the first i at which the sums became absolute.
struct DcLocal {
    uint32_t r[kDcPer];
    int32_t v[kDcPer];
    DcElem e;
    uint32_t reset_at;
};
__device__ __forceinline__ void dc_local(const uint32_t* __restrict__ rec, uint32_t j0, uint32_t hi, uint32_t bpm, uint32_t ri, uint32_t blk_comp,
    DcLocal& L)
{
    uint32_t bim = j0 % bpm;
    uint32_t rim = ri ? j0 % ri : (j0 == 0u ? 0u : 1u); // position inside the restart interval; without restarts only block 0 resets
    L.e = DcElem{0, 0, 0, 0};
    L.reset_at = kDcPer;
#pragma unroll
    for (int i = 0; i < kDcPer; ++i) {
        const uint32_t j = j0 + i;
        L.r[i] = (j < hi) ? rec[j] : 0u;
        if (rim == 0u && j < hi) {
            L.e = DcElem{1, 0, 0, 0};
            L.reset_at = min(L.reset_at, uint32_t(i));
        }
        const uint32_t c = (blk_comp >> (2u * bim)) & 3u;
        const int32_t diff = int32_t(L.r[i]) >> 16;
        L.e.d0 += (c == 0u) ? diff : 0;
        L.e.d1 += (c == 1u) ? diff : 0;
        L.e.d2 += (c == 2u) ? diff : 0;
        L.v[i] = (c == 0u) ? L.e.d0 : (c == 1u ? L.e.d1 : L.e.d2);
        bim = (bim + 1u == bpm) ? 0u : bim + 1u;
        ++rim;
        if (ri && rim == ri)
            rim = 0u;
    }
}

__global__ void __launch_bounds__(kDcThreads) dc_sum_kernel(BatchDev bd, int redo_only)
{
    __shared__ DcElem s_warp[kDcThreads / 32];
    const uint32_t cta = blockIdx.x;
    const uint32_t img = bd.dc_cta_img[cta];
    const ImageDev& im = bd.img[img];
    const uint32_t flags = bd.state[img].flags;
    if ((flags & (kFlagDirty | kFlagBadMarker)) || (redo_only && !(flags & kFlagRedo)))
        return;
    const uint32_t* rec = bd.blkrec + im.coef_off / 64;
    const uint32_t j0 = im.dc_blk_lo + (cta - im.dc_cta_off) * kDcChunk + threadIdx.x * kDcPer;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    DcLocal L;
    dc_local(rec, j0, im.write_blk_hi, im.bpm, im.restart_interval * im.bpm, im.blk_comp, L);
    const DcElem inc = dc_warp_scan(L.e, lane);
    if (lane == 31)
        s_warp[warp] = inc;
    __syncthreads();
    if (threadIdx.x == 0) {
        DcElem t = s_warp[0];
        for (int i = 1; i < kDcThreads / 32; ++i)
            t = dc_combine(t, s_warp[i]);
        bd.dc_sum[cta] = make_int4(t.f, t.d0, t.d1, t.d2);
    }
}

__global__ void __launch_bounds__(kDcThreads, 2) dc_apply_kernel(BatchDev bd, int redo_only)
{
    __shared__ DcElem s_warp[kDcThreads / 32];
    __shared__ DcElem s_carry;
    const uint32_t cta = blockIdx.x;
    const uint32_t img = bd.dc_cta_img[cta];
    const ImageDev& im = bd.img[img];
    const uint32_t flags = bd.state[img].flags;
    if ((flags & (kFlagDirty | kFlagBadMarker)) || (redo_only && !(flags & kFlagRedo)))
        return;
    uint32_t* rec = bd.blkrec + im.coef_off / 64;
    const uint32_t chunk = cta - im.dc_cta_off;
    const uint32_t j0 = im.dc_blk_lo + chunk * kDcChunk + threadIdx.x * kDcPer;
    const uint32_t hi = im.write_blk_hi, bpm = im.bpm, blk_comp = im.blk_comp;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    // predictors at the start of this chunk: the sums of the chunks before it, combined in order by warp 0
    if (warp == 0) {
        DcElem carry{0, 0, 0, 0};
        for (uint32_t c0 = 0; c0 < chunk; c0 += 32) {
            DcElem e{0, 0, 0, 0};
            if (c0 + lane < chunk) {
                const int4 t = bd.dc_sum[im.dc_cta_off + c0 + lane];
                e = DcElem{t.x, t.y, t.z, t.w};
            }
            e = dc_warp_scan(e, lane);
            DcElem last;
            last.f = __shfl_sync(0xffffffffu, e.f, 31);
            last.d0 = __shfl_sync(0xffffffffu, e.d0, 31);
            last.d1 = __shfl_sync(0xffffffffu, e.d1, 31);
            last.d2 = __shfl_sync(0xffffffffu, e.d2, 31);
            carry = dc_combine(carry, last);
        }
        if (lane == 0)
            s_carry = carry;
    }
    DcLocal L;
    dc_local(rec, j0, hi, bpm, im.restart_interval * bpm, blk_comp, L);
    const DcElem inc = dc_warp_scan(L.e, lane);
    if (lane == 31)
        s_warp[warp] = inc;
    __syncthreads();
    DcElem pre = s_carry;
    for (int i = 0; i < warp; ++i)
        pre = dc_combine(pre, s_warp[i]);
    const DcElem prev = dc_shfl_up(inc, 1);
    const DcElem ex = (lane == 0) ? pre : dc_combine(pre, prev); // everything before this thread's blocks
    // second walk: add the incoming predictors to the blocks that precede this thread's first reset
    uint32_t bim = j0 % bpm;
#pragma unroll
    for (int i = 0; i < kDcPer; ++i) {
        const uint32_t j = j0 + i;
        const uint32_t c = (blk_comp >> (2u * bim)) & 3u;
        const int32_t add = (uint32_t(i) < L.reset_at) ? (c == 0u ? ex.d0 : (c == 1u ? ex.d1 : ex.d2)) : 0;
        if (j < hi)
            rec[j] = (L.r[i] & 0xFFFFu) | (uint32_t(L.v[i] + add) << 16);
        bim = (bim + 1u == bpm) ? 0u : bim + 1u;
    }
}

// ------------------------------------------------------------------------------------------------
// K4a: dequantise + ISLOW IDCT (jidctint.c arithmetic), one thread per 8x8 block
// ------------------------------------------------------------------------------------------------
template <int kShift, uint32_t kBias>
__device__ __forceinline__ void idct8(int32_t i0, int32_t i1, int32_t i2, int32_t i3, int32_t i4, int32_t i5, int32_t i6, int32_t i7,
    int32_t* o)
{
    int32_t z1 = (i2 + i6) * 4433;
    int32_t t2 = z1 - i6 * 15137;
    int32_t t3 = z1 + i2 * 6270;
    int32_t t0 = (i0 + i4) * 8192;
    int32_t t1 = (i0 - i4) * 8192;
    const int32_t a10 = t0 + t3, a13 = t0 - t3, a11 = t1 + t2, a12 = t1 - t2;
    t0 = i7;
    t1 = i5;
    t2 = i3;
    t3 = i1;
    z1 = t0 + t3;
    int32_t z2 = t1 + t2, z3 = t0 + t2, z4 = t1 + t3;
    const int32_t z5 = (z3 + z4) * 9633;
    t0 *= 2446;
    t1 *= 16819;
    t2 *= 25172;
    t3 *= 12299;
    z1 *= -7373;
    z2 *= -20995;
    z3 = z3 * -16069 + z5;
    z4 = z4 * -3196 + z5;
    t0 += z1 + z3;
    t1 += z2 + z4;
    t2 += z2 + z3;
    t3 += z1 + z4;
    // DESCALE; kBias (a multiple of 1 << kShift) rides along with the rounding constant.  Unsigned adds: with the
    // bias the sum may wrap, and only bits kShift.. of it are used by the caller then.
    constexpr uint32_t rnd = (1u << (kShift - 1)) + kBias;
    auto descale = [](int32_t a, int32_t b) { return int32_t(uint32_t(a) + uint32_t(b) + rnd) >> kShift; };
    o[0] = descale(a10, t3);
    o[7] = descale(a10, -t3);
    o[1] = descale(a11, t2);
    o[6] = descale(a11, -t2);
    o[2] = descale(a12, t1);
    o[5] = descale(a12, -t1);
    o[3] = descale(a13, t0);
    o[4] = descale(a13, -t0);
}

// libjpeg's range_limit table indexed with (x & 1023), i.e. clamp(x + 128) with a 10-bit wrap:
//   i = x & 1023;  i < 128 -> i + 128;  i < 512 -> 255;  i < 896 -> 0;  else i - 896.
// With u = (x + 128) & 1023 that is u for u < 256, 255 for u < 640, 0 above; shifting the wrap
// point to 640 turns it into one clamp: t = ((x + 512) & 1023) - 384, result = min(max(t, 0), 255).
// The + 512 is added by the caller: pass 2 of the IDCT folds 512 << 18 into its rounding constant.
__device__ __forceinline__ uint32_t range_limit(int32_t x)
{
    return uint32_t(__vimin_s32_relu((x & 1023) - 384, 255)); // x already carries the + 512
}

#include "pixels.cuh"
#include "progressive.cuh"

__global__ void __launch_bounds__(128, 6) idct_kernel(BatchDev bd, int redo_only)
{
    __shared__ uint16_t s_q[3][64];
    __shared__ uint32_t s_qz[3][32]; // 8-bit quantisers paired in zig-zag order (px_idct_rows)
    __shared__ uint32_t s_scatter[32 * 128]; // entry-stream form: one column of 32 words per thread (sym_fetch_block)
    const uint32_t cta = blockIdx.x;
    const uint32_t img = bd.idct_cta_img[cta];
    const ImageDev& im = bd.img[img];
    const uint32_t flags = bd.state[img].flags;
    if ((flags & (kFlagDirty | kFlagBadMarker)) || (redo_only && !(flags & kFlagRedo)))
        return;
    for (int i = threadIdx.x; i < 3 * 64; i += blockDim.x)
        s_q[i / 64][i % 64] = im.qt[i / 64][i % 64];
    if (threadIdx.x < 96) {
        constexpr uint8_t kNatural[64] = {0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6, 7,
            14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62,
            63};
        const int c = threadIdx.x >> 5, j = threadIdx.x & 31;
        s_qz[c][j] = uint32_t(im.qt[c][kNatural[2 * j]] & 0xFFu) | (uint32_t(im.qt[c][kNatural[2 * j + 1]] & 0xFFu) << 24);
    }
    const bool q8 = im.q8 != 0; // every quantiser fits 8 bits: dequantise with dot-product instructions straight from the packed pairs
    __syncthreads();
    // block index over the wanted rows of the planes of the needed components, row-major per plane; a thread takes
    // kIdctPer blocks, blockDim.x apart (neighbouring threads keep neighbouring blocks: the row stores coalesce)
    const uint32_t first = (cta - im.idct_cta_off) * (kIdctPer * blockDim.x) + threadIdx.x;
    int c = 0;
    uint32_t bw = 0, by_end = 0, by = 0, bx = 0, lh = 0, lv = 0; // component, its width in blocks, end of its wanted rows, position
    bool locate = true;
#pragma unroll 1
    for (int k = 0; k < kIdctPer; ++k) {
    if (locate) { // from the linear index (first block; later ones only when they cross into the next component)
        uint32_t idx = first + uint32_t(k) * blockDim.x;
        for (c = 0; c < im.ncomp_needed; ++c) {
            bw = uint32_t(im.mcus_x) * im.h[c];
            const uint32_t rows = uint32_t(im.mcu_row_hi - im.mcu_row_lo) * im.v[c];
            if (idx < bw * rows) {
                by = uint32_t(im.mcu_row_lo) * im.v[c];
                by_end = by + rows;
                break;
            }
            idx -= bw * rows;
        }
        if (c >= im.ncomp_needed)
            return;
        by += idx / bw;
        bx = idx % bw;
        lh = 31u - uint32_t(__clz(int(im.h[c]))); // sampling factors are 1, 2 or 4
        lv = 31u - uint32_t(__clz(int(im.v[c])));
        locate = false;
    }
    const uint32_t mx = bx >> lh, my = by >> lv;
    const uint32_t bim = im.blk_first[c] + ((by - (my << lv)) << lh) + (bx - (mx << lh));
    const size_t blk_in_img = size_t(my * im.mcus_x + mx) * im.bpm + bim;
    // dense blocks: the sequential path's arena, or the progressive one (whose images never have an entry stream)
    const int16_t* dense = im.prog ? bd.pcoef + im.pcoef_off : bd.coef + im.coef_off;
    const size_t blk = im.coef_off / 64 + blk_in_img;
    const uint32_t rec = __ldg(bd.blkrec + blk);
    uint32_t w[32]; // the block as packed zig-zag pairs (DC slot unused: the DC value is in the record)
    if (im.prog) {
        px_fetch_block<2>(dense, rec, blk_in_img, w);
    } else if (bd.sym_mode) {
        SymSource S;
        S.sym = bd.sym + im.sym_off;
        S.chunks = bd.sym_chunks;
        S.shift = bd.sub_shift;
        S.n_sub = bd.state[img].n_sub;
        static_assert(kScatterStride == 128u * 4u, "the entries carry offsets into this column layout");
        sym_fetch_block(S, rec, __ldg(bd.blkpos + blk), smem_addr(s_scatter) + threadIdx.x * 4u, w);
    } else {
        px_fetch_block<2>(bd.coef, rec, blk, w);
    }
    if (q8) {
        uint8_t* dst8 = bd.planes + im.plane_off[c] + size_t(by) * 8 * (bw * 8) + bx * 8;
        const uint32_t pitch = bw * 8;
        px_idct_rows<0>(w, int32_t(rec) >> 16, s_qz[c], [](int) {}, [&](int y, uint32_t lo, uint32_t hi) {
            *reinterpret_cast<uint2*>(dst8 + size_t(y) * pitch) = make_uint2(lo, hi);
        });
    } else {
    // 16-bit quantisers (extended-sequential / progressive streams only): 64 coefficients, zig-zag order
    int32_t zz[64];
    {
#pragma unroll
        for (int j = 0; j < 32; ++j) {
            zz[2 * j] = int32_t(int16_t(w[j] & 0xFFFFu));
            zz[2 * j + 1] = int32_t(int16_t(w[j] >> 16));
        }
        zz[0] = int32_t(rec) >> 16;
    }
    // zig-zag position of each natural-order index (inverse of T.81 Figure A.6)
    constexpr int kZzPos[64] = {0, 1, 5, 6, 14, 15, 27, 28, 2, 4, 7, 13, 16, 26, 29, 42, 3, 8, 12, 17, 25, 30, 41, 43, 9, 11, 18, 24,
        31, 40, 44, 53, 10, 19, 23, 32, 39, 45, 52, 54, 20, 22, 33, 38, 46, 51, 55, 60, 21, 34, 37, 47, 50, 56, 59, 61, 35, 36, 48,
        49, 57, 58, 62, 63};
    const uint16_t* q = s_q[c];
    int32_t ws[64];
    // pass 1: columns, results scaled up by 2^PASS1_BITS
#pragma unroll
    for (int x = 0; x < 8; ++x) {
        int32_t col[8], o[8];
#pragma unroll
        for (int y = 0; y < 8; ++y)
            col[y] = zz[kZzPos[y * 8 + x]] * int32_t(q[y * 8 + x]);
        idct8<11, 0u>(col[0], col[1], col[2], col[3], col[4], col[5], col[6], col[7], o);
#pragma unroll
        for (int y = 0; y < 8; ++y)
            ws[y * 8 + x] = o[y];
    }
    // pass 2: rows, descale by CONST_BITS + PASS1_BITS + 3, range-limit, store 8 bytes per row
    uint8_t* dst = bd.planes + im.plane_off[c] + size_t(by) * 8 * (bw * 8) + bx * 8;
#pragma unroll
    for (int y = 0; y < 8; ++y) {
        int32_t o[8];
        idct8<18, (512u << 18)>(ws[y * 8 + 0], ws[y * 8 + 1], ws[y * 8 + 2], ws[y * 8 + 3], ws[y * 8 + 4], ws[y * 8 + 5], ws[y * 8 + 6], ws[y * 8 + 7], o);
        uint2 pk;
        pk.x = range_limit(o[0]) | (range_limit(o[1]) << 8) | (range_limit(o[2]) << 16) | (range_limit(o[3]) << 24);
        pk.y = range_limit(o[4]) | (range_limit(o[5]) << 8) | (range_limit(o[6]) << 16) | (range_limit(o[7]) << 24);
        *reinterpret_cast<uint2*>(dst + size_t(y) * (bw * 8)) = pk;
    }
    } // 16-bit quantisers
    // the next block of this thread lies blockDim.x further on
    bx += blockDim.x;
    while (bx >= bw) {
        bx -= bw;
        ++by;
    }
    locate = by >= by_end;
    } // blocks of this thread
}

// ------------------------------------------------------------------------------------------------
// K4b: upsample + colour convert + store
// ------------------------------------------------------------------------------------------------
struct PlaneView {
    const uint8_t* p;
    int stride, dw, dh, hx, vx, fancy_h, fancy_v;
};

// libjpeg jdsample.c semantics: which of the triangle filters applies is decided
// per component from its expansion factors; everything else replicates.
__device__ __forceinline__ PlaneView make_plane(const BatchDev& bd, const ImageDev& im, int c)
{
    PlaneView pv;
    pv.p = bd.planes + im.plane_off[c];
    pv.stride = int(im.plane_stride_px[c]);
    pv.dw = int(im.comp_w[c]);
    pv.dh = int(im.comp_h[c]);
    pv.hx = im.hmax / im.h[c];
    pv.vx = im.vmax / im.v[c];
    const bool f = im.fancy != 0;
    const bool h2v1 = f && pv.hx == 2 && pv.vx == 1 && pv.dw > 2;
    const bool h1v2 = f && pv.hx == 1 && pv.vx == 2;
    const bool h2v2 = f && pv.hx == 2 && pv.vx == 2 && pv.dw > 2;
    pv.fancy_h = h2v1 || h2v2;
    pv.fancy_v = h1v2 || h2v2;
    return pv;
}

__device__ __forceinline__ int sample_plane(const PlaneView& pv, int x, int y)
{
    if (pv.hx == 1 && pv.vx == 1)
        return pv.p[size_t(y) * pv.stride + x];
    if (!pv.fancy_h && !pv.fancy_v)
        return pv.p[size_t(y / pv.vx) * pv.stride + x / pv.hx];
    if (pv.fancy_h && !pv.fancy_v) { // h2v1
        const uint8_t* s = pv.p + size_t(y) * pv.stride;
        const int i = x >> 1, c = s[i];
        if (x & 1)
            return (i == pv.dw - 1) ? c : (3 * c + s[i + 1] + 2) >> 2;
        return (i == 0) ? c : (3 * c + s[i - 1] + 1) >> 2;
    }
    const int i = y >> 1;
    int nb = (y & 1) ? i + 1 : i - 1;
    nb = nb < 0 ? 0 : (nb > pv.dh - 1 ? pv.dh - 1 : nb);
    const uint8_t* s0 = pv.p + size_t(i) * pv.stride;
    const uint8_t* s1 = pv.p + size_t(nb) * pv.stride;
    if (!pv.fancy_h) // h1v2
        return (3 * s0[x] + s1[x] + ((y & 1) ? 2 : 1)) >> 2;
    const int j = x >> 1; // h2v2
    const int cur = 3 * s0[j] + s1[j];
    if (x & 1) {
        if (j == pv.dw - 1)
            return (cur * 4 + 7) >> 4;
        return (3 * cur + 3 * s0[j + 1] + s1[j + 1] + 7) >> 4;
    }
    if (j == 0)
        return (cur * 4 + 8) >> 4;
    return (3 * cur + 3 * s0[j - 1] + s1[j - 1] + 8) >> 4;
}

// jdcolor.c fixed-point YCbCr -> RGB (SURVEY.md 8c "Colour"); __viaddmin_s32_relu(a, b, c) is
// max(min(a + b, c), 0) in one instruction, i.e. the add and the [0,255] clamp together.
__device__ __forceinline__ void ycc_to_rgb(int y, int cb, int cr, uint32_t& r, uint32_t& g, uint32_t& b)
{
    cb -= 128;
    cr -= 128;
    r = uint32_t(__viaddmin_s32_relu(y, (91881 * cr + 32768) >> 16, 255));
    g = uint32_t(__viaddmin_s32_relu(y, (-22554 * cb - 46802 * cr + 32768) >> 16, 255));
    b = uint32_t(__viaddmin_s32_relu(y, (116130 * cb + 32768) >> 16, 255));
}

// byte 2 of a, b, c, d -> one little-endian word
__device__ __forceinline__ uint32_t pack_b2(uint32_t a, uint32_t b, uint32_t c, uint32_t d)
{
    return __byte_perm(__byte_perm(a, b, 0x0062), __byte_perm(c, d, 0x0062), 0x5410);
}
constexpr int kColorTx = 32; // threads in x, 8 pixels each
constexpr int kColorTy = 8;  // thread rows, color_unit_rows output rows each

// EXIF orientation: where pixel (sx, sy) of the rw x rh (cropped) decode lands in the stored image.
// Codes as in the EXIF standard / src/parsers/exif_orientation.h; checked against OpenCV and Pillow.
__device__ __forceinline__ void orient_xy(uint32_t code, int sx, int sy, int rw, int rh, int& dx, int& dy)
{
    switch (code) {
    case 2: dx = rw - 1 - sx; dy = sy; break;            // mirror horizontal
    case 3: dx = rw - 1 - sx; dy = rh - 1 - sy; break;   // rotate 180
    case 4: dx = sx; dy = rh - 1 - sy; break;            // mirror vertical
    case 5: dx = sy; dy = sx; break;                     // transpose
    case 6: dx = rh - 1 - sy; dy = sx; break;            // rotate 90 CW
    case 7: dx = rh - 1 - sy; dy = rw - 1 - sx; break;   // transverse
    case 8: dx = sy; dy = rw - 1 - sx; break;            // rotate 270 CW
    default: dx = sx; dy = sy; break;
    }
}

// Stores one pixel (three channel values, already in output channel order) at its oriented position.
__device__ __forceinline__ void store_pixel_oriented(const ImageDev& im, int sx, int sy, uint32_t v0, uint32_t v1, uint32_t v2)
{
    int dx, dy;
    orient_xy(im.orient, sx, sy, int(im.roi_x1) - int(im.roi_x0), int(im.roi_y1) - int(im.roi_y0), dx, dy);
    if (im.fmt == kFmtY) {
        im.out[size_t(dy) * im.row_stride + dx] = uint8_t(v0);
    } else if (im.fmt == kFmtPRGB || im.fmt == kFmtPBGR || im.fmt == kFmtPYUV) {
        uint8_t* o = im.out + size_t(dy) * im.row_stride + dx;
        o[0] = uint8_t(v0);
        o[im.plane_stride] = uint8_t(v1);
        o[2 * im.plane_stride] = uint8_t(v2);
    } else {
        uint8_t* o = im.out + size_t(dy) * im.row_stride + size_t(dx) * 3;
        o[0] = uint8_t(v0);
        o[1] = uint8_t(v1);
        o[2] = uint8_t(v2);
    }
}

// Generic (any sampling, any alignment) path: one pixel at a time through sample_plane().
__device__ __forceinline__ void color_pixel_generic(const BatchDev& bd, const ImageDev& im, int x, int y)
{
    const PlaneView p0 = make_plane(bd, im, 0);
    uint32_t r, g, b;
    if (im.fmt == kFmtY || im.ncomp == 1) {
        r = g = b = uint32_t(sample_plane(p0, x, y));
    } else {
        const PlaneView p1 = make_plane(bd, im, 1), p2 = make_plane(bd, im, 2);
        if (im.fmt == kFmtIYUV || im.fmt == kFmtPYUV) {
            r = uint32_t(sample_plane(p0, x, y));
            g = uint32_t(sample_plane(p1, x, y));
            b = uint32_t(sample_plane(p2, x, y));
        } else {
            ycc_to_rgb(sample_plane(p0, x, y), sample_plane(p1, x, y), sample_plane(p2, x, y), r, g, b);
        }
    }
    const int ox = x - im.roi_x0, oy = y - im.roi_y0;
    if (im.fmt == kFmtIBGR || im.fmt == kFmtPBGR) {
        const uint32_t t = r;
        r = b;
        b = t;
    }
    if (im.orient > 1) {
        store_pixel_oriented(im, ox, oy, r, g, b);
        return;
    }
    if (im.fmt == kFmtY) {
        im.out[size_t(oy) * im.row_stride + ox] = uint8_t(r);
        return;
    }
    if (im.fmt == kFmtPRGB || im.fmt == kFmtPBGR || im.fmt == kFmtPYUV) {
        uint8_t* o = im.out + size_t(oy) * im.row_stride + ox;
        o[0] = uint8_t(r);
        o[im.plane_stride] = uint8_t(g);
        o[2 * im.plane_stride] = uint8_t(b);
    } else {
        uint8_t* o = im.out + size_t(oy) * im.row_stride + size_t(ox) * 3;
        o[0] = uint8_t(r);
        o[1] = uint8_t(g);
        o[2] = uint8_t(b);
    }
}

// Stores 4 * NW contiguous bytes (little-endian words w[0..NW)) at ANY byte address.  Tightly pitched rows of
// odd widths put most rows of an image off the 8-byte grid: aligned words go out as words (funnel-shifted by
// the misalignment, which is the same for every lane of the warp), only the 4 bytes of head and tail singly.
template <int NW>
__device__ __forceinline__ void store_words_any(uint8_t* o, const uint32_t* w)
{
    const uint32_t a = uint32_t(reinterpret_cast<uintptr_t>(o)) & 3u;
    if (a == 0u) {
        if ((NW % 2) == 0 && (reinterpret_cast<uintptr_t>(o) & 7u) == 0) {
#pragma unroll
            for (int k = 0; k < NW / 2; ++k)
                reinterpret_cast<uint2*>(o)[k] = make_uint2(w[2 * k], w[2 * k + 1]);
        } else {
#pragma unroll
            for (int k = 0; k < NW; ++k)
                reinterpret_cast<uint32_t*>(o)[k] = w[k];
        }
        return;
    }
    const uint32_t head = 4u - a, sh = 8u * head; // bytes before the first aligned word; its position in w[0]
#pragma unroll
    for (int j = 0; j < 3; ++j)
        if (uint32_t(j) < head)
            o[j] = uint8_t(w[0] >> (8 * j));
    uint32_t* o32 = reinterpret_cast<uint32_t*>(o + head);
#pragma unroll
    for (int k = 0; k + 1 < NW; ++k)
        o32[k] = __funnelshift_r(w[k], w[k + 1], sh);
    uint8_t* tail = o + head + 4 * (NW - 1);
#pragma unroll
    for (int j = 0; j < 3; ++j)
        if (uint32_t(j) < a)
            tail[j] = uint8_t(w[NW - 1] >> (sh + 8 * j));
}

// Streams whose components came in three scans: the image proper has no entropy data, its planes were decoded as
// one-component images of their own (BatchDev::gate).  Their state flags become the image's before K4b looks at them.
__global__ void gate_flags_kernel(BatchDev bd)
{
    const int img = blockIdx.x * blockDim.x + threadIdx.x;
    if (img >= bd.n_images || !bd.img[img].multi)
        return;
    bd.state[img].flags = bd.state[bd.gate[3 * img]].flags | bd.state[bd.gate[3 * img + 1]].flags | bd.state[bd.gate[3 * img + 2]].flags;
}

// K4b.  Specialised sampling modes without EXIF orientation: a thread owns 16 pixels x (4 unit rows of 1 or 2 output rows),
// a CTA 512 x (32 or 64) pixels (px_color_global).  Everything else (generic sampling, orientation): one pixel at a time.
__global__ void __launch_bounds__(kColorTx* kColorTy, 4) color_kernel(BatchDev bd, int redo_only)
{
    const uint32_t cta = blockIdx.x;
    const uint32_t img = bd.color_cta_img[cta];
    const ImageDev& im = bd.img[img];
    const uint32_t flags = bd.state[img].flags;
    if ((flags & (kFlagDirty | kFlagBadMarker)) || (redo_only && !(flags & kFlagRedo)))
        return;
    const uint32_t local = cta - im.color_cta_off;
    const uint32_t cy = im.n_color_cta_x == 1u ? local : __umulhi(local, im.color_cta_x_rcp); // exact for local < 2^16 * n_color_cta_x
    const uint32_t cx = local - cy * im.n_color_cta_x;
    const int tx = threadIdx.x % kColorTx, ty = threadIdx.x / kColorTx;
    const int urows = im.color_unit_rows;
    if (im.color_mode >= 1 && im.orient <= 1) {
        const int X0 = (int(im.roi_x0) & ~(kPxUnit - 1)) + (int(cx) * kColorTx + tx) * kPxUnit;
        const int i_first = int(im.roi_y0) / urows + (int(cy) * kColorTy + ty) * kColorUnitsY;
        if (X0 >= int(im.roi_x1) || i_first * urows >= int(im.roi_y1))
            return;
        const PxColor K = px_color_setup(im.fmt);
        PxOut po;
        po.out = im.out;
        po.plane_stride = im.plane_stride;
        po.row_stride = im.row_stride;
        po.roi_x0 = im.roi_x0;
        po.roi_x1 = im.roi_x1;
        po.roi_y0 = im.roi_y0;
        po.dw = im.comp_w[1];
        po.dh = im.comp_h[1];
        PxPlanes P;
        P.y = bd.planes + im.plane_off[0];
        P.cb = bd.planes + im.plane_off[1];
        P.cr = bd.planes + im.plane_off[2];
        P.ystride = int(im.plane_stride_px[0]);
        P.cstride = int(im.plane_stride_px[1]);
        P.dw = po.dw;
        P.dh = po.dh;
        const int y_hi = im.roi_y1;
        switch (im.color_mode) {
        case 1: px_color_global<1, 1, false, true>(po, P, K, X0, i_first, kColorUnitsY, y_hi); break;
        case 2: px_color_global<1, 1, false, false>(po, P, K, X0, i_first, kColorUnitsY, y_hi); break;
        case 3: px_color_global<2, 1, false, false>(po, P, K, X0, i_first, kColorUnitsY, y_hi); break;
        case 4: px_color_global<1, 2, false, false>(po, P, K, X0, i_first, kColorUnitsY, y_hi); break;
        case 5: px_color_global<2, 2, false, false>(po, P, K, X0, i_first, kColorUnitsY, y_hi); break;
        case 6: px_color_global<2, 1, true, false>(po, P, K, X0, i_first, kColorUnitsY, y_hi); break;
        case 7: px_color_global<1, 2, true, false>(po, P, K, X0, i_first, kColorUnitsY, y_hi); break;
        default: px_color_global<2, 2, true, false>(po, P, K, X0, i_first, kColorUnitsY, y_hi); break;
        }
        return;
    }
    const int X0 = (int(im.roi_x0) & ~7) + (int(cx) * kColorTx + tx) * 8;
    const int Y0 = (int(im.roi_y0) & ~(urows - 1)) + ((int(cy) * kColorTy + ty) * kColorUnitsY) * urows; // urows is 1 or 2; kColorUnitsY units per thread
    if (X0 >= int(im.roi_x1) || Y0 >= int(im.roi_y1))
        return;
    for (int r = 0; r < urows * kColorUnitsY; ++r)
        for (int i = 0; i < 8; ++i) {
            const int x = X0 + i, y = Y0 + r;
            if (x >= int(im.roi_x0) && x < int(im.roi_x1) && y >= int(im.roi_y0) && y < int(im.roi_y1))
                color_pixel_generic(bd, im, x, y);
        }
}

constexpr size_t kSyncSmem = kSyncPairOffset + 2 * sizeof(HuffTableDev);
constexpr size_t kWriteSmem = sizeof(HuffSetDev) + sizeof(BlockTable);

} // namespace

int launch_scan(const BatchDev& b, cudaStream_t s)
{
    if (b.n_tiles == 0)
        return 0;
    scan_count_kernel<<<b.n_tiles, kScanThreads, 0, s>>>(b);
    scan_prefix_kernel<<<(b.n_images + 3) / 4, 128, 0, s>>>(b);
    scan_scatter_kernel<<<b.n_tiles, kScanThreads, 0, s>>>(b);
    return 3;
}

int launch_sync(const BatchDev& b, bool fix_mode, bool no_warm, cudaStream_t s)
{
    if (b.n_seq == 0)
        return 0;
    // dynamic shared memory opt-in is per device
    static bool attr_set[64] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev >= 0 && dev < 64 && !attr_set[dev]) {
        cudaFuncSetAttribute(huff_sync_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(kSyncSmem));
        cudaFuncSetAttribute(huff_sync_kernel<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(kSyncSmem));
        cudaFuncSetAttribute(huff_write_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(kWriteSmem));
        cudaFuncSetAttribute(huff_emit_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(kWriteSmem));
        attr_set[dev] = true;
    }
    const int mode = (fix_mode ? 1 : 0) | (no_warm ? 2 : 0);
    if (b.seq_warm == 16) // the two values the host uses (decoder.cpp: 8 Kbit of warm-up, at least 8 subsequences)
        huff_sync_kernel<16><<<b.n_seq, kSeqThreads, kSyncSmem, s>>>(b, mode);
    else
        huff_sync_kernel<8><<<b.n_seq, kSeqThreads, kSyncSmem, s>>>(b, mode);
    return 1;
}

int launch_verify_scan(const BatchDev& b, bool fix_mode, cudaStream_t s)
{
    if (b.n_images == 0)
        return 0;
    verify_scan_kernel<<<b.n_images, 256, 0, s>>>(b, fix_mode ? 1 : 0);
    return 1;
}

int launch_write(const BatchDev& b, bool redo_only, cudaStream_t s)
{
    if (b.n_wseq == 0)
        return 0;
    if (b.sym_mode)
        huff_emit_kernel<<<b.n_wseq, kSeqThreads, kWriteSmem, s>>>(b, redo_only ? 1 : 0);
    else
        huff_write_kernel<<<b.n_wseq, kSeqThreads, kWriteSmem, s>>>(b, redo_only ? 1 : 0);
    return 1;
}

int launch_dc_scan(const BatchDev& b, bool redo_only, cudaStream_t s)
{
    if (b.n_dc_cta == 0)
        return 0;
    dc_sum_kernel<<<b.n_dc_cta, kDcThreads, 0, s>>>(b, redo_only ? 1 : 0);
    dc_apply_kernel<<<b.n_dc_cta, kDcThreads, 0, s>>>(b, redo_only ? 1 : 0);
    return 2;
}

int launch_idct(const BatchDev& b, bool redo_only, cudaStream_t s)
{
    if (b.n_idct_cta == 0)
        return 0;
    idct_kernel<<<b.n_idct_cta, 128, 0, s>>>(b, redo_only ? 1 : 0);
    return 1;
}

int launch_progressive(const BatchDev& b, cudaStream_t s)
{
    if (b.n_prog_img == 0 || b.n_prog_chain == 0)
        return 0;
    cudaMemsetAsync(b.prog_progress, 0, 4ull * size_t(b.n_prog_chain), s);
    prog_scan_kernel<<<b.n_prog_chain, 32, 0, s>>>(b);
    prog_record_kernel<<<dim3((b.prog_max_blocks * 8u + 255u) / 256u, unsigned(b.n_prog_img)), 256, 0, s>>>(b);
    return 2;
}

int launch_color(const BatchDev& b, bool redo_only, cudaStream_t s)
{
    if (b.n_color_cta == 0)
        return 0;
    int launched = 1;
    if (b.n_multi) {
        gate_flags_kernel<<<(b.n_images + 127) / 128, 128, 0, s>>>(b);
        ++launched;
    }
    color_kernel<<<b.n_color_cta, kColorTx * kColorTy, 0, s>>>(b, redo_only ? 1 : 0);
    return launched;
}

// Shared memory of one tile of the fused pixel kernel (the same arithmetic as pixels_kernel's set-up).
size_t pixels_tile_smem(int ncomp, const int* h, const int* v, int vmax, int strip_mcus, bool fancy_h, bool fancy_v)
{
    (void)vmax;
    const int carry = fancy_v ? 2 : 0;
    size_t total = 0;
    for (int c = 0; c < ncomp; ++c) {
        const int cols = (strip_mcus + ((c > 0 && fancy_h) ? 2 : 0)) * h[c];
        total += size_t(carry + 8 * v[c]) * size_t((cols * 8 + 15) / 16 * 16 + 32); // rows padded to 16 bytes + kPxApron on either side
    }
    return (total + 15) / 16 * 16;
}

// A loop iteration of K3 consumes at least one bit (the shortest code) or hops over a restart boundary (at most one per byte);
// one more entry ends the list.
uint32_t sym_chunks_for(uint32_t sub_bits) { return (sub_bits + sub_bits / 8 + 16 + 3) / 4; } // four entries to a chunk

size_t pixels_smem_budget() { return (size_t(227 - 5 * kPxCtasPerSm) * 1024 / (kPxCtasPerSm * kPxWarps)) / 256 * 256; } // 8704 bytes per warp at 3 x 8 warps
int pixels_strip_mcus(int bpm_held, int halo_blocks) { return std::max(1, (96 - halo_blocks) / std::max(1, bpm_held)); }

int launch_pixels(const BatchDev& b, bool redo_only, cudaStream_t s)
{
    if (b.n_px_task == 0)
        return 0;
    static bool attr_set[64] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev >= 0 && dev < 64 && !attr_set[dev]) {
        cudaFuncSetAttribute(pixels_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(pixels_smem_budget() * kPxWarps));
        attr_set[dev] = true;
    }
    static const int debug = std::getenv("B200JPEG_PX_DEBUG") ? std::atoi(std::getenv("B200JPEG_PX_DEBUG")) : 0; // 2: no phase B, 4: no phase A (timing experiments)
    pixels_kernel<<<(b.n_px_task + kPxWarps - 1) / kPxWarps, kPxThreads, size_t(b.px_smem) * kPxWarps, s>>>(b, (redo_only ? 1 : 0) | debug);
    return 1;
}

} // namespace b200jpeg

// ------------------------------------------------------------------------------------------------
// b200jpegScanSegmentNoAlloc: the K0 scan for callers that must not allocate
// ------------------------------------------------------------------------------------------------
// Everything K0 needs besides the caller's input and outputs -- the segment descriptor, the
// tile->segment map and the per-tile counts / prefixes -- lives in this fixed array, which every
// GPU that loads the library carries until the process exits.
namespace b200jpeg {
namespace {

constexpr size_t kNoAllocMaxTiles = B200JPEG_NOALLOC_MAX_SEGMENT_BYTES / kScanTileBytes;
constexpr size_t kNoAllocDescBytes = 1024;
static_assert(sizeof(ImageDev) <= kNoAllocDescBytes, "descriptor slot too small");
constexpr size_t kNoAllocOffTileImg = kNoAllocDescBytes;
constexpr size_t kNoAllocOffInfo = kNoAllocOffTileImg + 4 * kNoAllocMaxTiles;
constexpr size_t kNoAllocOffBase = kNoAllocOffInfo + 8 * kNoAllocMaxTiles;
constexpr size_t kNoAllocScratchBytes = kNoAllocOffBase + 8 * kNoAllocMaxTiles; // 1 KiB + 20 B per 16 KiB tile = 11 KiB

static __device__ __align__(256) unsigned char g_noalloc_scan_scratch[kNoAllocScratchBytes];

struct NoAllocParams {
    uint32_t raw_len, n_tiles, n_rst_expected, n_bnd_cap;
};

// Rewrites the part of the shared scratch that has to start from a known value: the descriptor
// (all fields; the array still holds the previous call's) and the tile->segment map (all zero).
__global__ void noalloc_scan_setup_kernel(unsigned char* scratch, NoAllocParams p)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t* desc_words = reinterpret_cast<uint32_t*>(scratch);
    if (i < kNoAllocDescBytes / 4)
        desc_words[i] = 0u;
    if (i < p.n_tiles)
        reinterpret_cast<uint32_t*>(scratch + kNoAllocOffTileImg)[i] = 0u;
    __syncthreads();
    if (i == 0) {
        ImageDev* im = reinterpret_cast<ImageDev*>(scratch);
        im->raw_len = p.raw_len;
        im->n_tiles = p.n_tiles;
        im->n_rst_expected = p.n_rst_expected;
        im->n_bnd_cap = p.n_bnd_cap;
        im->n_sub_max = 0xFFFFFFFFu; // (no subsequence arrays here: the count is only reported)
    }
}

// Device address of the scratch array, resolved once per device (the array's name is only a
// host-side handle; kernels need the address cudaGetSymbolAddress reports for the current device).
std::mutex g_noalloc_mu;
unsigned char* g_noalloc_ptr[64] = {};

int noalloc_scratch_for_current_device(unsigned char** out)
{
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64)
        return B200JPEG_ERR_CUDA;
    std::lock_guard<std::mutex> lock(g_noalloc_mu);
    if (!g_noalloc_ptr[dev]) {
        unsigned char* ptr = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(reinterpret_cast<void**>(&ptr), g_noalloc_scan_scratch);
        if (e != cudaSuccess || !ptr) {
            cudaGetLastError();
            return B200JPEG_ERR_CUDA;
        }
        g_noalloc_ptr[dev] = ptr;
    }
    *out = g_noalloc_ptr[dev];
    return B200JPEG_OK;
}

void noalloc_message(char* message, size_t capacity, const char* text)
{
    if (message && capacity) {
        std::strncpy(message, text, capacity - 1);
        message[capacity - 1] = '\0';
    }
}

} // namespace
} // namespace b200jpeg

extern "C" B200JPEG_API int b200jpegScanSegmentNoAlloc(const void* dev_raw, size_t length, uint32_t expected_restart_markers,
    void* dev_out_words, uint32_t* dev_out_boundaries, uint32_t boundary_capacity, b200jpegScanResult_t* dev_result, void* cuda_stream,
    char* message, size_t message_capacity)
{
    using namespace b200jpeg;
    noalloc_message(message, message_capacity, "");
    if (!dev_raw || !dev_out_words || !dev_out_boundaries || !dev_result || length == 0 || (reinterpret_cast<uintptr_t>(dev_raw) & 15u) ||
        (reinterpret_cast<uintptr_t>(dev_out_words) & 3u) || boundary_capacity < expected_restart_markers + 1) {
        noalloc_message(message, message_capacity, "b200jpegScanSegmentNoAlloc: null / misaligned pointer, empty segment or boundary array too small");
        return B200JPEG_ERR_ARG;
    }
    if (length > size_t(B200JPEG_NOALLOC_MAX_SEGMENT_BYTES)) {
        noalloc_message(message, message_capacity,
            "b200jpegScanSegmentNoAlloc: segment longer than B200JPEG_NOALLOC_MAX_SEGMENT_BYTES (8 MiB); the fixed scratch cannot hold its tile counts");
        return B200JPEG_ERR_TOO_LARGE;
    }
    unsigned char* scratch = nullptr;
    if (noalloc_scratch_for_current_device(&scratch) != B200JPEG_OK) {
        noalloc_message(message, message_capacity, "b200jpegScanSegmentNoAlloc: cudaGetSymbolAddress failed on the current device");
        return B200JPEG_ERR_CUDA;
    }
    cudaStream_t stream = static_cast<cudaStream_t>(cuda_stream);
    NoAllocParams p;
    p.raw_len = uint32_t(length);
    p.n_tiles = uint32_t((length + kScanTileBytes - 1) / kScanTileBytes);
    p.n_rst_expected = expected_restart_markers;
    p.n_bnd_cap = boundary_capacity;
    // block 0 clears and refills the descriptor (one barrier apart); all blocks clear the tile map
    const uint32_t setup_threads = p.n_tiles > uint32_t(kNoAllocDescBytes / 4) ? p.n_tiles : uint32_t(kNoAllocDescBytes / 4);
    noalloc_scan_setup_kernel<<<(setup_threads + 1023) / 1024, 1024, 0, stream>>>(scratch, p);
    BatchDev b{};
    b.n_images = 1;
    b.img = reinterpret_cast<const ImageDev*>(scratch);
    b.tile_img = reinterpret_cast<const uint32_t*>(scratch + kNoAllocOffTileImg);
    b.tile_info = reinterpret_cast<uint2*>(scratch + kNoAllocOffInfo);
    b.tile_base = reinterpret_cast<uint2*>(scratch + kNoAllocOffBase);
    b.state = reinterpret_cast<ImageState*>(dev_result);
    b.raw = static_cast<const uint8_t*>(dev_raw);
    b.cmp = static_cast<uint32_t*>(dev_out_words);
    b.bnd = dev_out_boundaries;
    b.n_tiles = int(p.n_tiles);
    b.sub_bits = 1024; // only sizes the n_sub field of the result (documented as 1024-bit units)
    launch_scan(b, stream);
    if (cudaGetLastError() != cudaSuccess) {
        noalloc_message(message, message_capacity, "b200jpegScanSegmentNoAlloc: kernel launch failed");
        return B200JPEG_ERR_CUDA;
    }
    return B200JPEG_OK;
}
