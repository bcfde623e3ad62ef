// K4 (fused): dequantise + ISLOW IDCT -> component tiles in shared memory -> chroma upsampling + YCbCr->RGB ->
// 128-bit stores into the caller's buffer.  Included by kernels.cu (inside its anonymous namespace, after the
// IDCT / colour helpers).
//
// Replaces the scanline loop of the reference's CPU path (extensions/libjpeg_turbo/jpeg_mem.cpp:311-379 driving
// libjpeg-turbo's jidctint / jdsample / jdcolor); the component planes never leave the SM.
//
// One WARP owns a tile: a strip of MCU columns x a band of MCU rows of one image, and walks the band top to bottom, one
// MCU row per step:
//   phase A  every lane takes 8x8 blocks of that MCU row: coefficients -> registers -> IDCT -> 8-byte row stores
//            into the warp's component tiles in shared memory;
//   phase B  every lane takes units of 16 pixels x (1 or 2) rows: luma and chroma from shared memory, triangle or box
//            upsampling, colour transform, 16-byte + 8-byte stores (interleaved RGB: 48 bytes per row).
// Warps never wait for each other (the phases are separated by __syncwarp only): the 24 warps of an SM drift apart, so
// that at any time some transform, some colour-convert and some wait for their coefficients -- which is what keeps the
// integer pipes issuing.  A strip is as wide as gives a warp three full passes of blocks per step (96 blocks).
// The vertical triangle filter of 4:2:0 / 4:4:0 looks one chroma row up and down, so with it the output trails the
// IDCT by one chroma row (two luma rows): the last two rows of every tile are carried over to the next step, and a
// band starts / ends with a step that only transforms the one chroma row its first / last output rows look at (a
// one-row IDCT: the column pass computes one output per column).  The horizontal filter looks one chroma sample left
// and right: a strip keeps the chroma of one more MCU column on each side.
#pragma once

// defined further down in kernels.cu, next to the colour helpers
__device__ __forceinline__ uint32_t pack_b2(uint32_t a, uint32_t b, uint32_t c, uint32_t d);
template <int NW>
__device__ __forceinline__ void store_words_any(uint8_t* o, const uint32_t* w);

#ifndef B200_PX_THREADS
#define B200_PX_THREADS 256
#endif
constexpr int kPxThreads = B200_PX_THREADS;
constexpr int kPxCtasPerSm = kPxThreads >= 256 ? 3 : 6;
constexpr int kPxUnit = 16; // pixels per thread and row in phase B
constexpr int kPxApron = 16; // bytes before and after the samples of every plane row

struct PxComp {
    int bx0, nbx;     // block columns held: [bx0, bx0 + nbx)
    int lh, lv;       // log2 of the sampling factors
    int off, stride;  // plane in shared memory: byte offset, bytes per row
    int first;        // block-in-MCU index of the component's first block
};

struct PxTile {
    PxComp c[3];
    int ncomp;        // components held (1: luma only)
    int mx0, mx1;     // MCU columns of the strip
    int r_lo, r_hi;   // MCU rows of the band
    int lag;          // 1: the output trails the IDCT by one unit row (vertical triangle filter)
    int carry;        // rows carried over from the previous MCU row at the top of every plane (2 * lag)
    int urows;        // unit rows per MCU row (chroma rows per MCU row; luma rows when VX == 1)
};

// Coefficients of one block (dense layout: 16-byte groups in zig-zag order, a mask of the groups K3 wrote, DC in the record),
// as they lie in memory: w[j] = positions 2j (low half) and 2j + 1 (high half) of the zig-zag sequence; position 0 is not
// stored there, the DC value comes back separately.
// kHalf: 0 = groups 0..3, 1 = groups 4..7, 2 = all eight
template <int kHalf>
__device__ __forceinline__ void px_fetch_block(const int16_t* __restrict__ coef, uint32_t rec, size_t blk, uint32_t (&w)[32])
{
    const uint4* s4 = reinterpret_cast<const uint4*>(coef) + blk * 8;
#pragma unroll
    for (int i = (kHalf == 1 ? 4 : 0); i < (kHalf == 0 ? 4 : 8); ++i) {
        // (volatile: these loads are issued where the code says -- in the middle of the previous block's row pass)
        asm volatile("{\n\t.reg .pred q;\n\t"
                     "setp.ne.u32 q, %5, 0;\n\t"
                     "mov.u32 %0, 0;\n\tmov.u32 %1, 0;\n\tmov.u32 %2, 0;\n\tmov.u32 %3, 0;\n\t"
                     "@q ld.global.nc.v4.u32 {%0,%1,%2,%3}, [%4];\n\t}"
                     : "=r"(w[4 * i]), "=r"(w[4 * i + 1]), "=r"(w[4 * i + 2]), "=r"(w[4 * i + 3])
                     : "l"(s4 + i), "r"(rec & (1u << i)));
    }
}

// Coefficients of one block from the entry stream K3 wrote (huff_emit_kernel; format in kernels.cuh): the block's entries
// start behind the slot its record names, in the subsequence its DC code word lies in, and may continue in the next
// subsequence.  They are scattered into a per-thread column of shared memory (word j of thread t at scat + j * stride:
// consecutive threads hit consecutive banks) in the packed zig-zag pair layout the dense form has in memory, and read back.
// Every address is bounded by what the host reserved, whatever the entries say (damaged streams are reported, never
// followed out of bounds).
struct SymSource {
    const uint4* sym;      // entry stream of the image
    uint32_t chunks;       // chunks reserved per subsequence
    uint32_t shift;        // log2 of the subsequence length in bits
    uint32_t n_sub;        // subsequences of the image
};

__device__ __forceinline__ void sym_fetch_block(const SymSource& S, uint32_t rec, uint32_t pos, uint32_t scat, uint32_t (&w)[32])
{
#pragma unroll
    for (int j = 0; j < 32; ++j)
        asm volatile("st.shared.u32 [%0], %1;" ::"r"(scat + uint32_t(j) * kScatterStride), "r"(0u) : "memory");
    uint32_t k = pos >> S.shift, slot = (rec & 0xFFFFu) + 1u; // the entry behind the DC symbol's
    bool done = k >= S.n_sub || (slot >> 2) >= S.chunks;
    auto chunk_at = [&](uint32_t kk, uint32_t c) { return __ldg(S.sym + (size_t(kk >> 5) * S.chunks + c) * 32 + (kk & 31u)); };
    uint4 ch = make_uint4(0u, 0u, 0u, 0u);
    if (!done)
        ch = chunk_at(k, slot >> 2);
#pragma unroll 1
    for (int guard = 0; !done && guard < 320; ++guard) {
        const uint32_t c = slot >> 2;
        // the next chunk of the column is requested before this one is worked through (most blocks span several)
        uint4 nx = make_uint4(0u, 0u, 0u, 0u);
        const bool have_next = c + 1u < S.chunks;
        if (have_next)
            nx = chunk_at(k, c + 1u);
        const uint32_t cw[4] = {ch.x, ch.y, ch.z, ch.w};
        bool hop = false;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            if (done || hop || uint32_t(j) < (slot & 3u))
                continue;
            // every entry may be stored as it is: those without a coefficient put a zero at the unused DC position
            const uint32_t e = cw[j];
            // (the mask keeps the word row and the half bit only: whatever a damaged stream left in an entry, the store stays
            // aligned and inside this thread's column)
            asm volatile("st.shared.u16 [%0], %1;" ::"r"(scat + ((e >> 16) & 0x3E02u)), "h"(uint16_t(e)) : "memory");
            done = (e & 0x40000000u) != 0u; // the block ends with this entry
            hop = (e & 0x80000000u) != 0u;  // the subsequence's entries end: the block continues in the next one
        }
        if (hop) {
            ++k;
            slot = 0;
            done = k >= S.n_sub;
            if (!done)
                ch = chunk_at(k, 0u);
        } else {
            slot = (c + 1u) * 4u;
            ch = nx;
            done = done || !have_next;
        }
    }
#pragma unroll
    for (int j = 0; j < 32; ++j)
        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(w[j]) : "r"(scat + uint32_t(j) * kScatterStride) : "memory");
}

// signed 16-bit halves of a times unsigned bytes of b: lo = a.lo * b.b0 + a.hi * b.b1, hi = a.lo * b.b2 + a.hi * b.b3
__device__ __forceinline__ int32_t px_dp2a_lo(uint32_t a, uint32_t b)
{
    int32_t d;
    asm("dp2a.lo.s32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(0));
    return d;
}
__device__ __forceinline__ int32_t px_dp2a_hi(uint32_t a, uint32_t b)
{
    int32_t d;
    asm("dp2a.hi.s32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(0));
    return d;
}

__device__ __forceinline__ void px_sts64(uint32_t addr, uint32_t a, uint32_t b)
{
    asm volatile("st.shared.v2.u32 [%0], {%1, %2};" ::"r"(addr), "r"(a), "r"(b) : "memory");
}

// Dequantisation straight from the packed pairs: one dot-product instruction per coefficient picks the 16-bit half, sign-extends
// it and multiplies by the 8-bit quantiser (qz[j] = q(2j) | q(2j+1) << 24 in zig-zag order: the lo form sees q(2j) and a zero
// byte, the hi form a zero byte and q(2j+1)).  Baseline tables are 8 bits wide; streams with 16-bit tables take the idct kernel.
// dst: shared-window address of the block's first row.
// kRows: 0 = all eight rows, 1 = only row 7, 2 = only row 0 (the rows a band's first / last step needs); what is not
// asked for is never computed (everything is unrolled, unused column outputs are dead code).
// mid(0) / mid(1): called when four / six of the eight rows are out (the packed words are long dead by then) -- where the two
// halves of the next block's loads are issued: by then enough registers of this block are free to receive them.
// store(y, lo, hi): row y of the block, eight pixels in two words.
template <int kRows, typename Mid, typename Store>
__device__ __forceinline__ void px_idct_rows(const uint32_t (&w)[32], int32_t dc, const uint32_t* __restrict__ qz, Mid mid, Store store)
{
    constexpr int kZzPos[64] = {0, 1, 5, 6, 14, 15, 27, 28, 2, 4, 7, 13, 16, 26, 29, 42, 3, 8, 12, 17, 25, 30, 41, 43, 9, 11, 18, 24,
        31, 40, 44, 53, 10, 19, 23, 32, 39, 45, 52, 54, 20, 22, 33, 38, 46, 51, 55, 60, 21, 34, 37, 47, 50, 56, 59, 61, 35, 36, 48,
        49, 57, 58, 62, 63};
    constexpr int kNat[64] = {0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6, 7, 14, 21,
        28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};
    (void)kZzPos;
    // dequantised coefficients in natural order, produced word by word so that the packed words die as they are consumed
    int32_t ws[64];
    ws[0] = dc * int32_t(qz[0] & 0xFFu);
    ws[kNat[1]] = px_dp2a_hi(w[0], qz[0]);
#pragma unroll
    for (int j = 1; j < 32; ++j) {
        ws[kNat[2 * j]] = px_dp2a_lo(w[j], qz[j]);
        ws[kNat[2 * j + 1]] = px_dp2a_hi(w[j], qz[j]);
    }
    // pass 1: columns, in place
#pragma unroll
    for (int x = 0; x < 8; ++x) {
        int32_t o[8];
        idct8<11, 0u>(ws[x], ws[8 + x], ws[16 + x], ws[24 + x], ws[32 + x], ws[40 + x], ws[48 + x], ws[56 + x], o);
#pragma unroll
        for (int y = 0; y < 8; ++y)
            ws[y * 8 + x] = o[y];
    }
#pragma unroll
    for (int y = 0; y < 8; ++y) {
        if (y == 4)
            mid(0);
        if (y == 6)
            mid(1);
        if ((kRows == 1 && y != 7) || (kRows == 2 && y != 0))
            continue;
        int32_t o[8];
        idct8<18, (512u << 18)>(ws[y * 8 + 0], ws[y * 8 + 1], ws[y * 8 + 2], ws[y * 8 + 3], ws[y * 8 + 4], ws[y * 8 + 5], ws[y * 8 + 6], ws[y * 8 + 7], o);
        uint2 pk;
        pk.x = range_limit(o[0]) | (range_limit(o[1]) << 8) | (range_limit(o[2]) << 16) | (range_limit(o[3]) << 24);
        pk.y = range_limit(o[4]) | (range_limit(o[5]) << 8) | (range_limit(o[6]) << 16) | (range_limit(o[7]) << 24);
        store(y, pk.x, pk.y);
    }
}

template <int kRows, typename Mid>
__device__ __forceinline__ void px_idct_block(const uint32_t (&w)[32], int32_t dc, const uint32_t* __restrict__ qz, uint32_t dst, int stride, Mid mid)
{
    px_idct_rows<kRows>(w, dc, qz, mid, [&](int y, uint32_t lo, uint32_t hi) { px_sts64(dst + uint32_t(y * stride), lo, hi); });
}

// Phase A of one step: the blocks of MCU row r.  kind 0: every component, all rows; 1: chroma only, last row of the
// MCU row (the step above a band); 2: chroma only, first row (the step below a band).
struct PxSource {          // what phase A needs of the batch and the image, read once per CTA
    const int16_t* coef;
    const uint32_t* blkrec;
    size_t blk0;           // first block of the image in the batch arenas
    uint32_t mcus_x, bpm;
};

__device__ __forceinline__ void px_idct_phase(const PxSource src, const PxTile& T, const uint32_t (*s_qz)[32], uint32_t smem, int r, int kind)
{
    int n[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        const bool on = c < T.ncomp && (kind == 0 || c > 0);
        n[c] = on ? (T.c[c].nbx << (kind == 0 ? T.c[c].lv : 0)) : 0;
    }
    const int total = n[0] + n[1] + n[2];
    // item -> (component, block row, block column, block index)
    auto locate = [&](int t, int& c, int& by, int& local) -> uint32_t {
        c = 0;
        local = t;
        if (local >= n[0]) {
            local -= n[0];
            c = 1;
            if (local >= n[1]) {
                local -= n[1];
                c = 2;
            }
        }
        const PxComp& pc = T.c[c];
        by = 0; // block row inside the MCU row
        if (kind == 0) {
            if (local >= pc.nbx) { // sampling factors are 1 or 2 on this path
                local -= pc.nbx;
                by = 1;
            }
        } else if (kind == 1) {
            by = (1 << pc.lv) - 1;
        }
        const int bx = pc.bx0 + local;
        const int mcu = bx >> pc.lh;
        const uint32_t bim = uint32_t(pc.first + (by << pc.lh) + (bx - (mcu << pc.lh)));
        return (uint32_t(r) * src.mcus_x + uint32_t(mcu)) * src.bpm + bim;
    };
    const int lane = threadIdx.x & 31;
    if (lane >= total)
        return;
    // Software pipeline over the lane's items: while a block is transformed, the record of the next one is already
    // here and its coefficient words are requested as soon as the registers of the current block's words are free.
    int c, by, local;
    uint32_t blk = locate(lane, c, by, local);
    uint32_t rec = __ldg(src.blkrec + src.blk0 + blk);
    uint32_t w[32];
    px_fetch_block<2>(src.coef, rec, src.blk0 + blk, w);
#pragma unroll 1
    for (int t = lane; t < total; t += 32) {
        const int stride = T.c[c].stride;
        const uint32_t dst = smem + uint32_t(T.c[c].off + (T.carry + by * 8) * stride + kPxApron + local * 8);
        const uint32_t* qz = s_qz[c];
        const int32_t dc = int32_t(rec) >> 16;
        const bool more = t + 32 < total;
        if (more) {
            blk = locate(t + 32, c, by, local);
            rec = __ldg(src.blkrec + src.blk0 + blk);
        }
        auto next = [&](int half) {
            if (more) {
                if (half == 0)
                    px_fetch_block<0>(src.coef, rec, src.blk0 + blk, w);
                else
                    px_fetch_block<1>(src.coef, rec, src.blk0 + blk, w);
            }
        };
        if (kind == 0)
            px_idct_block<0>(w, dc, qz, dst, stride, next);
        else if (kind == 1)
            px_idct_block<1>(w, dc, qz, dst, stride, next);
        else
            px_idct_block<2>(w, dc, qz, dst, stride, next);
    }
}

// ---- phase B helpers
// Output buffers are device memory: say so, the pointer comes out of a descriptor and would otherwise be stored through generically.
__device__ __forceinline__ void px_stg128(uint8_t* p, uint32_t a, uint32_t b, uint32_t c, uint32_t d)
{
    asm volatile("st.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(__cvta_generic_to_global(p)), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ void px_stg64(uint8_t* p, uint32_t a, uint32_t b)
{
    asm volatile("st.global.v2.u32 [%0], {%1, %2};" ::"l"(__cvta_generic_to_global(p)), "r"(a), "r"(b) : "memory");
}
__device__ __forceinline__ int px_byte(uint32_t w, int k) { return int(__byte_perm(w, 0u, 0x4440 + k)); } // byte k of w

// Output layout of an image, uniform over the CTA.  The colour transform itself always runs in R, G, B terms with
// immediate constants (16 bits up, the -128 and the rounding folded in); BGR only changes the order of packing.
struct PxColor {
    int bgr;              // 1: channel 0 is blue
    int yuv;              // 1: no transform (Y, Cb, Cr as they are)
    int planar;           // 1: one plane per channel (or the single plane of Y output)
    int nch;              // planes to write
};

__device__ __forceinline__ PxColor px_color_setup(int fmt)
{
    PxColor k;
    k.bgr = (fmt == kFmtIBGR || fmt == kFmtPBGR) ? 1 : 0;
    k.yuv = (fmt == kFmtIYUV || fmt == kFmtPYUV) ? 1 : 0;
    k.planar = (fmt == kFmtY || fmt == kFmtPRGB || fmt == kFmtPBGR || fmt == kFmtPYUV) ? 1 : 0;
    k.nch = (fmt == kFmtY) ? 1 : 3;
    return k;
}

// Stores eight pixels of one output row: half `hf` (0 / 1) of a 16-pixel unit whose first byte lies at `o`.
// Interleaved: pk[0..5] = the 24 bytes; with `o` 16-byte aligned half 0 goes out as 16 + 8 bytes, half 1 (at o + 24) as
// 8 + 16 bytes, so that a unit row is two 128-bit and two 64-bit stores.  Planar: pk[2 * ch + 0..1] = the 8 bytes of channel ch.
// x_lo / x_hi: the part of the eight pixels inside the region of interest (0 / 8 when all of them are).
struct PxOut {             // where the pixels go, read once per CTA (stores through `out` would otherwise force re-reads of the descriptor)
    uint8_t* out;
    size_t plane_stride;
    uint32_t row_stride;
    int roi_x0, roi_x1, roi_y0;
    int dw, dh;            // real size of the chroma components
};

__device__ __forceinline__ void px_store_half(const PxOut& im, const PxColor& K, uint8_t* orow, int ox, int hf, const uint32_t* pk, int x_lo, int x_hi)
{
    const bool full = x_lo == 0 && x_hi == 8;
    if (K.planar) {
#pragma unroll
        for (int ch = 0; ch < 3; ++ch) {
            if (ch >= K.nch)
                break;
            uint8_t* o = orow + size_t(ch) * im.plane_stride + ox;
            if (full) {
                if ((reinterpret_cast<uintptr_t>(o) & 7u) == 0)
                    px_stg64(o, pk[2 * ch], pk[2 * ch + 1]);
                else
                    store_words_any<2>(o, pk + 2 * ch);
            } else {
#pragma unroll
                for (int i = 0; i < 8; ++i)
                    if (i >= x_lo && i < x_hi)
                        o[i] = uint8_t(pk[2 * ch + (i >> 2)] >> (8 * (i & 3)));
            }
        }
        return;
    }
    uint8_t* o = orow + ptrdiff_t(ox) * 3;
    if (full) {
        const uint32_t a = uint32_t(reinterpret_cast<uintptr_t>(o)) & 15u;
        if (hf == 0 && a == 0) {
            px_stg128(o, pk[0], pk[1], pk[2], pk[3]);
            px_stg64(o + 16, pk[4], pk[5]);
        } else if (hf == 1 && a == 8) {
            px_stg64(o, pk[0], pk[1]);
            px_stg128(o + 8, pk[2], pk[3], pk[4], pk[5]);
        } else {
            store_words_any<6>(o, pk);
        }
    } else {
#pragma unroll
        for (int b = 0; b < 24; ++b)
            if (b >= 3 * x_lo && b < 3 * x_hi)
                o[b] = uint8_t(pk[b >> 2] >> (8 * (b & 3)));
    }
}

// Phase B of one step for one sampling mode.  The step's output rows are Y0 = VX * (urows * r - lag) onwards; output row y
// belongs to chroma row i = y / VX, which sits (with its vertical neighbours i - 1 / i + 1, clamped to the component's real
// height, where libjpeg replicates) at plane row (chroma row) - (urows * r - carry); luma row y sits at plane row y - Y0.
// A lane takes one item at a time: 16 pixels of one output row, consecutive lanes consecutive pixels of the row.  The loop
// body is kept small on purpose (it shares the instruction cache with the IDCT): everything an item needs comes from
// shared memory when it needs it, worked through in two halves of eight pixels.  The aprons of the planes hold the
// replicated edge samples, so no address is clamped horizontally.
template <int HX, int VX, bool kFancy, bool kLumaOnly>
__device__ __forceinline__ void px_color_phase(const PxOut& im, const PxTile& T, const PxColor& K, const uint8_t* smem, int r, int y_lo, int y_hi)
{
    constexpr bool fancy_v = kFancy && VX == 2;
    const int lane = threadIdx.x & 31;
    const int X_strip = T.mx0 * (8 << T.c[0].lh);
    const int strip_px = T.c[0].nbx * 8;
    const int ucols = (strip_px + kPxUnit - 1) / kPxUnit;
    const uint32_t ucols_rcp = 65536u / uint32_t(ucols) + 1u; // u / ucols == (u * rcp) >> 16 for the few hundred items of a step
    const int dh = im.dh;
    const int i0 = T.urows * r - T.lag;
    const int Y0 = VX * i0;
    const int ya = max(y_lo, Y0), yb = min(y_hi, Y0 + VX * T.urows);
    if (ya >= yb)
        return;
    const int crow0 = T.urows * r - T.carry; // chroma row held at plane row 0
    const int ystride = T.c[0].stride;
    const uint8_t* p0 = smem + T.c[0].off + kPxApron;
    const int cstride = kLumaOnly ? 0 : T.c[1].stride;
    // plane a = Cb, plane b = Cr; pa points at the sample column of the strip's first pixel
    const uint8_t* pa = kLumaOnly ? p0 : smem + T.c[1].off + kPxApron + (X_strip / HX - T.c[1].bx0 * 8);
    const ptrdiff_t a_to_b = kLumaOnly ? 0 : ptrdiff_t(T.c[2].off) - ptrdiff_t(T.c[1].off);
    const int roi_x0 = im.roi_x0, roi_x1 = im.roi_x1, roi_y0 = im.roi_y0;
    const int n_items = (yb - ya) * ucols;
#pragma unroll 1
    for (int u = lane; u < n_items; u += 32) {
        const int row = int((uint32_t(u) * ucols_rcp) >> 16);
        const int ux = u - row * ucols;
        const int y = ya + row;
        const int i = VX == 2 ? (y >> 1) : y, rr = VX == 2 ? (y & 1) : 0;
        const int X0 = X_strip + ux * kPxUnit;
        if (X0 >= roi_x1 || X0 + kPxUnit <= roi_x0)
            continue;
        const uint8_t* yp = p0 + (y - Y0) * ystride + ux * kPxUnit;
        const uint8_t* ca = pa + (i - crow0) * cstride + ux * (kPxUnit / HX); // Cb, current row
        int nb = rr == 0 ? i - 1 : i + 1; // vertical neighbour, replicated at the component's top and bottom
        nb = nb < 0 ? 0 : (nb > dh - 1 ? dh - 1 : nb);
        const ptrdiff_t to_nb = fancy_v ? ptrdiff_t(nb - i) * cstride : 0;
        uint8_t* const orow = im.out + size_t(y - roi_y0) * im.row_stride;
#pragma unroll
        for (int hf = 0; hf < 2; ++hf) {
            const int X = X0 + 8 * hf;
            if (X >= roi_x1 || X + 8 <= roi_x0 || X - X_strip >= strip_px)
                continue;
            const int x_lo = max(roi_x0 - X, 0), x_hi = min(roi_x1 - X, 8);
            int ua[8], ub[8]; // the eight upsampled Cb / Cr samples
            if (!kLumaOnly) {
#pragma unroll
                for (int ab = 0; ab < 2; ++ab) {
                    const uint8_t* q = ca + (ab ? a_to_b : 0) + hf * (8 / HX);
                    int* out = ab ? ub : ua;
                    if (HX == 2) {
                        int s[6]; // left neighbour, four samples, right neighbour, vertically filtered
                        const uint32_t cw = *reinterpret_cast<const uint32_t*>(q);
                        s[0] = q[-1];
                        s[5] = q[4];
#pragma unroll
                        for (int k = 0; k < 4; ++k)
                            s[1 + k] = px_byte(cw, k);
                        if (fancy_v) {
                            const uint32_t nw = *reinterpret_cast<const uint32_t*>(q + to_nb);
                            s[0] = 3 * s[0] + q[to_nb - 1];
                            s[5] = 3 * s[5] + q[to_nb + 4];
#pragma unroll
                            for (int k = 0; k < 4; ++k)
                                s[1 + k] = 3 * s[1 + k] + px_byte(nw, k);
                        }
#pragma unroll
                        for (int k = 0; k < 4; ++k) {
                            if (fancy_v) { // h2v2 triangle (jdsample.c)
                                out[2 * k] = (3 * s[k + 1] + 8 + s[k]) >> 4;
                                out[2 * k + 1] = (3 * s[k + 1] + 7 + s[k + 2]) >> 4;
                            } else if (kFancy) { // h2v1 triangle
                                out[2 * k] = (3 * s[k + 1] + 1 + s[k]) >> 2;
                                out[2 * k + 1] = (3 * s[k + 1] + 2 + s[k + 2]) >> 2;
                            } else {
                                out[2 * k] = out[2 * k + 1] = s[k + 1];
                            }
                        }
                    } else {
                        const uint2 cw = *reinterpret_cast<const uint2*>(q);
#pragma unroll
                        for (int k = 0; k < 8; ++k)
                            out[k] = px_byte(k < 4 ? cw.x : cw.y, k & 3);
                        if (fancy_v) { // h1v2 triangle
                            const uint2 nw = *reinterpret_cast<const uint2*>(q + to_nb);
#pragma unroll
                            for (int k = 0; k < 8; ++k)
                                out[k] = (3 * out[k] + px_byte(k < 4 ? nw.x : nw.y, k & 3) + 1 + rr) >> 2;
                        }
                    }
                }
            }
            const uint2 yv = *reinterpret_cast<const uint2*>(yp + 8 * hf);
            uint32_t pk[6];
#pragma unroll
            for (int g = 0; g < 2; ++g) {
                uint32_t c0[4], c1[4], c2[4];
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    c0[k] = __byte_perm(g ? yv.y : yv.x, 0u, 0x4044 + 0x100 * k); // luma << 16
                if (kLumaOnly) {
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        c1[k] = c2[k] = c0[k];
                } else if (K.yuv) {
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        c1[k] = uint32_t(ua[4 * g + k]) << 16;
                        c2[k] = uint32_t(ub[4 * g + k]) << 16;
                    }
                } else {
                    // jdcolor.c 16 bits up: ua = Cb, ub = Cr
                    constexpr int kR = 32768 - 128 * 91881, kG = 32768 + 128 * (22554 + 46802), kB = 32768 - 128 * 116130;
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const int yk = int(c0[k]), cb = ua[4 * g + k], cr = ub[4 * g + k];
                        c0[k] = uint32_t(__viaddmin_s32_relu(91881 * cr + yk, kR, 0xFFFFFF));
                        c1[k] = uint32_t(__viaddmin_s32_relu(-46802 * cr + (-22554 * cb + yk), kG, 0xFFFFFF));
                        c2[k] = uint32_t(__viaddmin_s32_relu(116130 * cb + yk, kB, 0xFFFFFF));
                    }
                }
                if (K.planar) {
                    const uint32_t p0w = pack_b2(c0[0], c0[1], c0[2], c0[3]), p2w = pack_b2(c2[0], c2[1], c2[2], c2[3]);
                    pk[g] = K.bgr ? p2w : p0w;
                    pk[2 + g] = pack_b2(c1[0], c1[1], c1[2], c1[3]);
                    pk[4 + g] = K.bgr ? p0w : p2w;
                } else if (K.bgr) {
                    pk[3 * g + 0] = pack_b2(c2[0], c1[0], c0[0], c2[1]);
                    pk[3 * g + 1] = pack_b2(c1[1], c0[1], c2[2], c1[2]);
                    pk[3 * g + 2] = pack_b2(c0[2], c2[3], c1[3], c0[3]);
                } else {
                    pk[3 * g + 0] = pack_b2(c0[0], c1[0], c2[0], c0[1]);
                    pk[3 * g + 1] = pack_b2(c1[1], c2[1], c0[2], c1[2]);
                    pk[3 * g + 2] = pack_b2(c2[2], c0[3], c1[3], c2[3]);
                }
            }
            px_store_half(im, K, orow, X - roi_x0, hf, pk, x_lo, x_hi);
        }
    }
}

// The same colour stage over component planes in global memory (the color kernel: images that do not take the fused
// kernel, and the pixel stage when fused_pixels is off).  A thread takes 16 pixels of `n_rows` consecutive unit rows
// starting at unit row i_first; neighbours are clamped to the component's real size here (the planes have no aprons).
struct PxPlanes {
    const uint8_t* y;
    const uint8_t* cb;
    const uint8_t* cr;
    int ystride, cstride;
    int dw, dh; // real chroma size
};

template <int HX, int VX, bool kFancy, bool kLumaOnly, bool kPlainRgb>
__device__ __forceinline__ void px_color_global_impl(const PxOut& im, const PxPlanes& P, const PxColor& Kin, int X0, int i_first, int n_rows, int y_hi)
{
    PxColor K = Kin;
    if (kPlainRgb) { // interleaved R, G, B: the layout branches fold away
        K.bgr = 0;
        K.yuv = 0;
        K.planar = 0;
        K.nch = 3;
    }
    constexpr bool fancy_v = kFancy && VX == 2;
    const int roi_x0 = im.roi_x0, roi_x1 = im.roi_x1, roi_y0 = im.roi_y0;
    if (X0 >= roi_x1 || X0 + kPxUnit <= roi_x0)
        return;
#pragma unroll 1
    for (int row = 0; row < n_rows * VX; ++row) {
        const int y = VX * i_first + row;
        if (y < roi_y0 || y >= y_hi)
            continue;
        const int i = VX == 2 ? (y >> 1) : y, rr = VX == 2 ? (y & 1) : 0;
        int nb = rr == 0 ? i - 1 : i + 1; // vertical neighbour, replicated at the component's top and bottom
        nb = nb < 0 ? 0 : (nb > P.dh - 1 ? P.dh - 1 : nb);
        const uint8_t* yp = P.y + size_t(y) * P.ystride + X0;
        const uint8_t* ca = kLumaOnly ? P.y : P.cb + size_t(i) * P.cstride + X0 / HX;
        const ptrdiff_t a_to_b = kLumaOnly ? 0 : P.cr - P.cb;
        const ptrdiff_t to_nb = fancy_v ? ptrdiff_t(nb - i) * P.cstride : 0;
        uint8_t* const orow = im.out + size_t(y - roi_y0) * im.row_stride;
#pragma unroll
        for (int hf = 0; hf < 2; ++hf) {
            const int X = X0 + 8 * hf;
            if (X >= roi_x1 || X + 8 <= roi_x0)
                continue;
            const int x_lo = max(roi_x0 - X, 0), x_hi = min(roi_x1 - X, 8);
            int ua[8], ub[8];
            if (!kLumaOnly) {
                const int j = X / HX; // first chroma sample of this half
#pragma unroll
                for (int ab = 0; ab < 2; ++ab) {
                    const uint8_t* q = ca + (ab ? a_to_b : 0) + hf * (8 / HX);
                    int* out = ab ? ub : ua;
                    if (HX == 2) {
                        int s[6];
                        const int lo = j > 0 ? -1 : 0, hi = j + 4 < P.dw ? 4 : P.dw - 1 - j; // where the filter replicates
                        const uint32_t cw = __ldg(reinterpret_cast<const uint32_t*>(q));
#pragma unroll
                        for (int k = 0; k < 4; ++k)
                            s[1 + k] = px_byte(cw, k);
                        if (kFancy) {
                            s[0] = __ldg(q + lo);
                            s[5] = __ldg(q + hi);
                        }
                        if (fancy_v) {
                            const uint32_t nw = __ldg(reinterpret_cast<const uint32_t*>(q + to_nb));
                            s[0] = 3 * s[0] + __ldg(q + to_nb + lo);
                            s[5] = 3 * s[5] + __ldg(q + to_nb + hi);
#pragma unroll
                            for (int k = 0; k < 4; ++k)
                                s[1 + k] = 3 * s[1 + k] + px_byte(nw, k);
                        }
                        if (kFancy && j + 3 > P.dw - 1) { // the right image edge lies inside these samples: the last real one repeats
                            int last = s[1];
#pragma unroll
                            for (int k = 1; k < 4; ++k)
                                if (j + k <= P.dw - 1)
                                    last = s[1 + k];
#pragma unroll
                            for (int k = 0; k < 4; ++k)
                                if (j + k > P.dw - 1)
                                    s[1 + k] = last;
                        }
#pragma unroll
                        for (int k = 0; k < 4; ++k) {
                            if (fancy_v) {
                                out[2 * k] = (3 * s[k + 1] + 8 + s[k]) >> 4;
                                out[2 * k + 1] = (3 * s[k + 1] + 7 + s[k + 2]) >> 4;
                            } else if (kFancy) {
                                out[2 * k] = (3 * s[k + 1] + 1 + s[k]) >> 2;
                                out[2 * k + 1] = (3 * s[k + 1] + 2 + s[k + 2]) >> 2;
                            } else {
                                out[2 * k] = out[2 * k + 1] = s[k + 1];
                            }
                        }
                    } else {
                        const uint2 cw = __ldg(reinterpret_cast<const uint2*>(q));
#pragma unroll
                        for (int k = 0; k < 8; ++k)
                            out[k] = px_byte(k < 4 ? cw.x : cw.y, k & 3);
                        if (fancy_v) {
                            const uint2 nw = __ldg(reinterpret_cast<const uint2*>(q + to_nb));
#pragma unroll
                            for (int k = 0; k < 8; ++k)
                                out[k] = (3 * out[k] + px_byte(k < 4 ? nw.x : nw.y, k & 3) + 1 + rr) >> 2;
                        }
                    }
                }
            }
            const uint2 yv = __ldg(reinterpret_cast<const uint2*>(yp + 8 * hf));
            uint32_t pk[6];
#pragma unroll
            for (int g = 0; g < 2; ++g) {
                uint32_t c0[4], c1[4], c2[4];
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    c0[k] = __byte_perm(g ? yv.y : yv.x, 0u, 0x4044 + 0x100 * k);
                if (kLumaOnly) {
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        c1[k] = c2[k] = c0[k];
                } else if (K.yuv) {
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        c1[k] = uint32_t(ua[4 * g + k]) << 16;
                        c2[k] = uint32_t(ub[4 * g + k]) << 16;
                    }
                } else {
                    constexpr int kR = 32768 - 128 * 91881, kG = 32768 + 128 * (22554 + 46802), kB = 32768 - 128 * 116130;
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const int yk = int(c0[k]), cb = ua[4 * g + k], cr = ub[4 * g + k];
                        c0[k] = uint32_t(__viaddmin_s32_relu(91881 * cr + yk, kR, 0xFFFFFF));
                        c1[k] = uint32_t(__viaddmin_s32_relu(-46802 * cr + (-22554 * cb + yk), kG, 0xFFFFFF));
                        c2[k] = uint32_t(__viaddmin_s32_relu(116130 * cb + yk, kB, 0xFFFFFF));
                    }
                }
                if (K.planar) {
                    const uint32_t p0w = pack_b2(c0[0], c0[1], c0[2], c0[3]), p2w = pack_b2(c2[0], c2[1], c2[2], c2[3]);
                    pk[g] = K.bgr ? p2w : p0w;
                    pk[2 + g] = pack_b2(c1[0], c1[1], c1[2], c1[3]);
                    pk[4 + g] = K.bgr ? p0w : p2w;
                } else if (K.bgr) {
                    pk[3 * g + 0] = pack_b2(c2[0], c1[0], c0[0], c2[1]);
                    pk[3 * g + 1] = pack_b2(c1[1], c0[1], c2[2], c1[2]);
                    pk[3 * g + 2] = pack_b2(c0[2], c2[3], c1[3], c0[3]);
                } else {
                    pk[3 * g + 0] = pack_b2(c0[0], c1[0], c2[0], c0[1]);
                    pk[3 * g + 1] = pack_b2(c1[1], c2[1], c0[2], c1[2]);
                    pk[3 * g + 2] = pack_b2(c2[2], c0[3], c1[3], c2[3]);
                }
            }
            px_store_half(im, K, orow, X - roi_x0, hf, pk, x_lo, x_hi);
        }
    }
}

template <int HX, int VX, bool kFancy, bool kLumaOnly>
__device__ __forceinline__ void px_color_global(const PxOut& im, const PxPlanes& P, const PxColor& K, int X0, int i_first, int n_rows, int y_hi)
{
    if (!K.bgr && !K.yuv && !K.planar)
        px_color_global_impl<HX, VX, kFancy, kLumaOnly, true>(im, P, K, X0, i_first, n_rows, y_hi);
    else
        px_color_global_impl<HX, VX, kFancy, kLumaOnly, false>(im, P, K, X0, i_first, n_rows, y_hi);
}

// After phase A, before phase B: the samples the horizontal triangle filter replicates at the left and right edge of the
// image go into the aprons of the chroma rows written in this step (column -1 = column 0, column dw = column dw - 1).
__device__ __forceinline__ void px_edge_phase(const PxTile& T, uint8_t* smem, int dw, int first_row, int n_rows)
{
    for (int t = threadIdx.x & 31; t < 2 * n_rows * 2; t += 32) {
        const int c = 1 + (t & 1), side = (t >> 1) & 1, row = first_row + (t >> 2);
        const PxComp& pc = T.c[c];
        uint8_t* q = smem + pc.off + row * pc.stride + kPxApron;
        const int xorg = pc.bx0 * 8;
        if (side == 0) {
            if (xorg == 0)
                q[-1] = q[0];
        } else if (dw - xorg >= 1 && dw - xorg <= pc.nbx * 8) {
            q[dw - xorg] = q[dw - 1 - xorg];
        }
    }
}

constexpr int kPxWarps = kPxThreads / 32;

__global__ void __launch_bounds__(kPxThreads, kPxCtasPerSm) pixels_kernel(BatchDev bd, int redo_only)
{
    extern __shared__ uint4 s_px_raw[];
    __shared__ uint32_t s_qz_all[kPxWarps][3][32];
    __shared__ PxTile s_tile_all[kPxWarps];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t task = blockIdx.x * kPxWarps + warp; // one tile per warp
    if (task >= uint32_t(bd.n_px_task))
        return;
    uint8_t* smem = reinterpret_cast<uint8_t*>(s_px_raw) + size_t(warp) * bd.px_smem;
    uint32_t (*s_qz)[32] = s_qz_all[warp];
    const uint32_t img = bd.px_cta_img[task];
    const ImageDev& im = bd.img[img];
    const uint32_t flags = bd.state[img].flags;
    if ((flags & (kFlagDirty | kFlagBadMarker)) || ((redo_only & 1) && !(flags & kFlagRedo)))
        return;
    { // quantisers paired in zig-zag order: q(2j) | q(2j+1) << 24
        constexpr uint8_t kNatural[64] = {0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6, 7,
            14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62,
            63};
#pragma unroll
        for (int c = 0; c < 3; ++c)
            s_qz[c][lane] = uint32_t(im.qt[c][kNatural[2 * lane]]) | (uint32_t(im.qt[c][kNatural[2 * lane + 1]]) << 24);
    }
    // ---- tile geometry (uniform over the warp; lives in shared memory: the phases index it dynamically)
    if (lane == 0) {
        PxTile& T = s_tile_all[warp];
        const int local = int(task - im.px_cta_off);
        const int band = local / int(im.px_n_strips), strip = local - band * int(im.px_n_strips);
        const int mode = im.color_mode; // 1: luma only, 2..5: (hx, vx) = (1,1), (2,1), (1,2), (2,2) replicated, 6..8: 3..5 with the triangle filter
        const int vx = (mode == 4 || mode == 5 || mode == 7 || mode == 8) ? 2 : 1;
        const bool fancy_h = (mode == 6 || mode == 8), fancy_v = (mode == 7 || mode == 8);
        T.ncomp = (mode == 1) ? 1 : 3;
        T.mx0 = int(im.px_mcu_col_lo) + strip * int(im.px_strip_mcus);
        T.mx1 = min(T.mx0 + int(im.px_strip_mcus), int(im.px_mcu_col_hi));
        T.r_lo = int(im.px_mcu_row_lo) + band * int(im.px_band_rows);
        T.r_hi = min(T.r_lo + int(im.px_band_rows), int(im.px_mcu_row_hi));
        T.lag = fancy_v ? 1 : 0;
        T.carry = 2 * T.lag;
        T.urows = (8 * int(im.vmax)) / vx;
        int off = 0;
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            PxComp& pc = T.c[c];
            const int cc = c < T.ncomp ? c : 0;
            pc.lh = 31 - __clz(int(im.h[cc]));
            pc.lv = 31 - __clz(int(im.v[cc]));
            pc.first = im.blk_first[cc];
            const int m0 = (c > 0 && fancy_h) ? max(T.mx0 - 1, 0) : T.mx0;
            const int m1 = (c > 0 && fancy_h) ? min(T.mx1 + 1, int(im.mcus_x)) : T.mx1;
            pc.bx0 = m0 << pc.lh;
            pc.nbx = (m1 - m0) << pc.lh;
            pc.stride = ((pc.nbx * 8 + 15) & ~15) + 2 * kPxApron; // an apron on either side: the samples the horizontal filter replicates at the image edges
            pc.off = off;
            if (c < T.ncomp)
                off += (T.carry + (8 << pc.lv)) * pc.stride;
        }
    }
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
    PxSource src;
    src.coef = bd.coef;
    src.blkrec = bd.blkrec;
    src.blk0 = size_t(im.coef_off / 64);
    src.mcus_x = im.mcus_x;
    src.bpm = im.bpm;
    const int color_mode = im.color_mode, mcus_y = im.mcus_y;
    const bool fancy_h_mode = color_mode == 6 || color_mode == 8;
    const uint32_t smem_addr32 = smem_addr(smem);
    __syncwarp();
    const PxTile& T = s_tile_all[warp];
    const int mcu_h = 8 * int(im.vmax);
    const int y_lo = max(T.r_lo * mcu_h, int(im.roi_y0)), y_hi = min(T.r_hi * mcu_h, int(im.roi_y1));
    const int r_first = (T.lag && T.r_lo > 0) ? T.r_lo - 1 : T.r_lo;
    // with the lag the last unit row of the band is still pending after its last MCU row: one more step, which
    // transforms the chroma row below it (none at the bottom of the image: the filter replicates there)
    const int r_last = T.lag ? T.r_hi : T.r_hi - 1;
#pragma unroll 1
    for (int r = r_first; r <= r_last; ++r) {
        const int kind = r < T.r_lo ? 1 : (r >= T.r_hi ? 2 : 0);
        if (r < mcus_y && !(redo_only & 4))
            px_idct_phase(src, T, s_qz, smem_addr32, r, kind);
        __syncwarp();
        if (fancy_h_mode && r < mcus_y) {
            // rows written in this step: all of them, or only the last / the first chroma row (steps above / below the band)
            const int rows_c = 8 << T.c[1].lv;
            px_edge_phase(T, smem, po.dw, T.carry + (kind == 1 ? rows_c - 1 : 0), kind == 0 ? rows_c : 1);
            __syncwarp();
        }
        if (kind != 1 && !(redo_only & 2)) {
            switch (color_mode) {
            case 1: px_color_phase<1, 1, false, true>(po, T, K, smem, r, y_lo, y_hi); break;
            case 2: px_color_phase<1, 1, false, false>(po, T, K, smem, r, y_lo, y_hi); break;
            case 3: px_color_phase<2, 1, false, false>(po, T, K, smem, r, y_lo, y_hi); break;
            case 4: px_color_phase<1, 2, false, false>(po, T, K, smem, r, y_lo, y_hi); break;
            case 5: px_color_phase<2, 2, false, false>(po, T, K, smem, r, y_lo, y_hi); break;
            case 6: px_color_phase<2, 1, true, false>(po, T, K, smem, r, y_lo, y_hi); break;
            case 7: px_color_phase<1, 2, true, false>(po, T, K, smem, r, y_lo, y_hi); break;
            default: px_color_phase<2, 2, true, false>(po, T, K, smem, r, y_lo, y_hi); break;
            }
        }
        if (T.lag && r < r_last) {
            // the last two rows of every plane become the first two of the next step
            __syncwarp();
            for (int c = 0; c < T.ncomp; ++c) {
                const PxComp& pc = T.c[c];
                const int rows = T.carry + (8 << pc.lv);
                const int words = pc.stride / 4;
                for (int k = lane; k < 2 * words; k += 32) {
                    const int row = k >= words ? 1 : 0, w = k - row * words;
                    reinterpret_cast<uint32_t*>(smem + pc.off + row * pc.stride)[w] =
                        reinterpret_cast<const uint32_t*>(smem + pc.off + (rows - 2 + row) * pc.stride)[w];
                }
            }
        }
        __syncwarp();
    }
}
