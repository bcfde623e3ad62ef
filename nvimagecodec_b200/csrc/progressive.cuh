// Progressive (SOF2) scans on the device.  Included by kernels.cu inside its anonymous namespace.
//
// What is decoded: libjpeg's jdphuff.c semantics (decode_mcu_DC_first / _AC_first / _DC_refine / _AC_refine,
// process_restart), which the reference reaches through extensions/libjpeg_turbo/jpeg_mem.cpp:154-219; the scans of a
// stream accumulate into dense coefficient blocks (BatchDev::pcoef), and once the last scan is through, a record pass
// gives the IDCT stage the same per-block records the sequential path has (group mask + DC).
//
// How.  A refinement scan cannot be cut into self-synchronising subsequences: how many correction bits follow a symbol
// depends on which coefficients of the CURRENT block the earlier scans left non-zero, so a decoder that does not know
// its absolute block index cannot even parse.  The unit of parallel work is therefore the chain -- one restart interval
// of one scan (the whole scan when there are no restarts) -- decoded front to back by ONE WARP whose lanes all run the
// same parse (same bit window, same state: no divergence, table lookups broadcast from shared memory) and share the
// per-block work: lane l owns the coefficients at zig-zag positions l and l + 32, so the history mask of a block is two
// ballots, a run of r zero-history positions is found with bit operations on that mask, and all correction bits of a
// symbol (or of a whole block inside an end-of-band run) are applied in one step.  Chains of different images, of
// different scans of one image that touch disjoint coefficients (host: ProgScan::wave), and of different restart
// intervals run side by side; waves are separate launches on the batch's stream.
//
// The scans' bytes went through K0 like any scan (unstuffed big-endian words, restart boundaries in BatchDev::bnd).
#pragma once

// ---- bit reader over the unstuffed words of a scan: warp-uniform
struct PBits {
    const uint32_t* w;   // words of the scan
    uint32_t wmax;       // last word index that may be read
    uint32_t p;          // bit position
    uint32_t base;       // word index of w0
    uint32_t w0, w1, w2;
};

__device__ __forceinline__ uint32_t pb_word(const PBits& b, uint32_t i) { return __ldg(b.w + min(i, b.wmax)); }

__device__ __forceinline__ void pb_seek(PBits& b, uint32_t p)
{
    b.p = p;
    b.base = p >> 5;
    b.w0 = pb_word(b, b.base);
    b.w1 = pb_word(b, b.base + 1u);
    b.w2 = pb_word(b, b.base + 2u);
}

// the 32 bits at the current position
__device__ __forceinline__ uint32_t pb_peek(PBits& b)
{
    while ((b.p >> 5) != b.base) { // at most two steps: never more than 32 bits are consumed between two peeks
        ++b.base;
        b.w0 = b.w1;
        b.w1 = b.w2;
        b.w2 = pb_word(b, b.base + 2u);
    }
    return __funnelshift_l(b.w1, b.w0, b.p & 31u);
}

__device__ __forceinline__ uint32_t pb_get(PBits& b, uint32_t n) // n <= 32
{
    if (n == 0u)
        return 0u;
    const uint32_t v = pb_peek(b) >> (32u - n);
    b.p += n;
    return v;
}

// One Huffman symbol.  Codes of up to 9 bits come out of the first-level table, longer ones out of the canonical search.
__device__ __forceinline__ uint32_t pb_symbol(PBits& b, const ProgHuffDev& t, uint32_t& err)
{
    const uint32_t win = pb_peek(b);
    const uint32_t e = t.lut[win >> 23];
    if (e) {
        b.p += e >> 8;
        return e & 0xFFu;
    }
#pragma unroll 1
    for (int l = 10; l <= 16; ++l) {
        const int32_t code = int32_t(win >> (32 - l));
        if (code <= t.maxcode[l]) {
            b.p += uint32_t(l);
            return t.vals[(t.valoff[l] + code) & 0xFF];
        }
    }
    err |= kFlagBadHuffman;
    b.p += 16u;
    return 0u;
}

__device__ __forceinline__ int32_t pb_extend(uint32_t r, uint32_t s) // T.81 F.2.2.1 EXTEND
{
    return s == 0u ? 0 : (r >= (1u << (s - 1u)) ? int32_t(r) : int32_t(r) - int32_t((1u << s) - 1u));
}

// Block m of a one-component scan (raster order over the component's own wb x hb blocks) -> its place in the image's
// MCU-order block array.
struct PMap {
    uint32_t wb, h, v, mcus_x, bpm, first;
};
__device__ __forceinline__ uint32_t pmap_block(const PMap& g, uint32_t m)
{
    const uint32_t by = m / g.wb, bx = m - by * g.wb;
    const uint32_t my = by / g.v, mx = bx / g.h;
    return (my * g.mcus_x + mx) * g.bpm + g.first + (by - my * g.v) * g.h + (bx - mx * g.h);
}

// ---- the four scan types.  Each decodes MCUs [m0, m1) of its scan; everything is warp-uniform except the stores.

// DC, first pass (decode_mcu_DC_first): one Huffman symbol + magnitude bits per block, DPCM per component, value << Al.
__device__ __forceinline__ void prog_dc_first(PBits& b, const ProgHuffDev* tab, const ProgScanDev& sc, const ImageDev& im, const PMap& g,
    int16_t* coef, uint32_t m0, uint32_t m1, int lane, uint32_t& err)
{
    int32_t pred[4] = {0, 0, 0, 0};
#pragma unroll 1
    for (uint32_t m = m0; m < m1; ++m) {
        uint32_t blk = sc.ncomp == 1 ? pmap_block(g, m) : m * im.bpm;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            if (i >= sc.ncomp)
                break;
            const int nb = sc.ncomp == 1 ? 1 : im.h[sc.comp[i]] * im.v[sc.comp[i]];
#pragma unroll 1
            for (int j = 0; j < nb; ++j, ++blk) {
                const uint32_t s = pb_symbol(b, tab[i], err);
                if (s > 15u)
                    err |= kFlagBadHuffman;
                pred[i] += pb_extend(pb_get(b, s & 15u), s & 15u);
                if (lane == 0)
                    coef[size_t(blk) * 64] = int16_t(pred[i] << sc.al);
            }
        }
    }
}

// DC, refinement (decode_mcu_DC_refine): one raw bit per block; 32 blocks per step, one per lane.
__device__ __forceinline__ void prog_dc_refine(PBits& b, const ProgScanDev& sc, const ImageDev& im, const PMap& g, int16_t* coef, uint32_t m0,
    uint32_t m1, int lane)
{
    const uint32_t per = sc.ncomp == 1 ? 1u : im.bpm;
    const uint32_t nb = (m1 - m0) * per;
    const int16_t p1 = int16_t(1 << sc.al);
#pragma unroll 1
    for (uint32_t j0 = 0; j0 < nb; j0 += 32u) {
        const uint32_t win = pb_peek(b);
        const uint32_t j = j0 + uint32_t(lane);
        if (j < nb && ((win >> (31 - lane)) & 1u)) {
            const uint32_t blk = sc.ncomp == 1 ? pmap_block(g, m0 + j) : m0 * per + j;
            coef[size_t(blk) * 64] |= p1;
        }
        b.p += min(32u, nb - j0);
    }
}

// AC, first pass (decode_mcu_AC_first): run/size symbols inside the band, end-of-band runs over whole blocks.
__device__ __forceinline__ void prog_ac_first(PBits& b, const ProgHuffDev& tab, const ProgScanDev& sc, const PMap& g, int16_t* coef, uint32_t m0,
    uint32_t m1, int lane, uint32_t& err)
{
    uint32_t eobrun = 0;
    const uint32_t ss = sc.ss, se = sc.se;
#pragma unroll 1
    for (uint32_t m = m0; m < m1; ++m) {
        if (eobrun) { // the band of these blocks stays zero
            const uint32_t skip = min(eobrun, m1 - m);
            eobrun -= skip;
            m += skip - 1u;
            continue;
        }
        int16_t* cb = coef + size_t(pmap_block(g, m)) * 64;
        uint32_t k = ss;
#pragma unroll 1
        while (k <= se) {
            const uint32_t sym = pb_symbol(b, tab, err);
            const uint32_t r = sym >> 4, s = sym & 15u;
            if (s) {
                k += r;
                const int32_t v = pb_extend(pb_get(b, s), s);
                if (k > se) {
                    err |= kFlagOverrun;
                    break;
                }
                if (lane == 0)
                    cb[k] = int16_t(v << sc.al);
                ++k;
            } else if (r == 15u) {
                k += 16u;
            } else {
                eobrun = (1u << r) + pb_get(b, r) - 1u; // this block is the first of the run
                break;
            }
        }
    }
}

// Correction bits (decode_mcu_AC_refine, the `if (*thiscoef != 0)` arm) for the positions of set C, lowest position first:
// a set bit moves the coefficient one step (1 << Al) away from zero unless its bit Al is already set.  lo / hi: this
// lane's coefficients at positions lane and lane + 32.
__device__ __forceinline__ void prog_correct(PBits& b, uint64_t C, int32_t& lo, int32_t& hi, bool& dlo, bool& dhi, int lane, int32_t p1)
{
    const uint32_t clo = uint32_t(C), chi = uint32_t(C >> 32);
    const uint32_t n = uint32_t(__popc(clo) + __popc(chi));
    if (n == 0u)
        return;
    const uint32_t n1 = min(n, 32u);
    const uint32_t v1 = pb_get(b, n1) << (32u - n1); // first bit read = bit 31
    const uint32_t v2 = n > 32u ? pb_get(b, n - 32u) << (64u - n) : 0u;
    const uint32_t below = (1u << lane) - 1u;
    auto bit_of = [&](uint32_t idx) { return idx < 32u ? (v1 >> (31u - idx)) & 1u : (v2 >> (63u - idx)) & 1u; };
    if ((clo >> lane) & 1u) {
        if (bit_of(uint32_t(__popc(clo & below))) && (lo & p1) == 0) {
            lo += lo >= 0 ? p1 : -p1;
            dlo = true;
        }
    }
    if ((chi >> lane) & 1u) {
        if (bit_of(uint32_t(__popc(clo) + __popc(chi & below))) && (hi & p1) == 0) {
            hi += hi >= 0 ? p1 : -p1;
            dhi = true;
        }
    }
}

// AC, refinement (decode_mcu_AC_refine).
__device__ __forceinline__ void prog_ac_refine(PBits& b, const ProgHuffDev& tab, const ProgScanDev& sc, const PMap& g, int16_t* coef, uint32_t m0,
    uint32_t m1, int lane, uint32_t& err)
{
    uint32_t eobrun = 0;
    const uint32_t ss = sc.ss, se = sc.se;
    const int32_t p1 = 1 << sc.al;
    const uint64_t band = (se == 63u ? ~uint64_t(0) : ((uint64_t(1) << (se + 1u)) - 1u)) & ~((uint64_t(1) << ss) - 1u);
    // the block after this one is requested while this one is worked on
    int16_t* cb = coef + size_t(pmap_block(g, m0)) * 64;
    int32_t nlo = cb[lane], nhi = cb[lane + 32];
#pragma unroll 1
    for (uint32_t m = m0; m < m1; ++m) {
        int32_t lo = nlo, hi = nhi;
        int16_t* cur = cb;
        if (m + 1u < m1) {
            cb = coef + size_t(pmap_block(g, m + 1u)) * 64;
            nlo = cb[lane];
            nhi = cb[lane + 32];
        }
        // history: which coefficients of the band the earlier scans left non-zero
        const uint64_t nz = (uint64_t(__ballot_sync(0xffffffffu, lo != 0)) | (uint64_t(__ballot_sync(0xffffffffu, hi != 0)) << 32)) & band;
        bool dlo = false, dhi = false; // only what this scan changed is stored: other bands may be in other scans' hands
        uint32_t k = ss;
        if (eobrun == 0u) {
#pragma unroll 1
            while (k <= se) {
                const uint32_t sym = pb_symbol(b, tab, err);
                uint32_t r = sym >> 4;
                const uint32_t s = sym & 15u;
                int32_t newv = 0;
                if (s) {
                    if (s != 1u)
                        err |= kFlagBadHuffman; // (libjpeg warns and goes on)
                    newv = pb_get(b, 1u) ? p1 : -p1;
                } else if (r != 15u) {
                    eobrun = (1u << r) + pb_get(b, r);
                    break; // the rest of the band: correction bits only, below
                }
                // pass r zero-history positions from k on and stop at the next one (position t, se + 1 when the band ends
                // first); the non-zero ones passed on the way take a correction bit each
                const uint64_t from_k = band & ~((uint64_t(1) << k) - 1u);
                uint64_t z = ~nz & from_k;
#pragma unroll 1
                for (; r && z; --r)
                    z &= z - 1u;
                const uint32_t t = z ? uint32_t(__ffsll((long long)z)) - 1u : se + 1u;
                const uint64_t upto_t = t >= 64u ? ~uint64_t(0) : (uint64_t(1) << t) - 1u;
                prog_correct(b, nz & from_k & upto_t, lo, hi, dlo, dhi, lane, p1);
                if (newv) {
                    if (t > se) {
                        err |= kFlagOverrun;
                        break;
                    }
                    if (uint32_t(lane) == (t & 31u)) {
                        if (t < 32u) {
                            lo = newv;
                            dlo = true;
                        } else {
                            hi = newv;
                            dhi = true;
                        }
                    }
                }
                k = t + 1u;
            }
        }
        if (eobrun > 0u) {
            if (k <= se)
                prog_correct(b, nz & ~((uint64_t(1) << k) - 1u), lo, hi, dlo, dhi, lane, p1);
            --eobrun;
        }
        if (dlo)
            cur[lane] = int16_t(lo);
        if (dhi)
            cur[lane + 32] = int16_t(hi);
    }
}

// One warp per chain.
__global__ void __launch_bounds__(32) prog_scan_kernel(BatchDev bd, uint32_t chain0)
{
    __shared__ ProgHuffDev s_tab[4];
    const int lane = threadIdx.x;
    const uint32_t chain = chain0 + blockIdx.x;
    const ProgScanDev& sc = bd.prog_scan[bd.prog_chain[chain]];
    const ImageDev& im = bd.img[sc.parent];
    const ImageDev& ch = bd.img[sc.child];
    const ImageState cst = bd.state[sc.child];
    if (cst.flags) { // K0 found the restart markers of this scan out of order / miscounted
        if (lane == 0)
            atomicOr(&bd.state[sc.parent].flags, cst.flags);
        return;
    }
    const bool huffman = !(sc.ss == 0 && sc.ah != 0);
    if (huffman) {
        for (int i = 0; i < sc.ncomp; ++i) {
            const uint4* src = reinterpret_cast<const uint4*>(bd.prog_huff + sc.table[i]);
            uint4* dst = reinterpret_cast<uint4*>(&s_tab[i]);
            for (int j = lane; j < int(sizeof(ProgHuffDev) / 16); j += 32)
                dst[j] = __ldg(src + j);
        }
    }
    __syncwarp();
    // the chain's MCUs and bytes
    const uint32_t j = chain - sc.chain_off;
    const uint32_t m0 = sc.ri ? j * sc.ri : 0u, m1 = sc.ri ? min(sc.n_mcu, m0 + sc.ri) : sc.n_mcu;
    const uint32_t byte0 = j ? bd.bnd[ch.bnd_off + j - 1u] : 0u;
    const uint32_t byte1 = j + 1u < sc.n_chain ? bd.bnd[ch.bnd_off + j] : cst.cmp_len;
    PBits b;
    b.w = bd.cmp + ch.cmp_off / 4;
    b.wmax = (ch.raw_len + 15u) / 4u; // inside what the host reserved for this scan (the words past cmp_len are never trusted)
    pb_seek(b, byte0 * 8u);
    PMap g;
    const int c0 = sc.comp[0];
    g.wb = max(1u, uint32_t(sc.wb));
    g.h = im.h[c0];
    g.v = im.v[c0];
    g.mcus_x = im.mcus_x;
    g.bpm = im.bpm;
    g.first = im.blk_first[c0];
    int16_t* coef = bd.pcoef + im.pcoef_off;
    uint32_t err = 0;
    if (sc.ss == 0 && sc.ah == 0)
        prog_dc_first(b, s_tab, sc, im, g, coef, m0, m1, lane, err);
    else if (sc.ss == 0)
        prog_dc_refine(b, sc, im, g, coef, m0, m1, lane);
    else if (sc.ah == 0)
        prog_ac_first(b, s_tab[0], sc, g, coef, m0, m1, lane, err);
    else
        prog_ac_refine(b, s_tab[0], sc, g, coef, m0, m1, lane, err);
    if (b.p > byte1 * 8u)
        err |= kFlagOverrun; // the chain read past its restart interval / the end of the scan
    if (err && lane == 0)
        atomicOr(&bd.state[sc.parent].flags, err);
}

// After the last wave: the per-block records the IDCT stage reads (bits 0-7: which 16-byte groups hold a non-zero AC
// coefficient, bits 16-31: the DC value), and the DC slot of the block cleared, as in the dense form of the sequential
// path.  Eight threads per block, one per group.
__global__ void __launch_bounds__(256) prog_record_kernel(BatchDev bd)
{
    const uint32_t img = bd.prog_img[blockIdx.y];
    const ImageDev& im = bd.img[img];
    const uint32_t t = blockIdx.x * 256u + threadIdx.x;
    const uint32_t blk = t >> 3, grp = t & 7u;
    const bool live = blk < im.total_blocks;
    uint4 v = make_uint4(0u, 0u, 0u, 0u);
    uint4* src = reinterpret_cast<uint4*>(bd.pcoef + im.pcoef_off) + size_t(blk) * 8 + grp;
    if (live)
        v = *src;
    const int32_t dc = int32_t(int16_t(v.x & 0xFFFFu)); // meaningful in group 0
    if (grp == 0u)
        v.x &= 0xFFFF0000u;
    const uint32_t any = __ballot_sync(0xffffffffu, (v.x | v.y | v.z | v.w) != 0u);
    if (live && grp == 0u) {
        const uint32_t mask = (any >> (threadIdx.x & 24u)) & 0xFFu;
        bd.blkrec[im.coef_off / 64 + blk] = mask | (uint32_t(dc) << 16);
        if (dc != 0)
            *reinterpret_cast<uint32_t*>(src) = v.x;
    }
}
