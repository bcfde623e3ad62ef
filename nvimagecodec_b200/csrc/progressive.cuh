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
// ballots, the positions of its zero-history coefficients are listed once per block in shared memory (a run of r of them
// is one read of that list), and the correction bits of a block are fetched and applied by all lanes at once.
//
// All chains of a batch run in ONE launch.  A scan that refines what an earlier scan of the same image wrote does not wait
// for that scan to end: block m of a scan depends on block m of its predecessor only, so every chain publishes how far it
// has come (BatchDev::prog_progress, release store every few blocks) and a successor chain follows it a few blocks behind
// (relaxed polls and one fence per successful wait; coefficient reads go to L2, `ld.global.cg`, since another SM wrote
// them in the same launch; the last scan of a component, which nobody follows, publishes nothing).  The ten
// scans of a typical stream thus decode side by side, and an image takes about as long as its longest scan instead of the
// sum.  Chains are numbered so that a predecessor always has the lower index (image by image, scans in file order): with
// CTAs dispatched in index order a waiting chain's predecessor is resident or finished, whatever the grid size.
//
// The scans' bytes went through K0 like any scan (unstuffed big-endian words, restart boundaries in BatchDev::bnd).
#pragma once

// ---- bit reader over the unstuffed words of a scan: warp-uniform
// The words pass through a ring of four 32-word chunks in shared memory that the warp fills a chunk at a time (one coalesced
// load, requested a whole chunk before it is needed), so that a window after a jump -- a run of correction bits was skipped --
// costs shared-memory reads, not three exposed global loads.  The ring always holds the words from (at least) 45 before to
// 51 behind the current one: enough for one block's symbols ahead and for the correction bits of the block under way behind.
struct PBits {
    const uint32_t* w;   // words of the scan
    uint32_t wmax;       // last word index that may be read
    uint32_t p;          // bit position
    uint32_t base;       // word index of w0
    uint32_t w0, w1, w2;
    uint32_t ring;       // shared-memory address of the ring
    uint32_t n_chunk;    // chunks [n_chunk - 4, n_chunk) are in the ring
    uint32_t ahead;      // this lane's word of chunk n_chunk, already requested
    uint32_t lane;
};
constexpr uint32_t kRingAhead = 51u;

__device__ __forceinline__ void pb_fill(PBits& b, uint32_t word) // the ring must hold the words up to `word`
{
    if (b.n_chunk * 32u > word)
        return;
    __syncwarp();
    do {
        asm volatile("st.shared.u32 [%0], %1;" ::"r"(b.ring + (((b.n_chunk & 3u) * 32u + b.lane) << 2)), "r"(b.ahead) : "memory");
        ++b.n_chunk;
        b.ahead = __ldg(b.w + min(b.n_chunk * 32u + b.lane, b.wmax));
    } while (b.n_chunk * 32u <= word);
    __syncwarp();
}
__device__ __forceinline__ uint32_t pb_word(const PBits& b, uint32_t i)
{
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(b.ring + ((i & 127u) << 2)));
    return v;
}

__device__ __forceinline__ void pb_seek(PBits& b, uint32_t p)
{
    b.p = p;
    b.base = p >> 5;
    pb_fill(b, b.base + kRingAhead);
    b.w0 = pb_word(b, b.base);
    b.w1 = pb_word(b, b.base + 1u);
    b.w2 = pb_word(b, b.base + 2u);
}
__device__ __forceinline__ void pb_open(PBits& b, const uint32_t* words, uint32_t wmax, uint32_t ring, uint32_t lane, uint32_t p)
{
    b.w = words;
    b.wmax = wmax;
    b.ring = ring;
    b.lane = lane;
    b.n_chunk = p >> 10; // the chunk the first word lies in
    b.ahead = __ldg(b.w + min(b.n_chunk * 32u + lane, wmax));
    pb_seek(b, p);
}

// the 32 bits at the current position
__device__ __forceinline__ uint32_t pb_peek(PBits& b)
{
    const uint32_t d = (b.p >> 5) - b.base;
    if (d) {
        if (d == 1u) { // the usual step
            ++b.base;
            pb_fill(b, b.base + kRingAhead);
            b.w0 = b.w1;
            b.w1 = b.w2;
            b.w2 = pb_word(b, b.base + 2u);
        } else {       // a run of correction bits was skipped
            pb_seek(b, b.p);
        }
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
// A table as the loops see it: the shared-space address of its first-level entries, pinned in a register (a generic pointer,
// or an address the compiler may recompute, costs a special-register read per symbol), and the table itself for long codes.
struct PTab {
    uint32_t lut;
    const ProgHuffDev* t;
};
__device__ __forceinline__ uint32_t pinned_smem_addr(const void* p)
{
    uint32_t a;
    asm volatile("mov.u32 %0, %1;" : "=r"(a) : "r"(smem_addr(p)));
    return a;
}
__device__ __forceinline__ uint32_t pb_symbol(PBits& b, const PTab& pt, uint32_t& err)
{
    const ProgHuffDev& t = *pt.t;
    const uint32_t win = pb_peek(b);
    uint32_t e;
    asm volatile("ld.shared.u16 %0, [%1];" : "=r"(e) : "r"(pt.lut + ((win >> 22) & 0x3FEu)));
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

// Blocks of a one-component scan run in raster order over the component's own wb x hb blocks; PMap turns (bx, by) into
// the block's place in the image's MCU-order array.  Sampling factors are 1, 2 or 4: shifts, no divisions in the loops.
struct PMap {
    uint32_t wb, lh, lv, mcus_x, bpm, first;
    uint32_t bx, by;
};
__device__ __forceinline__ void pmap_seek(PMap& g, uint32_t m)
{
    g.by = m / g.wb;
    g.bx = m - g.by * g.wb;
}
__device__ __forceinline__ void pmap_advance(PMap& g, uint32_t n)
{
    g.bx += n;
    if (g.bx >= g.wb) {
        if (g.bx < 2u * g.wb) {
            g.bx -= g.wb;
            ++g.by;
        } else {
            const uint32_t rows = g.bx / g.wb;
            g.by += rows;
            g.bx -= rows * g.wb;
        }
    }
}
__device__ __forceinline__ uint32_t pmap_index(const PMap& g, uint32_t bx, uint32_t by)
{
    const uint32_t mx = bx >> g.lh, my = by >> g.lv;
    return (my * g.mcus_x + mx) * g.bpm + g.first + ((by - (my << g.lv)) << g.lh) + (bx - (mx << g.lh));
}
__device__ __forceinline__ uint32_t pmap_block(const PMap& g) { return pmap_index(g, g.bx, g.by); }
__device__ __forceinline__ uint32_t pmap_block_at(const PMap& g, uint32_t m) // any block (one division: not for inner loops)
{
    const uint32_t by = m / g.wb;
    return pmap_index(g, m - by * g.wb, by);
}

// ---- progress of a chain / following a predecessor chain (warp-uniform)
struct PSync {
    const uint32_t* pred;  // progress word of the predecessor chain (nullptr: none)
    uint32_t avail;        // blocks [.., avail) of the predecessor are known to be finished
    uint32_t* mine;        // progress word of this chain (nullptr: nobody follows it)
    uint32_t published;
    bool one_writer;       // only lane 0 stores coefficients in this scan type: its release store alone orders them
};
__device__ __forceinline__ void ps_need(PSync& y, uint32_t m_excl) // the predecessor must have finished its MCUs below m_excl
{
    if (y.avail >= m_excl) // (no predecessor: avail is all ones)
        return;
    for (;;) { // relaxed polls (an acquire load empties this SM's L1 every time), one fence once the predecessor is far enough
        uint32_t v;
        asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(y.pred) : "memory");
        if (v >= m_excl) {
            asm volatile("fence.acq_rel.gpu;" ::: "memory");
            y.avail = v;
            return;
        }
        __nanosleep(200);
    }
}
__device__ __forceinline__ void ps_publish(PSync& y, uint32_t m_done, bool force, uint32_t lane)
{
    if (y.mine == nullptr || (!force && m_done < y.published + 16u)) // (nobody follows the last scan of a component)
        return;
    y.published = m_done;
    if (!y.one_writer) { // every lane stored: all of it has to be visible before the progress word says so
        __syncwarp();
        __threadfence();
    }
    if (lane == 0u)
        asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(y.mine), "r"(m_done) : "memory");
}
__device__ __forceinline__ int32_t pc_load(const int16_t* p) { return int32_t(__ldcg(p)); } // coefficients: L2, never this SM's L1
__device__ __forceinline__ void pc_store_if(bool on, int16_t* p, int32_t v) // a predicated store: no divergent branch for one lane's sake
{
    asm volatile("{\n\t.reg .pred q;\n\tsetp.ne.u32 q, %2, 0;\n\t@q st.global.u16 [%0], %1;\n\t}" ::"l"(p), "r"(v), "r"(uint32_t(on)) : "memory");
}

// ---- the four scan types.  Each decodes MCUs [m0, m1) of its scan; everything is warp-uniform except the stores.

// Everything the scan loops read is held in registers (PScan, PMap): the descriptors live in global memory, and a loop that
// stores coefficients would otherwise re-read them after every store.
struct PScan {
    int ncomp, ss, se, al;
    int nb[4];   // interleaved scan: blocks per MCU of each scan component
    uint32_t bpm;
};

// DC, first pass (decode_mcu_DC_first): one Huffman symbol + magnitude bits per block, DPCM per component, value << Al.
__device__ __forceinline__ void prog_dc_first(PBits& b, const PTab* tab, const PScan& sc, PMap& g, int16_t* coef, uint32_t m0, uint32_t m1,
    int lane, uint32_t& err, PSync& y)
{
    int32_t pred[4] = {0, 0, 0, 0};
    pmap_seek(g, m0);
#pragma unroll 1
    for (uint32_t m = m0; m < m1; ++m) {
        ps_need(y, m + 1u);
        ps_publish(y, m, false, uint32_t(lane));
        uint32_t blk = m * sc.bpm;
        if (sc.ncomp == 1) {
            blk = pmap_block(g);
            pmap_advance(g, 1u);
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            if (i >= sc.ncomp)
                break;
            const int nb = sc.ncomp == 1 ? 1 : sc.nb[i];
#pragma unroll 1
            for (int j = 0; j < nb; ++j, ++blk) {
                const uint32_t s = pb_symbol(b, tab[i], err);
                if (s > 15u)
                    err |= kFlagBadHuffman;
                pred[i] += pb_extend(pb_get(b, s & 15u), s & 15u);
                pc_store_if(lane == 0, coef + size_t(blk) * 64, pred[i] << sc.al);
            }
        }
    }
}

// DC, refinement (decode_mcu_DC_refine): one raw bit per block; 32 blocks per step, one per lane.
__device__ __forceinline__ void prog_dc_refine(PBits& b, const PScan& sc, const PMap& g, int16_t* coef, uint32_t m0, uint32_t m1, int lane, PSync& y)
{
    const uint32_t per = sc.ncomp == 1 ? 1u : sc.bpm;
    const uint32_t nb = (m1 - m0) * per;
    const int16_t p1 = int16_t(1 << sc.al);
#pragma unroll 1
    for (uint32_t j0 = 0; j0 < nb; j0 += 32u) {
        ps_need(y, min(m1, m0 + (j0 + 32u + per - 1u) / per));
        ps_publish(y, m0 + j0 / per, false, uint32_t(lane));
        const uint32_t win = pb_peek(b);
        const uint32_t j = j0 + uint32_t(lane);
        if (j < nb && ((win >> (31 - lane)) & 1u)) {
            const uint32_t blk = sc.ncomp == 1 ? pmap_block_at(g, m0 + j) : m0 * per + j;
            int16_t* dc = coef + size_t(blk) * 64;
            *dc = int16_t(pc_load(dc) | p1);
        }
        b.p += min(32u, nb - j0);
    }
}

// AC, first pass (decode_mcu_AC_first): run/size symbols inside the band, end-of-band runs over whole blocks.
// The common symbols go through a pre-decoded table like those of the refinement scans (s_fl: one word per nine-bit window --
// byte 0 = code length + magnitude bits, bits 8-11 = magnitude bits, byte 3 = run; an end of band for this block only is a run
// of 64, i.e. past the band) in a loop whose recurrence is window -> table entry -> next window; codes longer than nine bits
// and end-of-band runs over several blocks (kRefSlow) take the plain path.
constexpr uint32_t kRefSlow = 1u << 20;
__device__ __forceinline__ uint32_t first_entry(uint32_t len, uint32_t sym)
{
    const uint32_t r = sym >> 4, s = sym & 15u;
    if (s == 0u && r != 15u)
        return r ? kRefSlow : (len | (64u << 24));
    return (len + s) | (s << 8) | (r << 24);
}

__device__ __forceinline__ void prog_ac_first(PBits& b, const PTab& tab, const uint32_t* s_fl, const PScan& sc, PMap& g, int16_t* coef,
    uint32_t m0, uint32_t m1, int lane, uint32_t& err, PSync& y)
{
    uint32_t eobrun = 0;
    const uint32_t ss = sc.ss, se = sc.se;
    const uint32_t fl_addr = pinned_smem_addr(s_fl);
    bool overrun = false;
    pmap_seek(g, m0);
#pragma unroll 1
    for (uint32_t m = m0; m < m1; ++m) {
        ps_publish(y, m, false, uint32_t(lane));
        if (eobrun) { // the band of these blocks stays zero
            const uint32_t skip = min(eobrun, m1 - m);
            ps_need(y, m + skip); // (nothing of them is read, but successors of this scan rely on the order)
            eobrun -= skip;
            m += skip - 1u;
            pmap_advance(g, skip);
            continue;
        }
        ps_need(y, m + 1u);
        int16_t* cb = coef + size_t(pmap_block(g)) * 64;
        pmap_advance(g, 1u);
        uint32_t k = ss;
        (void)pb_peek(b);                // window registers brought to b.p,
        pb_fill(b, b.base + kRingAhead); // and the ring far enough ahead for the whole block
        uint32_t w0 = b.w0, w1 = b.w1, w2 = b.w2, base = b.base, o = b.p & 31u;
#pragma unroll 1
        for (;;) {
            bool slow = false;
            do {
                const uint32_t win = __funnelshift_l(w1, w0, o);
                uint32_t e;
                asm volatile("ld.shared.u32 %0, [%1];" : "=r"(e) : "r"(fl_addr + ((win >> 21) & 0x7FCu)));
                if (e & kRefSlow) {
                    slow = true;
                    break;
                }
                const uint32_t total = e & 0xFFu, s = (e >> 8) & 15u;
                k += e >> 24;
                const uint32_t raw = (win >> ((32u - total) & 31u)) & ((1u << s) - 1u);
                const int32_t v = pb_extend(raw, s);
                overrun = overrun || (s != 0u && k > se);
                pc_store_if(s != 0u && k <= se && lane == 0, cb + min(k, 63u), v << sc.al);
                ++k;
                o += total;
            } while (k <= se && o < 32u);
            if (!slow) {
                if (k > se)
                    break;
                ++base; // the window moves on by a word (a symbol takes 24 bits at most)
                w0 = w1;
                w1 = w2;
                w2 = pb_word(b, base + 2u);
                o -= 32u;
                continue;
            }
            // ---- a rare symbol, the plain way
            b.p = base * 32u + o;
            b.base = base;
            b.w0 = w0;
            b.w1 = w1;
            b.w2 = w2;
            const uint32_t sym = pb_symbol(b, tab, err);
            const uint32_t r = sym >> 4, s = sym & 15u;
            if (s) {
                k += r;
                const int32_t v = pb_extend(pb_get(b, s), s);
                overrun = overrun || k > se;
                pc_store_if(k <= se && lane == 0, cb + min(k, 63u), v << sc.al);
                ++k;
            } else if (r == 15u) {
                k += 16u;
            } else {
                eobrun = (1u << r) + pb_get(b, r) - 1u; // this block is the first of the run
                k = se + 1u;
            }
            (void)pb_peek(b);
            w0 = b.w0, w1 = b.w1, w2 = b.w2, base = b.base, o = b.p & 31u;
            if (k > se)
                break;
        }
        b.p = base * 32u + o;
        b.base = base;
        b.w0 = w0;
        b.w1 = w1;
        b.w2 = w2;
    }
    if (overrun)
        err |= kFlagOverrun;
}

// ---- AC, refinement (decode_mcu_AC_refine)
// A symbol passes r zero-history positions and stops at the next one; the non-zero-history ones passed on the way take a
// correction bit each (a set bit moves the coefficient one step (1 << Al) away from zero unless its bit Al is already set).
//
// The parse is one dependent chain per symbol (window -> table -> which zero -> how many correction bits -> next
// window), so everything else is kept off it:
// * the table is pre-decoded for this scan type (refine_entry): bits consumed including the sign bit, run, kind;
// * the positions of a block's zero-history coefficients are listed once per block (s_q, written by the lanes that own
//   them): "the (j + r)-th zero of the band" is one shared-memory read, and the number of correction bits follows from
//   positions and counts alone -- the parse never loops over coefficients;
// * correction bits are not extracted while parsing.  Every non-zero-history coefficient of the band takes exactly one,
//   in position order, so the bit of the coefficient at position q sits at
//       (block start) + (code, sign and run-length bits of the symbols that started at or before q) + (rank of q among them)
//   and each lane only notes the middle term for its two positions (one compare-and-select per symbol); the bits are
//   fetched and applied once per block.
constexpr uint32_t kRefCoef = 1u << 9, kRefEob = 1u << 10, kRefBad = 1u << 11;
// the same for the common-symbol loop, two 32-bit words per nine-bit window.  x: byte 0 = code length + sign bit, byte 3 = run
// (64 for an end of band that covers this block only, EOB0: the run then stops behind the band, the rest of the band takes
// correction bits and nothing carries over), kRefSlow = not for that loop (a longer code, an end-of-band run over several
// blocks, a new coefficient of more than one bit).  y: the sign bit's place in the 32-bit window when the symbol brings a new coefficient, else 0.
constexpr uint32_t kBehind = 128u; // mark on the zero-list entries that stand for "behind the band" (positions are < 64)
// refine_entry: bits 0-4: code length + sign bit; bits 5-8: run
__device__ __forceinline__ uint32_t refine_entry(uint32_t len, uint32_t sym)
{
    const uint32_t r = sym >> 4, s = sym & 15u;
    return (len + (s ? 1u : 0u)) | (r << 5) | (s ? kRefCoef : 0u) | ((s == 0u && r != 15u) ? kRefEob : 0u) | (s > 1u ? kRefBad : 0u);
}

__device__ __forceinline__ void prog_ac_refine(PBits& b, const PTab& tab, const uint2* s_rl, const PScan& sc, PMap& g, int16_t* coef,
    uint32_t m0, uint32_t m1, uint32_t lane, uint32_t& err, uint8_t* s_q, PSync& y)
{
    uint32_t eobrun = 0;
    const uint32_t ss = sc.ss, se = sc.se;
    const int32_t p1 = 1 << sc.al;
    const uint64_t band = (se == 63u ? ~uint64_t(0) : ((uint64_t(1) << (se + 1u)) - 1u)) & ~((uint64_t(1) << ss) - 1u);
    const uint32_t bandlo = uint32_t(band), bandhi = uint32_t(band >> 32);
    const uint32_t below = (1u << lane) - 1u;
    const uint32_t rl_addr = pinned_smem_addr(s_rl), q_addr = pinned_smem_addr(s_q);
    for (uint32_t q = 64u + lane; q < 144u; q += 32u) // (the entries behind any list: what an EOB0's run of 64 reads)
        asm volatile("st.shared.u8 [%0], %1;" ::"r"(q_addr + q), "r"(se + 1u + kBehind) : "memory");
    uint32_t seen = 0; // every entry of the plain path, or-ed: the "bad" flag is looked at once, after the scan
    // the window lives in registers for the whole scan: three words from word `base` on, the position o bits into them
    // (o may run past them while correction bits are skipped; it is brought back at the top of every block)
    (void)pb_peek(b);
    uint32_t w0 = b.w0, w1 = b.w1, w2 = b.w2, base = b.base, o = b.p & 31u;
    // the block after this one is requested while this one is worked on
    pmap_seek(g, m0);
    ps_need(y, min(m1, m0 + 1u));
    int16_t* cb = coef + size_t(pmap_block(g)) * 64;
    int32_t nlo = pc_load(cb + lane), nhi = pc_load(cb + lane + 32);
#pragma unroll 1
    for (uint32_t m = m0; m < m1; ++m) {
        int32_t lo = nlo, hi = nhi;
        int16_t* cur = cb;
        ps_publish(y, m, false, lane);
        if (m + 1u < m1) {
            ps_need(y, m + 2u);
            pmap_advance(g, 1u);
            cb = coef + size_t(pmap_block(g)) * 64;
            nlo = pc_load(cb + lane);
            nhi = pc_load(cb + lane + 32);
        }
        // history: which coefficients of the band the earlier scans left non-zero
        const uint32_t nzlo = __ballot_sync(0xffffffffu, lo != 0) & bandlo, nzhi = __ballot_sync(0xffffffffu, hi != 0) & bandhi;
        const uint32_t zlo = ~nzlo & bandlo, zhi = ~nzhi & bandhi;
        const uint32_t n_zlo = uint32_t(__popc(zlo)), n_z = n_zlo + uint32_t(__popc(zhi));
        const uint32_t n_nz = uint32_t(__popc(nzlo) + __popc(nzhi));
        const int32_t lo0 = lo, hi0 = hi;
        if (o >= 32u) {
            base += o >> 5;
            o &= 31u;
            pb_fill(b, base + kRingAhead); // the ring far enough ahead for the whole block
            w0 = pb_word(b, base);
            w1 = pb_word(b, base + 1u);
            w2 = pb_word(b, base + 2u);
        } else {
            pb_fill(b, base + kRingAhead);
        }
        const uint32_t pstart = base * 32u + o;
        uint32_t clo = 0, chi = 0; // symbol bits read up to (and including) the symbol that passes this lane's positions
        uint32_t k = ss, j = 0;    // j: how many zero-history positions of the band lie before k
        if (eobrun == 0u) {
            __syncwarp();
            // (predicated stores, no divergent branches)
            asm volatile("{\n\t.reg .pred q;\n\tsetp.ne.u32 q, %2, 0;\n\t@q st.shared.u8 [%0], %1;\n\t}" ::"r"(q_addr + uint32_t(__popc(zlo & below))), "r"(lane),
                         "r"((zlo >> lane) & 1u)
                         : "memory");
            asm volatile("{\n\t.reg .pred q;\n\tsetp.ne.u32 q, %2, 0;\n\t@q st.shared.u8 [%0], %1;\n\t}" ::"r"(q_addr + n_zlo + uint32_t(__popc(zhi & below))),
                         "r"(lane + 32u), "r"((zhi >> lane) & 1u)
                         : "memory");
            // "no such zero": a run that reaches past the last one stops behind the band
            asm volatile("{\n\t.reg .pred q;\n\tsetp.lt.u32 q, %2, 16;\n\t@q st.shared.u8 [%0], %1;\n\t}" ::"r"(q_addr + n_z + lane), "r"(se + 1u + kBehind), "r"(lane)
                         : "memory");
            __syncwarp();
            uint32_t c = 0; // symbol bits of this block so far
            bool overrun = false;
#pragma unroll 1
            for (;;) {
                // ---- the common symbols: one pass of the inner loop each.  A lone warp issues an instruction every other
                // clock at best, so the loop is kept to the parse recurrence (window -> table entry -> position of the zero
                // the run stops at -> next window) and a handful of selects; whatever is rare -- a code longer than nine
                // bits, an end-of-band run over several blocks, the window moving on by a word (every eight symbols or so) --
                // leaves it with the symbol not yet consumed.
                bool slow = false;
                uint32_t t, ey; // (looked at once more behind the loop)
                int32_t lim = 0;
                do {
                    const uint32_t win = __funnelshift_l(w1, w0, o);
                    uint32_t ex; // table entry: ex = bits consumed / run / flags, ey = the sign bit's place in the window
                    asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(ex), "=r"(ey) : "r"(rl_addr + ((win >> 20) & 0xFF8u)));
                    const uint32_t r = ex >> 24, total = ex & 0xFFu;
                    // where the symbol stops: the (j + r)-th zero-history position, or kBehind + the position behind the band
                    asm volatile("ld.shared.u8 %0, [%1];" : "=r"(t) : "r"(q_addr + j + r));
                    if (ex & kRefSlow) { // (looked at while the read above is under way)
                        slow = true;
                        break;
                    }
                    const uint32_t zp = min(r, n_z - j); // zero-history positions passed
                    c += total;
                    clo = lane >= k ? c : clo;
                    chi = lane + 32u >= k ? c : chi;
                    const int32_t newv = (win & ey) ? p1 : -p1;
                    lo = (ey != 0u && lane == t) ? newv : lo; // (no lane owns a position behind the band)
                    hi = (ey != 0u && lane + 32u == t) ? newv : hi;
                    uint32_t pre = o + total - (k + zp); // (known before t arrives: one addition is left behind the read,
                    asm volatile("" : "+r"(pre));        //  and the loop test compares t itself)
                    lim = 32 - int32_t(pre);
                    o = pre + t; // += total + (positions passed that are not zero-history ones)
                    k = t + 1u;
                    j += r + 1u;
                } while (int32_t(t) < lim && t < se); // i.e. o < 32 and k <= se
                if (!slow && k <= se && o < 64u) { // the usual way out: the window moves on by a word
                    ++base;
                    w0 = w1;
                    w1 = w2;
                    w2 = pb_word(b, base + 2u);
                    o -= 32u;
                    continue;
                }
                if (!slow) {
                    const uint32_t mark = t & kBehind; // the band ended behind its last position: take the mark out again
                    o -= mark;
                    k -= mark;
                    overrun = overrun || (mark != 0u && ey != 0u); // a new coefficient with no place left for it
                    if (k > se)
                        break;
                    // a run of correction bits longer than the window registers reach was skipped (the ring reaches far enough)
                    base += o >> 5;
                    o &= 31u;
                    w0 = pb_word(b, base);
                    w1 = pb_word(b, base + 1u);
                    w2 = pb_word(b, base + 2u);
                    continue;
                }
                j = min(j, n_z);
                b.p = base * 32u + o; // (the plain path works on the reader itself)
                b.base = base;
                b.w0 = w0;
                b.w1 = w1;
                b.w2 = w2;
                // ---- a rare symbol, the plain way
                const uint32_t p0 = b.p;
                const uint32_t sym = pb_symbol(b, tab, err);
                const uint32_t e = refine_entry(b.p - p0, sym);
                const uint32_t total = e & 31u, r = (e >> 5) & 15u;
                b.p = p0 + total;
                c += total;
                if (e & kRefEob) {
                    c += r;
                    eobrun = (1u << r) + pb_get(b, r);
                    (void)pb_peek(b);
                    w0 = b.w0, w1 = b.w1, w2 = b.w2, base = b.base, o = b.p & 31u;
                    break; // the rest of the band: correction bits only
                }
                seen |= e;
                const int32_t newv = (e & kRefCoef) ? ((pb_word(b, (b.p - 1u) >> 5) >> (31u - ((b.p - 1u) & 31u))) & 1u ? p1 : -p1) : 0;
                const uint32_t idx = j + r;
                const bool found = idx < n_z;
                asm volatile("ld.shared.u8 %0, [%1];" : "=r"(t) : "r"(q_addr + min(idx, 79u)));
                t = found ? t : se + 1u;
                b.p += (t - k) - (found ? r : n_z - j);
                clo = lane >= k ? c : clo;
                chi = lane + 32u >= k ? c : chi;
                if (newv) {
                    overrun = overrun || !found;
                    lo = (lane == t && found) ? newv : lo;
                    hi = (lane + 32u == t && found) ? newv : hi;
                }
                k = t + 1u;
                j = min(idx + 1u, n_z);
                (void)pb_peek(b);
                w0 = b.w0, w1 = b.w1, w2 = b.w2, base = b.base, o = b.p & 31u;
                if (k > se)
                    break;
            }
            if (overrun)
                err |= kFlagOverrun;
            if (eobrun > 0u && k <= se) { // the band's tail behind an end-of-band symbol
                clo = lane >= k ? c : clo;
                chi = lane + 32u >= k ? c : chi;
                o += (se + 1u - k) - (n_z - min(j, n_z));
            }
        } else {
            o += n_nz; // a block inside an end-of-band run: correction bits only
        }
        if (eobrun > 0u)
            --eobrun;
        // correction bits, fetched where they lie -- straight-line code for every lane (a lone warp pays a divergent branch
        // dearly): lanes without a non-zero-history coefficient read some word of the ring and select nothing
        {
            const uint32_t at_lo = pstart + clo + uint32_t(__popc(nzlo & below));
            const uint32_t at_hi = pstart + chi + uint32_t(__popc(nzlo)) + uint32_t(__popc(nzhi & below));
            const uint32_t bit_lo = (pb_word(b, at_lo >> 5) >> (31u - (at_lo & 31u))) & 1u;
            const uint32_t bit_hi = (pb_word(b, at_hi >> 5) >> (31u - (at_hi & 31u))) & 1u;
            const bool up_lo = ((nzlo >> lane) & bit_lo) != 0u && (lo & p1) == 0;
            const bool up_hi = ((nzhi >> lane) & bit_hi) != 0u && (hi & p1) == 0;
            lo = up_lo ? lo + (lo >= 0 ? p1 : -p1) : lo;
            hi = up_hi ? hi + (hi >= 0 ? p1 : -p1) : hi;
        }
        // only what this scan changed is stored: other bands may be in other scans' hands
        asm volatile("{\n\t.reg .pred q;\n\tsetp.ne.s32 q, %1, %2;\n\t@q st.global.u16 [%0], %1;\n\t}" ::"l"(cur + lane), "r"(lo), "r"(lo0) : "memory");
        asm volatile("{\n\t.reg .pred q;\n\tsetp.ne.s32 q, %1, %2;\n\t@q st.global.u16 [%0], %1;\n\t}" ::"l"(cur + lane + 32), "r"(hi), "r"(hi0) : "memory");
    }
    b.p = base * 32u + o;
    if (seen & kRefBad)
        err |= kFlagBadHuffman; // a new coefficient of more than one bit (libjpeg warns and goes on)
}

// One warp per chain.
__global__ void __launch_bounds__(32) prog_scan_kernel(BatchDev bd)
{
    __shared__ ProgHuffDev s_tab[4];
    __shared__ uint8_t s_q[144]; // zero-history positions of the block under way; "behind the band" from there on
    __shared__ uint32_t s_ring[128]; // the scan's words around the current position (PBits)
    __shared__ uint2 s_rl[512]; // refinement scans: the first-level table pre-decoded for the common-symbol loop
    const int lane = threadIdx.x;
    const uint32_t chain = blockIdx.x;
    const ProgScanDev sd = bd.prog_scan[bd.prog_chain[chain]];
    const ImageDev& im = bd.img[sd.parent];
    const ImageDev& ch = bd.img[sd.child];
    const ImageState cst = bd.state[sd.child];
    // the chain's MCUs
    const uint32_t j = chain - sd.chain_off;
    const uint32_t m0 = sd.ri ? j * sd.ri : 0u, m1 = sd.ri ? min(sd.n_mcu, m0 + sd.ri) : sd.n_mcu;
    const long long clk0 = clock64();
    PSync y;
    y.mine = sd.followed ? bd.prog_progress + chain : nullptr;
    y.published = m0;
    y.one_writer = sd.ah == 0; // first scans (DC and AC) store through lane 0 only
    y.pred = nullptr;
    y.avail = 0xFFFFFFFFu;
    if (sd.n_pred) {
        if (sd.pred_fine) { // same MCU order and restart interval: chain j follows chain j of the predecessor scan
            y.pred = bd.prog_progress + bd.prog_scan[sd.pred[0]].chain_off + j;
            y.avail = 0u;
        } else {            // otherwise every predecessor scan has to be through
            for (uint32_t q = 0; q < sd.n_pred; ++q) {
                const ProgScanDev& ps = bd.prog_scan[sd.pred[q]];
                for (uint32_t c = 0; c < ps.n_chain; ++c) {
                    PSync w;
                    w.pred = bd.prog_progress + ps.chain_off + c;
                    w.avail = 0u;
                    ps_need(w, ps.ri ? min(ps.n_mcu, (c + 1u) * ps.ri) : ps.n_mcu);
                }
            }
        }
    }
    uint32_t err = cst.flags; // K0 found the restart markers of this scan out of order / miscounted: nothing is decoded
    if (!err) {
        const bool huffman = !(sd.ss == 0 && sd.ah != 0);
        if (huffman) {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                if (i >= sd.ncomp)
                    break;
                const uint4* src = reinterpret_cast<const uint4*>(bd.prog_huff + sd.table[i]);
                uint4* dst = reinterpret_cast<uint4*>(&s_tab[i]);
                for (int q = lane; q < int(sizeof(ProgHuffDev) / 16); q += 32)
                    dst[q] = __ldg(src + q);
            }
        }
        __syncwarp();
        if (sd.ss != 0 && sd.ah == 0) {
            uint32_t* fl = reinterpret_cast<uint32_t*>(s_rl); // (the refinement table's storage: a chain is one or the other)
            for (int q = lane; q < 512; q += 32) {
                const uint32_t e = s_tab[0].lut[q];
                fl[q] = e ? first_entry(e >> 8, e & 0xFFu) : kRefSlow;
            }
            __syncwarp();
        }
        if (sd.ss != 0 && sd.ah != 0) {
            for (int q = lane; q < 512; q += 32) {
                const uint32_t e = s_tab[0].lut[q];
                const uint32_t f = e ? refine_entry(e >> 8, e & 0xFFu) : 0u;
                const uint32_t run = (f >> 5) & 15u;
                uint32_t v = (f & kRefBad) ? kRefSlow : ((f & 31u) | (run << 24));
                if (f & kRefEob) // an end of band for this block only runs "64 zeros" far, i.e. behind the band
                    v = run ? kRefSlow : ((f & 31u) | (64u << 24));
                s_rl[q] = e ? make_uint2(v, (f & kRefCoef) ? 1u << (32u - (f & 31u)) : 0u) : make_uint2(kRefSlow, 0u);
            }
            __syncwarp();
        }
        const uint32_t byte0 = j ? bd.bnd[ch.bnd_off + j - 1u] : 0u;
        const uint32_t byte1 = j + 1u < sd.n_chain ? bd.bnd[ch.bnd_off + j] : cst.cmp_len;
        PBits b; // (wmax: inside what the host reserved for this scan; the words past cmp_len are never trusted)
        pb_open(b, bd.cmp + ch.cmp_off / 4, (ch.raw_len + 15u) / 4u, pinned_smem_addr(s_ring), uint32_t(lane), byte0 * 8u);
        PMap g;
        const int c0 = sd.comp[0];
        g.wb = max(1u, uint32_t(sd.wb));
        g.lh = 31u - uint32_t(__clz(int(im.h[c0])));
        g.lv = 31u - uint32_t(__clz(int(im.v[c0])));
        g.bx = g.by = 0u;
        g.mcus_x = im.mcus_x;
        g.bpm = im.bpm;
        g.first = im.blk_first[c0];
        int16_t* coef = bd.pcoef + im.pcoef_off;
        PScan sc;
        sc.ncomp = sd.ncomp;
        sc.ss = sd.ss;
        sc.se = sd.se;
        sc.al = sd.al;
        sc.bpm = im.bpm;
#pragma unroll
        for (int i = 0; i < 4; ++i)
            sc.nb[i] = i < sd.ncomp ? int(im.h[sd.comp[i] & 3]) * int(im.v[sd.comp[i] & 3]) : 0;
        PTab tab[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            tab[i].lut = pinned_smem_addr(s_tab[i].lut);
            tab[i].t = &s_tab[i];
        }
        if (sd.ss == 0 && sd.ah == 0)
            prog_dc_first(b, tab, sc, g, coef, m0, m1, lane, err, y);
        else if (sd.ss == 0)
            prog_dc_refine(b, sc, g, coef, m0, m1, lane, y);
        else if (sd.ah == 0)
            prog_ac_first(b, tab[0], reinterpret_cast<const uint32_t*>(s_rl), sc, g, coef, m0, m1, lane, err, y);
        else
            prog_ac_refine(b, tab[0], s_rl, sc, g, coef, m0, m1, uint32_t(lane), err, s_q, y);
        if (b.p > byte1 * 8u)
            err |= kFlagOverrun; // the chain read past its restart interval / the end of the scan
    }
    if (err && lane == 0)
        atomicOr(&bd.state[sd.parent].flags, err);
    // whatever happened, the successors of this chain must not wait for ever; a fine-grained successor relies on this
    // chain having followed ITS predecessor to the end (the order is transitive)
    ps_need(y, m1);
    ps_publish(y, m1, true, uint32_t(lane));
    if (lane == 0)
        bd.prog_progress[bd.n_prog_chain + chain] = uint32_t((clock64() - clk0) >> 10); // (read back under B200JPEG_TRACE)
}

// After the scans: the per-block records the IDCT stage reads (bits 0-7: which 16-byte groups hold a non-zero AC
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
