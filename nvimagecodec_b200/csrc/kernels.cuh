// Device-side data layout shared by the kernels (kernels.cu) and the batch
// orchestration (decoder.cpp).  See DESIGN.md "Data layout in HBM".
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

#include "jpeg_stream.h"

namespace b200jpeg {

// ---- entropy-stage geometry
constexpr int kScanThreads = 256;     // K0: 64 B (four 16-byte chunks) of byte-stuffed scan data per thread,
constexpr int kScanChunks = 4;        //     16 KiB per CTA
constexpr int kScanTileBytes = kScanThreads * 16 * kScanChunks;
constexpr int kMinSubseqBits = 512;   // K1-K3: bits of unstuffed stream per thread ("subsequence"); the host picks
constexpr int kMaxSubseqBits = 4096;  //   512 .. 4096 per batch (BatchDev::sub_bits) so the GPU stays filled
constexpr int kSeqThreads = 256;      // threads per CTA of the entropy kernels
constexpr int kSeqMult = 1;           // K1: subsequences per thread ...
constexpr int kSeqLogical = kSeqThreads * kSeqMult; // ... i.e. per CTA ("sequence"), of which the
constexpr int kSeqWarmBits = 8192;    // leading ones only warm up the entry state: this many bits of them (BatchDev::seq_warm),
constexpr int kSeqWarmMin = 8;        // at least this many subsequences
#ifndef B200_IDCT_PER
#define B200_IDCT_PER 4
#endif
constexpr int kIdctPer = B200_IDCT_PER;        // K4a: blocks per thread (table load, barrier and descriptor reads amortised)
#ifndef B200_COLOR_UNITS_Y
#define B200_COLOR_UNITS_Y 4
#endif
constexpr int kColorUnitsY = B200_COLOR_UNITS_Y;       // K4b: vertically adjacent units per thread (descriptor reads, index arithmetic and the mode switch amortised)
constexpr uint32_t kScatterStride = 512; // bytes between consecutive words of a thread's scatter column in the IDCT kernel (128 threads x 4)
constexpr int kDcThreads = 512;       // K3b: threads per CTA, ...
constexpr int kDcPer = 8;             //      blocks per thread,
constexpr int kDcChunk = kDcThreads * kDcPer; // blocks per CTA

// ---- per-image error / state flags (ImageState::flags)
enum : uint32_t {
    kFlagBadMarker = 1u << 0,   // RST out of sequence / wrong restart count
    kFlagBadHuffman = 1u << 1,  // invalid code met by a synchronised decoder
    kFlagOverrun = 1u << 2,     // coefficient positions ran past the image or ended short
    kFlagDirty = 1u << 3,       // sequence entry states not yet consistent (needs the fix-up path)
    kFlagPhase = 1u << 4,       // prefix-summed block index disagrees with the decoder's MCU phase
    kFlagRedo = 1u << 5         // went through the fix-up path: coefficients / pixels must be produced again
};

// Output formats (same values as include/b200jpeg.h)
enum : int { kFmtIRGB = 0, kFmtIBGR = 1, kFmtPRGB = 2, kFmtPBGR = 3, kFmtY = 4, kFmtIYUV = 5, kFmtPYUV = 6 };

struct ImageDev {
    // stream geometry
    uint64_t raw_off;   // bytes into the raw (byte-stuffed) arena, 16-aligned
    uint64_t cmp_off;   // bytes into the unstuffed arena, 16-aligned
    uint64_t coef_off;  // int16 elements into the coefficient arena (block index * 64; also locates the block records)
    uint64_t sym_off;   // 16-byte chunks into the entry-stream arena (entry-stream form of K3)
    uint64_t pcoef_off; // progressive images: int16 elements into BatchDev::pcoef (dense zig-zag blocks, MCU order)
    uint64_t plane_off[3]; // bytes into the plane arena
    uint32_t raw_len;
    uint32_t tile_off, n_tiles;     // K0 tiles
    uint32_t sub_off, n_sub_max;    // subsequences (upper bound from the stuffed length)
    uint32_t seq_off, n_seq;        // sequences (CTAs of K1)
    uint32_t wseq_off, n_wseq;      // CTAs of K3 (kSeqThreads subsequences each)
    uint32_t bnd_off, n_bnd_cap;    // restart boundaries (+ end of stream)
    uint32_t restart_interval;      // MCUs, 0 = none
    uint32_t n_rst_expected;
    uint32_t total_blocks;          // n_mcu * blocks_per_mcu
    uint32_t write_blk_lo, write_blk_hi; // blocks (MCU decode order) of the MCU rows the requested ROI needs, incl. one row of context
    uint32_t dc_blk_lo;                  // start of the restart interval write_blk_lo lies in (0 without restarts): DC chain start
    uint16_t mcu_row_lo, mcu_row_hi;     // the same window as MCU rows
    uint32_t huff_set;
    uint32_t blk_comp;              // 2 bits per block-in-MCU: component
    uint32_t blk_tab;               // bit b: DC table id of block b, bit 16+b: AC table id
    uint32_t dc_cta_off, n_dc_cta;  // K3b chunks of kDcChunk blocks covering [dc_blk_lo, write_blk_hi)
    uint32_t idct_cta_off, n_idct_cta;
    uint32_t color_cta_off, n_color_cta_x, n_color_cta_y;
    uint32_t color_cta_x_rcp;       // ceil(2^32 / n_color_cta_x): cta -> (row, column) without a division
    // fused pixel kernel (K4): tiles (one per warp) = strips of MCU columns x bands of MCU rows over the MCUs the region of interest
    // touches; px_n_cta tiles from px_cta_off on; 0: the image goes through the idct + color kernels instead (generic sampling,
    // EXIF orientation, three-scan streams)
    uint32_t px_cta_off, px_n_cta;
    uint16_t px_strip_mcus, px_n_strips, px_band_rows, px_n_bands;
    uint16_t px_mcu_col_lo, px_mcu_col_hi, px_mcu_row_lo, px_mcu_row_hi;
    uint32_t plane_stride_px[3];    // row pitch of each MCU-padded component plane
    uint16_t comp_w[3], comp_h[3];  // real component size ceil(W*h/hmax) x ceil(H*v/vmax): where edges replicate
    uint16_t width, height, mcus_x, mcus_y;
    uint8_t ncomp, hmax, vmax, bpm;
    uint8_t h[3], v[3];
    uint8_t blk_first[3];           // first block-in-MCU index of each component
    uint8_t ncomp_needed;           // 1 when only luma is consumed (Y output)
    uint16_t qt[3][64];             // natural order, per component
    // output
    uint8_t* out;
    uint64_t plane_stride;          // bytes between planes of planar output
    uint32_t row_stride;
    uint16_t roi_x0, roi_y0, roi_x1, roi_y1;
    uint8_t fmt, fancy;
    uint8_t color_mode;      // K4b specialisation: 0 generic, 1 luma only, 2..5 chroma (hx,vx) = (1,1),(2,1),(1,2),(2,2) replicated, 6..8 = 3..5 with the triangle filter
    uint8_t color_unit_rows; // output rows per K4b thread (the chroma vertical factor on the fast paths)
    uint8_t orient;          // EXIF orientation code applied when storing (0/1 = none; 5..8 transpose)
    uint8_t q8;              // 1: every quantiser of the image fits 8 bits (all baseline tables do)
    uint8_t prog;            // 1: progressive stream -- its scans (device images of their own, K0 only) were decoded into
                             //    BatchDev::pcoef by the progressive kernels; the IDCT stage reads the dense blocks from there
    uint8_t multi;           // 1: the components came in three scans and were decoded as one-component images of their own;
                             //    BatchDev::gate names them (their state flags become this image's before K4b)
};

struct ImageState {
    uint32_t cmp_len;  // unstuffed bytes
    uint32_t n_rst;    // RST markers found
    uint32_t n_sub;    // ceil(cmp_len*8 / sub_bits)
    uint32_t flags;
};

// One scan of a progressive image (progressive.cuh).  Its entropy-coded segment went through K0 as a device image of its
// own (`child`: unstuffed words, restart boundaries); `parent` holds the geometry and the coefficient blocks.  A chain is
// one restart interval of the scan (the whole scan without restarts): one warp decodes it serially.
struct ProgScanDev {
    uint32_t parent, child;
    uint32_t chain_off, n_chain;
    uint32_t table[4];      // BatchDev::prog_huff index per scan component
    uint32_t pred[4];       // scans (BatchDev::prog_scan indices, all lower than this one's) that last wrote coefficients of the
    uint8_t n_pred;         //   same components and class (DC / AC) -- this scan follows them block by block when the MCU order
    uint8_t pred_fine;      //   and restart interval agree (pred_fine: exactly one predecessor), else waits for their end
    uint8_t followed;       // some later scan names this one as a predecessor: its chains publish their progress
    uint32_t ri, n_mcu;     // restart interval and size of the scan in ITS MCUs
    uint16_t wb, hb;        // one-component scan: the component's block columns / rows
    uint8_t ncomp, comp[4], ss, se, ah, al;
};

// All device pointers of one prepared batch.
struct BatchDev {
    int n_images;
    const ImageDev* img;
    ImageState* state;
    const uint8_t* raw;
    uint32_t* cmp;          // unstuffed stream, big-endian 32-bit words
    uint2* tile_info;       // K0a: x = kept | end<<31, y = rst count
    uint2* tile_base;       // K0b: exclusive prefix (x = kept or 0xFFFFFFFF if past the end, y = rst)
    const uint32_t* tile_img; // tile -> image
    uint32_t* bnd;          // restart boundaries (unstuffed byte positions) then end of stream
    const uint32_t* seq_img;  // sequence -> image
    const uint32_t* wseq_img; // K3 CTA -> image
    uint32_t* sub_in;       // entry state per subsequence
    uint32_t* sub_out;      // exit state per subsequence
    uint32_t* sub_info;     // blocks that end inside the subsequence (or absolute index | 1<<31 after a restart)
    uint32_t* sub_start;    // exclusive scan: absolute block index at entry
    const HuffSetDev* huff;
    int16_t* coef;
    // Entry-stream form of the coefficients (K3 huff_emit / the walker of the IDCT kernel): every subsequence thread of K3 writes
    // one 32-bit entry per decoded symbol -- low half: the coefficient; high half: the byte offset of its zig-zag position in
    // the IDCT kernel's scatter column (kScatterStride), bit 14: the block ends with this entry, bit 15: the subsequence's
    // entries end here; 0 for symbols that carry no coefficient -- in chunks of four (16 bytes).  The chunks of the 32
    // subsequences of a warp interleave, so that a warp's store of a chunk row is 512 contiguous bytes: chunk c of
    // subsequence k of an image lies at sym[sym_off + ((k >> 5) * sym_chunks + c) * 32 + (k & 31)].
    // blkpos[block] = bit position of the block's DC code word (-> the subsequence that holds it), the low half of
    // blkrec[block] = the entry slot of that DC symbol inside the subsequence.
    uint4* sym;
    uint32_t* blkpos;
    uint32_t sym_chunks;    // chunks reserved per subsequence
    uint32_t sub_shift;     // log2(sub_bits)
    int sym_mode;           // 1: entry-stream form, 0: dense 16-byte coefficient groups
    uint32_t* blkrec;       // per coefficient block: bits 0-7 which of its eight 16-byte groups K3 wrote (bit g = group g),
                            // bits 16-31 its DC: the coded difference out of K3, the value after K3b
    uint8_t* planes;
    const uint32_t* dc_cta_img;   // K3b CTA -> image
    int4* dc_sum;                 // K3b: per chunk (reset seen, DC sums of the three components)
    const uint32_t* idct_cta_img;
    const uint32_t* color_cta_img;
    const uint32_t* px_cta_img;   // fused pixel kernel: tile (one per warp) -> image
    int n_px_task;                // tiles
    uint32_t px_smem;             // shared memory per warp of the fused pixel kernel: the largest tile of the batch
    // progressive images
    const ProgScanDev* prog_scan;
    const uint32_t* prog_chain;   // chain -> scan
    uint32_t* prog_progress;      // per chain: MCUs of its scan finished so far (zeroed before the launch)
    int n_prog_chain;
    const ProgHuffDev* prog_huff;
    const uint32_t* prog_img;     // the progressive images of the batch (for the record pass)
    int16_t* pcoef;               // their coefficient blocks: 64 int16 in zig-zag order per block, MCU order, DC included
    int n_prog_img;
    uint32_t prog_max_blocks;     // largest total_blocks among them
    const uint32_t* gate;         // three image indices per image, see ImageDev::multi
    int n_tiles, n_seq, n_wseq, n_sub, n_dc_cta, n_idct_cta, n_color_cta;
    int n_multi;            // images with ImageDev::multi set
    uint32_t* sync_stats;   // optional K1 counters (4 x uint32), may be null
    uint32_t sub_bits;      // subsequence length in bits for this batch
    uint32_t seq_warm;      // warm-up subsequences at the head of every K1 CTA
};

// Kernel launchers (kernels.cu).  Each returns the number of kernels it launched.
int launch_scan(const BatchDev& b, cudaStream_t s);                 // K0a, K0b, K0c
int launch_sync(const BatchDev& b, bool fix_mode, bool no_warm, cudaStream_t s); // K1 (first pass or fix-up pass)
int launch_verify_scan(const BatchDev& b, bool fix_mode, cudaStream_t s); // K2
int launch_write(const BatchDev& b, bool redo_only, cudaStream_t s);  // K3 (dense groups or entry stream, by b.sym_mode)
uint32_t sym_chunks_for(uint32_t sub_bits); // chunks of eight entries a subsequence of that many bits can need at most
int launch_dc_scan(const BatchDev& b, bool redo_only, cudaStream_t s); // K3b
// Progressive scans: every chain of the batch in one launch, then the per-block records for the IDCT stage
int launch_progressive(const BatchDev& b, cudaStream_t s);
int launch_idct(const BatchDev& b, bool redo_only, cudaStream_t s);   // K4a
int launch_color(const BatchDev& b, bool redo_only, cudaStream_t s);  // K4b
int launch_pixels(const BatchDev& b, bool redo_only, cudaStream_t s); // K4 fused (IDCT + upsampling + colour)
size_t pixels_tile_smem(int ncomp, const int* h, const int* v, int vmax, int strip_mcus, bool fancy_h, bool fancy_v); // its shared memory for one tile shape
constexpr size_t kPixelsMaxSmem = 96 * 1024; // hard limit of a tile; strips are cut to pixels_smem_budget()
size_t pixels_smem_budget(); // shared memory a warp's tile may take so that the intended number of warps fits an SM
int pixels_strip_mcus(int bpm_held, int halo_blocks); // MCUs per strip: three full passes of blocks per warp and step

} // namespace b200jpeg
