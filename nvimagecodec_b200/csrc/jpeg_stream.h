// Host-side JPEG marker walk and Huffman table preparation for the b200 decoder.
//
// The reference's built-in parser (src/parsers/jpeg.cpp:269-345) stops at SOS and
// never reads DQT / DHT / DRI / the SOS component selectors, so the plugin does its
// own walk here (T.81 B.2).  Nothing in this file touches the GPU.
#pragma once
#include <cstddef>
#include <cstdint>
#include <string>
#include <vector>

namespace b200jpeg {

constexpr int kMaxComp = 4;
constexpr int kMaxBlocksPerMcu = 10;
constexpr int kLutBits = 11;    // first-level Huffman lookup width
constexpr int kLut2Size = 256;  // second-level entries (16-bit windows of the codes longer than kLutBits)

enum StreamStatus { kStreamOk = 0, kStreamNotJpeg, kStreamUnsupported, kStreamCorrupt, kStreamMultiScan, kStreamProgressive };

struct HuffSpec {
    uint8_t bits[17]; // bits[l] = number of codes of length l (1..16)
    uint8_t vals[256];
    bool present;
};

struct StreamHeader {
    int width = 0, height = 0, ncomp = 0, precision = 0, sof = 0;
    int comp_id[kMaxComp] = {0, 0, 0, 0};
    int h[kMaxComp] = {0, 0, 0, 0}, v[kMaxComp] = {0, 0, 0, 0}, tq[kMaxComp] = {0, 0, 0, 0};
    int td[kMaxComp] = {0, 0, 0, 0}, ta[kMaxComp] = {0, 0, 0, 0};
    int hmax = 0, vmax = 0, mcus_x = 0, mcus_y = 0, blocks_per_mcu = 0;
    int restart_interval = 0;
    int nscans = 0;
    size_t scan_offset = 0, scan_length = 0;
    uint16_t qt[4][64]; // natural order
    bool qt_present[4] = {false, false, false, false};
    HuffSpec huff[2][4]; // [class][id]
    int adobe_transform = -1;
    bool has_jfif = false;
};

// Walks SOI..SOS, fills hdr, and decides whether the CUDA path handles the stream.
StreamStatus parse_stream(const uint8_t* data, size_t len, StreamHeader* hdr);

// Sequential streams whose three components come in three scans of one component each (T.81 A.2.2, non-interleaved):
// parse_stream answers kStreamMultiScan for them, and this second walk fills `frame` (geometry of the whole image, no scan)
// and scans[c] for component c: a self-contained ONE-component stream description -- the component's own size, the
// quantisation / Huffman tables and the restart interval in force when its scan starts, where its entropy-coded data lies.
// Each scan then decodes like a grayscale image of the component's size; the colour stage reads the three planes.
StreamStatus parse_multiscan(const uint8_t* data, size_t len, StreamHeader* frame, StreamHeader* scans /* [3] */);

// Progressive streams (SOF2, Huffman, 8 bits): parse_stream answers kStreamProgressive for them and this walk lists the
// scans (T.81 G.1: spectral selection Ss..Se, successive approximation Ah/Al; jdphuff.c start_pass_phuff_decoder checks).
// Every scan keeps a copy of the Huffman tables in force at its SOS (they are usually redefined before every scan).
// Only complete progressions are taken (every coefficient of every component refined
// down to its last bit): libjpeg smooths blocks of incomplete ones (jdcoefct.c smoothing_ok), so those stay with
// the CPU decoders.  The scan-count limit is NVIMGCODEC_MAX_JPEG_SCANS (default 256), as in
// extensions/libjpeg_turbo/jpeg_handle.cpp:38-75.
struct ProgScan {
    int ncomp = 0;               // components in the scan: 1, or all of the frame's (interleaved; DC scans only)
    int comp[kMaxComp] = {0, 0, 0, 0};   // frame component index of each
    int table[kMaxComp] = {-1, -1, -1, -1}; // index into ProgStream::tables (-1: none -- DC refinement reads raw bits)
    int ss = 0, se = 0, ah = 0, al = 0;
    int restart_interval = 0;    // in MCUs of THIS scan (one block each for a one-component scan)
    int mcus = 0;                // MCUs of this scan
    int wb = 0, hb = 0;          // one-component scan: block columns / rows of the component, ceil(size / 8)
    size_t offset = 0, length = 0; // entropy-coded segment, up to and including the marker that ends it
};
struct ProgStream {
    std::vector<ProgScan> scans;
    std::vector<HuffSpec> tables;
};
StreamStatus parse_progressive(const uint8_t* data, size_t len, StreamHeader* frame, ProgStream* out);
// The SOFn marker byte of a stream (0xC0 .. 0xCF), 0 if none is found before the first scan: segment headers only, no tables read.
int sof_marker_of(const uint8_t* data, size_t len);

// Device image of one Huffman table for the progressive scan kernel (progressive.cuh): plain symbols, not the packed
// actions of the sequential kernels -- a progressive symbol's meaning depends on the scan type.
struct alignas(16) ProgHuffDev {
    uint16_t lut[512];   // 9-bit window -> (code length << 8) | symbol; 0 = longer than 9 bits or unassigned
    int32_t maxcode[18]; // largest code of length l (1..16), -1 if none; [17] = sentinel
    int32_t valoff[18];  // symbol index = valoff[l] + code
    uint8_t vals[256];
};
void build_prog_table(const HuffSpec& spec, ProgHuffDev* out);

// Device-side Huffman table image (one per table; a "set" is DC0, DC1, AC0, AC1).
struct alignas(16) HuffTableDev {
    uint16_t lut[1 << kLutBits]; // packed symbol action (see pack_symbol); 0 = longer than kLutBits or invalid
    int32_t maxcode[18];         // maxcode[l] for l in 1..16 (canonical), -1 if none; [17] = sentinel
    int32_t valoff[18];          // sym index = valoff[l] + code
    uint16_t sym[256];           // packed symbol actions in canonical order (last-resort canonical search)
    // Second level for codes longer than kLutBits: in a canonical code they are exactly the 16-bit windows
    // >= first16 (the first long code, left-aligned), usually a hundred-odd values, so one more lookup
    // indexed by (window16 - first16) resolves them; n2 == 0 means "too many, use the canonical search".
    uint32_t first16, n2;
    uint16_t lut2[kLut2Size];
};
struct alignas(16) HuffSetDev {
    HuffTableDev t[4]; // 0: DC id0, 1: DC id1, 2: AC id0, 3: AC id1
};

void build_huff_table(const HuffSpec& spec, bool is_dc, HuffTableDev* out);
void huff_set_key(const StreamHeader& hdr, std::string* key); // canonical bytes of the table specs: the cache key
void build_huff_set(const StreamHeader& hdr, HuffSetDev* out);

} // namespace b200jpeg
