/*
 * b200jpeg.h -- plain C ABI of the B200-native batched baseline-JPEG decoder.
 *
 * This is the boundary a foreign-language binding (ctypes, cgo, JNI ...) links
 * against: plain pointers and sizes, no C++ or torch types.  The nvImageCodec
 * extension entry point (nvimgcodecExtensionModuleEntry, exported by the same
 * shared object) is a thin adapter over these calls; see INTEGRATION.md.
 *
 * Each declaration cites the reference interface it stands in for
 * (paths relative to NVIDIA/nvImageCodec v0.8.0).
 */
#ifndef B200JPEG_H_
#define B200JPEG_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define B200JPEG_API __attribute__((visibility("default")))
#else
#define B200JPEG_API
#endif

typedef struct b200jpegDecoder_st* b200jpegDecoder_t;

/* status codes (per call and per image) */
enum {
    B200JPEG_OK = 0,
    B200JPEG_ERR_ARG = 1,
    B200JPEG_ERR_NOT_JPEG = 2,
    B200JPEG_ERR_UNSUPPORTED = 3, /* valid JPEG outside the baseline subset this decoder handles */
    B200JPEG_ERR_CORRUPT = 4,     /* header or entropy data damaged */
    B200JPEG_ERR_CUDA = 5,
    B200JPEG_ERR_STATE = 6,
    B200JPEG_ERR_TOO_LARGE = 7
};

/* output sample formats; mirror nvimgcodecSampleFormat_t subsets accepted by the
 * reference CPU decoder (extensions/libjpeg_turbo/libjpeg_turbo_decoder.cpp:165-178). */
enum {
    B200JPEG_FMT_I_RGB = 0,
    B200JPEG_FMT_I_BGR = 1,
    B200JPEG_FMT_P_RGB = 2,
    B200JPEG_FMT_P_BGR = 3,
    B200JPEG_FMT_Y = 4,     /* single plane: P_Y / I_Y */
    B200JPEG_FMT_I_YUV = 5, /* Y, Cb, Cr samples (chroma upsampled, no colour transform), interleaved */
    B200JPEG_FMT_P_YUV = 6  /* ... planar */
};

/* What the host marker walk found.  Superset of what the reference's header
 * parser reports (src/parsers/jpeg.cpp:238-404: dims, components, sampling, SOF). */
typedef struct {
    int width, height, num_components;
    int h[4], v[4];          /* sampling factors per component */
    int sof_marker;          /* 0xC0 baseline, 0xC1 extended sequential, 0xC2 progressive ... */
    int precision;
    int restart_interval;    /* in MCUs, 0 = none */
    int num_scans_seen;      /* 1 for single interleaved scan */
    int supported;           /* 1 if b200jpegDecodeBatch accepts the stream */
    int reason;              /* B200JPEG_* explaining supported == 0 */
    uint64_t scan_offset;    /* first entropy-coded byte */
    uint64_t scan_length;    /* entropy-coded bytes (byte-stuffed, markers included, EOI excluded) */
    int mcus_x, mcus_y, blocks_per_mcu;
} b200jpegStreamInfo_t;

/* Where and how one image is written.  Stands in for the nvimgcodecImageInfo_t the
 * plugin reads per sample (include/nvimgcodec.h:538-559) plus the ROI from
 * nvimgcodecCodeStreamView_t::region (include/nvimgcodec.h:503-521; start/end are [y,x]). */
typedef struct {
    void* buffer;       /* device pointer on the decoder's GPU, or host pointer if buffer_is_host */
    int format;         /* B200JPEG_FMT_* */
    int buffer_is_host; /* 0 = device (kernels store straight into it), 1 = host (decoder bounces + D2H) */
    size_t row_stride;  /* bytes per row of a plane; 0 = tight */
    int roi_x0, roi_y0, roi_x1, roi_y1; /* all zero = full image; else [x0,x1) x [y0,y1) inside the image */
    int orientation;    /* EXIF orientation code 1..8 to apply to the (cropped) result, 0 or 1 = none.  Codes 5..8
                           swap width and height: the buffer then holds roi_width rows of roi_height pixels.  What
                           nvimgcodecDecodeParams_t::apply_exif_orientation asks of a decoder (include/nvimgcodec.h:784);
                           the reference's CPU plugin declines it (libjpeg_turbo_decoder.cpp:206-209), nvjpeg_ext applies it. */
    size_t plane_stride; /* bytes between the planes of a planar output; 0 = row_stride * rows of the decoded region (planes back
                           to back).  Set when the region is embedded in a taller image, e.g. an out-of-bounds region whose
                           border the caller fills (src/imgproc/out_of_bound_roi_fill.cpp:90-171). */
} b200jpegOutput_t;

/* Host-only marker walk (SOI .. SOS): tables present, sampling, restart interval,
 * scan extent.  Replaces the acceptance predicate of
 * extensions/libjpeg_turbo/libjpeg_turbo_decoder.cpp:110-212 (canDecode) for the
 * stream-dependent part; never touches the GPU. */
B200JPEG_API int b200jpegProbe(const void* data, size_t length, b200jpegStreamInfo_t* info);

/* Create / destroy a decoder bound to one GPU.  options is the nvImageCodec
 * option string ("<id>:<name>=<value>" tokens, include/nvimgcodec.h:1967-1971);
 * understood: fancy_upsampling=<0|1> (default 1, as libjpeg_turbo_decoder.cpp:223-249); progressive=<0|1> (default 1: complete
 * progressive SOF2 streams are decoded -- a throughput path, one warp per scan; 0: declined with
 * B200JPEG_ERR_UNSUPPORTED), progressive_min_batch=<n> (default 1: progressive streams are taken only in batches that hold
 * at least n of them, fewer are declined).
 * Tuning knobs (defaults are measured choices, results do not depend on them): subseq_bits=<512|1024|2048|4096>
 * (bits of entropy-coded data per GPU thread; default by batch size), lanes=<1..8> (pipeline lanes, default 6),
 * chunk_mb / chunk_images (pipeline chunk, default 8 MB of bitstream or 64 images),
 * direct_kb (pinned bitstreams of at least this size are copied without staging, default 128), host_threads (helper
 * threads for marker walks and staging copies, default by the CPUs the process may use, at most 8), max_arena_mb (images whose
 * coefficient + plane arenas would exceed it are refused with B200JPEG_ERR_TOO_LARGE, default 32768), max_fix_rounds;
 * debug_no_warm=1 is a test hook that forces the re-synchronisation path.
 * Replaces nvimgcodecDecoderDesc_t::create / destroy (include/nvimgcodec.h:1466-1482). */
B200JPEG_API int b200jpegCreate(b200jpegDecoder_t* decoder, int device_id, const char* options);
B200JPEG_API int b200jpegDestroy(b200jpegDecoder_t decoder);

/* The same with the caller's allocators: every arena the decoder needs (device scratch, pinned staging) is then taken
 * from and returned to these callbacks, stream-ordered, the size handed back on free -- the contract of
 * nvimgcodecDeviceAllocator_t / nvimgcodecPinnedAllocator_t (include/nvimgcodec.h:246-318), which the extension
 * forwards from nvimgcodecExecutionParams_t (include/nvimgcodec.h:1095-1096).  `stream` is a cudaStream_t.  A NULL
 * struct or a NULL pair falls back to what the reference's own buffers do (src/imgproc/device_buffer.cpp:42-69):
 * cudaMallocAsync on the batch's stream where the device has memory pools, else cudaMalloc; cudaMallocHost. */
typedef struct {
    void* device_ctx;
    int (*device_malloc)(void* ctx, void** ptr, size_t size, void* stream);
    int (*device_free)(void* ctx, void* ptr, size_t size, void* stream);
    void* pinned_ctx;
    int (*pinned_malloc)(void* ctx, void** ptr, size_t size, void* stream);
    int (*pinned_free)(void* ctx, void* ptr, size_t size, void* stream);
} b200jpegAllocators_t;
B200JPEG_API int b200jpegCreateEx(b200jpegDecoder_t* decoder, int device_id, const char* options, const b200jpegAllocators_t* allocators);

/* Decode a batch: host bitstreams in, pixels out (device or host buffers).
 * Replaces nvimgcodecDecoderDesc_t::decodeBatch (include/nvimgcodec.h:1519-1533) /
 * DecoderImpl::decode (libjpeg_turbo_decoder.cpp:295-420).  status[i] receives
 * B200JPEG_* per image.  cuda_stream is a cudaStream_t.  The call returns once every status is
 * final, entropy-decode errors included.  When an output is a host buffer that is after its pixels have
 * arrived; when all outputs are device buffers the last pixel kernels may still be running: they are
 * ordered on cuda_stream, so work queued on that stream afterwards (or a synchronise on it) sees the
 * finished pixels, as with any stream-ordered CUDA call.  Option sync_finish=1 waits for the pixels too. */
B200JPEG_API int b200jpegDecodeBatch(b200jpegDecoder_t decoder, int batch_size, const void* const* data, const size_t* lengths,
    const b200jpegOutput_t* outputs, int* status, void* cuda_stream);

/* The same work split in three so callers can overlap and so the device-resident
 * part can be timed on its own:
 *   Prepare : host marker walk, staging, async H2D of bitstreams and descriptors
 *   Execute : the CUDA kernels only (everything already in HBM); asynchronous
 *   Finish  : wait for the statuses (for the pixels too when an output is a host buffer), run the rare
 *             re-synchronisation path, D2H of host outputs
 * Prepare/Execute/Finish must be called in that order on one decoder; Execute may be
 * repeated on the same prepared batch (it is idempotent). */
B200JPEG_API int b200jpegPrepareBatch(b200jpegDecoder_t decoder, int batch_size, const void* const* data, const size_t* lengths,
    const b200jpegOutput_t* outputs, void* cuda_stream);
B200JPEG_API int b200jpegExecuteBatch(b200jpegDecoder_t decoder, void* cuda_stream);
B200JPEG_API int b200jpegFinishBatch(b200jpegDecoder_t decoder, int* status, void* cuda_stream);

/* Kernel-level entry: the marker + restart-interval scan (K0) of ONE entropy-coded segment that is
 * already in device memory.  Drops the stuffed 0x00 after 0xFF, fill bytes and RSTn markers, stops at
 * the first other marker, and records where each restart interval starts.  What the reference gets
 * from libjpeg-turbo's bit reader + restart handling (jdhuff.c / jdmarker.c behind
 * extensions/libjpeg_turbo/jpeg_mem.cpp:311-379); exposed so the kernel can be tested on its own.
 *   dev_raw            device pointer, 16-byte aligned, readable for length + 16 bytes
 *   dev_out_words      device buffer of at least round_up(length, 16) + 32 bytes: the unstuffed stream as
 *                      big-endian 32-bit words (byte i of the stream is at byte address i ^ 3)
 *   dev_out_boundaries device uint32[boundary_capacity], boundary_capacity >= expected_restart_markers + 1:
 *                      unstuffed byte offset of the start of restart interval 1, 2, ...; the entry
 *                      after the last one is the unstuffed length
 *   dev_result         device pointer, filled asynchronously on cuda_stream
 * The decoder's own (cudaMalloc'ed, grow-only) scratch is used for the per-tile counts. */
typedef struct {
    uint32_t unstuffed_bytes;
    uint32_t num_restart_markers;
    uint32_t num_subsequences; /* ceil(unstuffed bits / 1024), 0 if flags != 0 */
    uint32_t flags;            /* bit 0: restart markers out of sequence or not the expected number */
} b200jpegScanResult_t;
B200JPEG_API int b200jpegScanSegment(b200jpegDecoder_t decoder, const void* dev_raw, size_t length, uint32_t expected_restart_markers,
    void* dev_out_words, uint32_t* dev_out_boundaries, uint32_t boundary_capacity, b200jpegScanResult_t* dev_result, void* cuda_stream);

/* The same scan for callers that cannot allocate.  (No counterpart in the reference: nvImageCodec
 * plugins always go through an allocator.)
 *
 * Who this is for: code that runs after the process has handed all of its device memory to something
 * else -- a training framework's caching pool, a fixed up-front budget in an ingest service -- so that
 * a later cudaMalloc can fail outright, and code on a latency-critical path where the device-wide
 * synchronisation cudaMalloc / cudaFree imply is not acceptable.  b200jpegScanSegment grows the
 * decoder's scratch with cudaMalloc the first time it sees a larger segment; this entry point never
 * calls cudaMalloc, cudaFree or any other allocating routine, and needs no decoder handle.
 *
 * How: the per-tile counts, the tile map and the segment descriptor that K0 needs live in one fixed
 * 11 KiB array inside the library (a static __device__ array; every GPU that loads the library
 * holds it until the process exits).  That bounds the segment: length must not exceed
 * B200JPEG_NOALLOC_MAX_SEGMENT_BYTES (8 MiB of byte-stuffed scan data); longer segments are refused
 * with B200JPEG_ERR_TOO_LARGE and nothing is launched.  Inputs and outputs are the caller's, on the device
 * that is current when the call is made, with the same requirements as for b200jpegScanSegment; in
 * particular dev_raw must be 16-byte aligned and readable for length + 16 bytes (the scan reads whole
 * 16-byte chunks, the last one past the end of the segment), and dev_out_words must hold
 * round_up(length, 16) + 32 bytes.
 *
 * The first call on each device looks up the array's device address (cudaGetSymbolAddress, with
 * that device current); make that first call before starting any stream capture.  All calls on a
 * device share the one array, and it keeps whatever the previous call left: the entry point
 * rewrites the parts that must start clean on the given stream, but two calls whose GPU work can
 * overlap (different streams, different host threads) race on it -- serialise them.
 *
 * message (optional, may be NULL) receives a human-readable reason when the return value is not
 * B200JPEG_OK. */
#define B200JPEG_NOALLOC_MAX_SEGMENT_BYTES (8u << 20)
B200JPEG_API int b200jpegScanSegmentNoAlloc(const void* dev_raw, size_t length, uint32_t expected_restart_markers, void* dev_out_words,
    uint32_t* dev_out_boundaries, uint32_t boundary_capacity, b200jpegScanResult_t* dev_result, void* cuda_stream, char* message,
    size_t message_capacity);

/* One batch across several GPUs of one box.  Images are independent, so there is no
 * collective: the host partitions the batch by compressed byte count (greedy longest-processing-time,
 * deterministic), and every device decodes its share from its own pinned staging on its own stream,
 * driven by its own host thread.  The reference has no multi-GPU notion (one decoder == one
 * exec_params.device_id, include/nvimgcodec.h:1101); this is the thin driver above the decoder that
 * its users would otherwise write themselves.
 *   Plan        : device_of_image[i] = index (into device_ids) of the device that will decode image i
 *   DecodeBatch : outputs[i].buffer must be host memory (buffer_is_host = 1) or device memory on the
 *                 planned device of image i; returns when every device has finished. */
typedef struct b200jpegMultiDecoder_st* b200jpegMultiDecoder_t;
B200JPEG_API int b200jpegMultiCreate(b200jpegMultiDecoder_t* multi, const int* device_ids, int num_devices, const char* options);
B200JPEG_API int b200jpegMultiDestroy(b200jpegMultiDecoder_t multi);
B200JPEG_API int b200jpegMultiPlan(b200jpegMultiDecoder_t multi, int batch_size, const size_t* lengths, int* device_of_image);
B200JPEG_API int b200jpegMultiDecodeBatch(b200jpegMultiDecoder_t multi, int batch_size, const void* const* data, const size_t* lengths,
    const b200jpegOutput_t* outputs, const int* device_of_image, int* status);

/* Introspection for benchmarks and tests. */
enum {
    B200JPEG_STAT_KERNEL_LAUNCHES = 0, /* kernels launched by the last Execute */
    B200JPEG_STAT_ALGO_BYTES = 1,      /* C + 4*N_coef + P of the last prepared batch */
    B200JPEG_STAT_COMPRESSED_BYTES = 2,
    B200JPEG_STAT_NUM_COEFS = 3,
    B200JPEG_STAT_OUTPUT_BYTES = 4,
    B200JPEG_STAT_RESYNC_IMAGES = 5,   /* images that needed the re-synchronisation path in the last Finish */
    B200JPEG_STAT_DEVICE_BYTES = 6,    /* device memory currently held */
    B200JPEG_STAT_NUM_SUBSEQ = 7,
    B200JPEG_STAT_SYNC_CTAS = 8,          /* K1 counters of the last Execute (lane 0): CTAs, */
    B200JPEG_STAT_SYNC_REPAIR_ROUNDS = 9, /*   fixed-point rounds after the first full pass (summed over CTAs), */
    B200JPEG_STAT_SYNC_REPAIRED = 10,     /*   subsequences re-decoded in those rounds, */
    B200JPEG_STAT_SYNC_MAX_ROUNDS = 11    /*   most rounds any CTA needed */
};
B200JPEG_API int64_t b200jpegGetStat(b200jpegDecoder_t decoder, int which);

/* Per-stage device times (ms) of the last Execute, measured with CUDA events on the
 * launch stream when enabled with b200jpegSetProfiling(decoder, 1).  names[i] are static strings. */
B200JPEG_API int b200jpegSetProfiling(b200jpegDecoder_t decoder, int enable);
B200JPEG_API int b200jpegGetStageTimes(b200jpegDecoder_t decoder, float* ms, const char** names, int capacity);

/* Intermediate results of image `index` of the last executed batch (tests only):
 * coefficients as int16 [mcu][block-in-mcu][64] in zig-zag order. */
B200JPEG_API int b200jpegDebugCopyCoefficients(b200jpegDecoder_t decoder, int index, int16_t* host_out, size_t capacity_elems);

B200JPEG_API const char* b200jpegGetLastError(b200jpegDecoder_t decoder);
B200JPEG_API const char* b200jpegVersion(void);

#ifdef __cplusplus
}
#endif
#endif /* B200JPEG_H_ */
