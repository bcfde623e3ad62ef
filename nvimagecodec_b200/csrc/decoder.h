// Batch orchestration of the b200 JPEG decode path: host marker walk, one staging
// upload per batch, kernel sequence, the rare re-synchronisation path, statuses.
// Mirrors the role of DecoderImpl in the reference's batched plugin
// (extensions/nvjpeg/hw_decoder.cpp:636-888) for a single device.
#pragma once
#include <cuda_runtime.h>

#include <condition_variable>
#include <cstdint>
#include <deque>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

#include "b200jpeg.h"
#include "jpeg_stream.h"
#include "kernels.cuh"

namespace b200jpeg {

// Where a decoder gets its memory and its helper threads from.  An nvImageCodec host hands plugins its own
// allocators (nvimgcodecExecutionParams_t::device_allocator / pinned_allocator, include/nvimgcodec.h:246-281:
// stream-ordered, the size is passed back on free) and its executor; the extension adapter forwards them here.
// Unset members fall back to what the reference's own buffers do (src/imgproc/device_buffer.cpp:42-69):
// cudaMallocAsync / cudaFreeAsync on the batch's stream where the device has memory pools, else cudaMalloc;
// cudaMallocHost for pinned memory; a small internal thread pool for host-side fan-out.
struct Services {
    void* device_ctx = nullptr;
    int (*device_malloc)(void* ctx, void** ptr, size_t size, cudaStream_t stream) = nullptr;
    int (*device_free)(void* ctx, void* ptr, size_t size, cudaStream_t stream) = nullptr;
    void* pinned_ctx = nullptr;
    int (*pinned_malloc)(void* ctx, void** ptr, size_t size, cudaStream_t stream) = nullptr;
    int (*pinned_free)(void* ctx, void* ptr, size_t size, cudaStream_t stream) = nullptr;
    // run fn(i, arg) for i in [0, n) on helper threads (the caller may take part); returns when all are done
    void* par_ctx = nullptr;
    void (*parallel_for)(void* ctx, int n, void (*fn)(int i, void* arg), void* arg) = nullptr;
};

// B200JPEG_TRACE=1 in the environment: one line per batch on stderr with the host-side phase times (microseconds).
bool trace_enabled();
double trace_now_us();

struct DeviceBuffer {
    void* ptr = nullptr;
    size_t cap = 0;
    cudaStream_t stream = nullptr; // the stream the block was allocated on (frees are ordered on it)
    int idle = 0;                  // consecutive batches that needed less than a quarter of the block
    bool ensure(size_t bytes, size_t* total, const Services& sv, cudaStream_t s); // grows; shrinks after a run of small batches
    void release(const Services& sv);
    template <typename T>
    T* as() const { return static_cast<T*>(ptr); }
};

struct PinnedBuffer {
    void* ptr = nullptr;
    size_t cap = 0;
    bool ensure(size_t bytes, const Services& sv, cudaStream_t s);
    void release(const Services& sv);
    template <typename T>
    T* as() const { return static_cast<T*>(ptr); }
};

// A few persistent helper threads for the host side of a batch (marker walks, staging copies).
class HostPool {
  public:
    explicit HostPool(int n_threads);
    ~HostPool();
    void run(int n, void (*fn)(int, void*), void* arg); // caller takes part; returns when all n are done
    int size() const { return int(threads_.size()); }

  private:
    struct Impl;
    Impl* impl_;
    std::vector<std::thread> threads_;
};

struct Options {
    bool fancy_upsampling = true; // libjpeg_turbo_ext default (libjpeg_turbo_decoder.cpp:223-249)
    bool debug_no_warm = false;   // test hook: forces the re-synchronisation path
    int max_fix_rounds = 1 << 20;
    int subseq_bits = 0;          // 0 = choose per batch; else 512 / 1024 / 2048 / 4096
    int chunk_mb = 0;             // pipeline chunk: this much bitstream (0 = 8 MB) ...
    int direct_kb = 128;          // pinned bitstreams of at least this many KiB are copied straight from the caller's buffer
    int lanes = 6;                // pipeline lanes (decoder + stream each) the chunks rotate over
    int chunk_images = 64;        // ... or this many images, whichever comes first
    int host_threads = 0;         // helper threads for marker walks / staging (0 = by the CPUs this process may run on, at most 8)
    bool entry_stream = true;     // K3 writes the coefficients as a stream of 16-bit entries (huff_emit) instead of dense 16-byte groups (huff_write)
    bool progressive = true;      // take progressive (SOF2) streams (progressive.cuh); off: they are declined and fall to the next decoder
    int progressive_min_batch = 1; // progressive streams are taken only in batches that hold at least this many of them (a lone
                                   // progressive image gains nothing here over a CPU core, and it holds back the batch it travels in): fewer -> declined
    bool fused_pixels = false;    // IDCT + upsampling + colour in one kernel (pixels_kernel); off: the idct and color kernels, which are faster as measured (profiles/)
    bool use_executor = false;    // (extension) fan host work out over the framework's executor instead of the decoder's own helper threads (measured: ~0.1 ms of dispatch latency per fan-out)
    bool sync_finish = false;     // Finish / DecodeBatch wait for the pixel kernels too, not only for the statuses
    int max_arena_mb = 32768;     // refuse (ERR_TOO_LARGE) images whose coefficient + plane arenas would exceed this many MiB
};

Options parse_options(const char* options, const char* decoder_id);
struct ParsedInput;
void parse_input(const void* data, size_t len, ParsedInput* out);

// What the marker walk found for one input of a batch; computed by Decoder::parse_inputs (in parallel when the
// batch is large) or handed in by a caller that walked the stream already (the extension's canDecode).
struct ParsedInput {
    StreamStatus status = kStreamNotJpeg;
    StreamHeader frame;          // kStreamOk: the stream; kStreamMultiScan: geometry of the whole image
    std::unique_ptr<StreamHeader[]> scans; // kStreamMultiScan only: the three one-component scans
    StreamStatus multi_status = kStreamOk;
    std::unique_ptr<ProgStream> prog;      // kStreamProgressive only: the scans (frame = geometry of the image)
};

class Decoder {
  public:
    Decoder(int device_id, const Options& opt, const Services& sv, HostPool* pool);
    ~Decoder();

    // parsed: nullptr, or n entries of which the non-null ones replace the marker walk of that input
    int prepare(int n, const void* const* data, const size_t* len, const b200jpegOutput_t* outs, cudaStream_t stream,
        const ParsedInput* const* parsed = nullptr, int sources_pinned = -1, bool defer_upload = false);
    // defer_upload: prepare() stops after the layout; the caller then runs stage_run() (on any host thread) and upload()
    void stage_run();
    int upload(cudaStream_t stream);
    int execute(cudaStream_t stream);
    int finish(int* status, cudaStream_t stream);
    int finish_async(cudaStream_t stream);          // queue read-backs / host-output copies
    int finish_wait(int* status, cudaStream_t stream); // wait (for the entropy stage only when no output is a host buffer), fix-up path, statuses

    int64_t stat(int which) const;
    void set_profiling(bool on) { profiling_ = on; }
    int stage_times(float* ms, const char** names, int cap);
    int copy_coefficients(int index, int16_t* host_out, size_t cap_elems);
    int scan_segment(const void* dev_raw, size_t length, uint32_t expected_rst, void* dev_words, uint32_t* dev_bnd, uint32_t bnd_cap,
        void* dev_result, cudaStream_t stream);
    const char* last_error() const { return err_.c_str(); }
    int device() const { return device_; }
    cudaEvent_t done_event() const { return done_event_; } // recorded behind the last kernel / copy of the last batch
    bool has_host_outputs() const { return any_host_out_; }

  private:
    int fail(int code, const std::string& msg);
    bool cuda_ok(cudaError_t e, const char* what);
    void mark(int stage, cudaStream_t s);
    void parallel_for(int n, void (*fn)(int, void*), void* arg);
    int run_pixel_stage(bool redo_only, cudaStream_t stream);
    bool copy_host_output(int d, cudaStream_t stream);

    int device_;
    Options opt_;
    Services sv_;
    HostPool* pool_;
    std::string err_;

    // batch bookkeeping (host)
    int n_in_ = 0;                      // images passed to prepare
    std::vector<int> host_status_;      // per input image, B200JPEG_*
    std::vector<int> dev_index_;        // input index -> device batch index or -1
    std::vector<int> in_index_;         // device batch index -> input index
    std::vector<uint32_t> gates_;       // three device images per device image whose state flags decide about its pixels
    std::vector<ImageDev> imgs_;        // host copy of descriptors
    std::vector<ImageState> states_;
    struct HostOut {
        void* host_ptr;
        size_t bounce_off, bytes;
        size_t stride, tight, rows; // rows of `tight` pixel bytes, `stride` apart (all planes of a planar image together)
    };
    std::vector<HostOut> host_outs_;    // per device image; host_ptr == nullptr if output is a device buffer
    std::vector<ParsedInput> parsed_own_;
    std::vector<uint8_t> direct_;            // per device image: copied straight from the caller's (pinned) buffer
    std::vector<const uint8_t*> src_scan_;   // per device image: first byte of its entropy-coded segment in the caller's buffer
    size_t staged_bytes_ = 0, off_raw_ = 0, upload_bytes_ = 0;
    bool staged_ = false;
    bool prepared_ = false, executed_ = false, finish_queued_ = false, any_host_out_ = false;

    // device memory
    DeviceBuffer upload_, scratch_, cmp_, coef_, sym_, pcoef_, planes_, bounce_, scan_scratch_;
    size_t pcoef_bytes_ = 0;
    PinnedBuffer upload_host_, state_host_, scan_host_;
    size_t device_bytes_ = 0;
    size_t coef_bytes_ = 0;
    BatchDev bd_{};
    cudaEvent_t entropy_event_ = nullptr; // behind the state read-back that follows the entropy stage
    cudaEvent_t done_event_ = nullptr;
    cudaStream_t side_stream_ = nullptr;  // carries the state read-back so that it does not sit between the kernels
    cudaStream_t last_stream_ = nullptr;
    bool have_done_ = false;  // done_event_ was recorded for the last batch
    bool inflight_ = false;   // a prepared batch has not been waited for yet (its uploads may still read the pinned staging)

    // huffman set cache, keyed by the canonical bytes of the table specs (huff_set_key)
    static constexpr size_t kSetCacheMax = 4096;
    std::unordered_map<const HuffSetDev*, int> set_index_; // per batch: cache node -> index in the upload
    std::unordered_map<std::string, HuffSetDev> set_cache_;
    std::string set_key_;

    // stats
    double t_walk_ = 0, t_layout_ = 0, t_stage_ = 0, t_h2d_ = 0; // host phases of the last prepare (B200JPEG_TRACE)
  public:
    void trace_times(double* walk, double* layout, double* stage, double* h2d) const
    {
        *walk = t_walk_;
        *layout = t_layout_;
        *stage = t_stage_;
        *h2d = t_h2d_;
    }
  private:
    int64_t launches_ = 0, algo_bytes_ = 0, comp_bytes_ = 0, n_coefs_ = 0, out_bytes_ = 0, resync_images_ = 0;

    // profiling
    bool profiling_ = false;
    static constexpr int kStages = 8;
    cudaEvent_t ev_[kStages + 1] = {};
    bool ev_valid_ = false;
    bool ev_recorded_ = false;
};

// Chunked, overlapped decode of a whole batch (what b200jpegDecodeBatch and the extension's decodeBatch use).
class Pipeline {
  public:
    Pipeline(int device_id, const Options& opt, int n_lanes = 6, const Services& sv = Services());
    ~Pipeline();
    int decode(int n, const void* const* data, const size_t* len, const b200jpegOutput_t* outs, int* status, cudaStream_t user_stream,
        const ParsedInput* const* parsed = nullptr);
    Decoder& lane0() { return *lanes_[0].dec; } // Prepare / Execute / Finish and the kernel-level entry points use this one
    int64_t last_launches() const { return last_launches_; }
    int64_t last_resync() const { return last_resync_; }
    const char* last_error() const { return err_.c_str(); }
    int device() const { return device_; }

  private:
    struct Lane {
        std::unique_ptr<Decoder> dec;
        cudaStream_t stream = nullptr;
        cudaStream_t active_stream = nullptr;
        int first = 0;
        bool busy = false;
        bool used = false; // took part in the current call (its done event gates the caller's stream)
        bool staged = true; // (under st_mu_) the background staging of its current chunk has finished
    };
    // Staging of pageable bitstreams runs on a background thread (which fans the copies out over the helper threads): the
    // calling thread meanwhile lays out the next chunk and uploads / launches the previous one.
    void stager_main();
    void submit_stage(Lane& l);
    void wait_staged(Lane& l);
    std::thread stager_;
    std::mutex st_mu_;
    std::condition_variable st_cv_, st_done_cv_;
    std::deque<Lane*> st_queue_;
    bool st_stop_ = false;

    int device_;
    Options opt_;
    Services sv_;
    std::unique_ptr<HostPool> pool_;
    std::vector<Lane> lanes_;
    cudaEvent_t user_event_ = nullptr;
    int64_t last_launches_ = 0, last_resync_ = 0;
    std::string err_;
};

} // namespace b200jpeg
