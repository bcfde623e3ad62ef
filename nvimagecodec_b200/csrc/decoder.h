// Batch orchestration of the b200 JPEG decode path: host marker walk, one staging
// upload per batch, kernel sequence, the rare re-synchronisation path, statuses.
// Mirrors the role of DecoderImpl in the reference's batched plugin
// (extensions/nvjpeg/hw_decoder.cpp:636-888) for a single device.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>
#include <memory>
#include <string>
#include <unordered_map>
#include <vector>

#include "b200jpeg.h"
#include "jpeg_stream.h"
#include "kernels.cuh"

namespace b200jpeg {

struct DeviceBuffer {
    void* ptr = nullptr;
    size_t cap = 0;
    bool ensure(size_t bytes, size_t* total); // grow-only
    void release();
    template <typename T>
    T* as() const { return static_cast<T*>(ptr); }
};

struct PinnedBuffer {
    void* ptr = nullptr;
    size_t cap = 0;
    bool ensure(size_t bytes);
    void release();
    template <typename T>
    T* as() const { return static_cast<T*>(ptr); }
};

struct Options {
    bool fancy_upsampling = true; // libjpeg_turbo_ext default (libjpeg_turbo_decoder.cpp:223-249)
    bool debug_no_warm = false;   // test hook: forces the re-synchronisation path
    int max_fix_rounds = 1 << 20;
    int subseq_bits = 0;          // 0 = choose per batch; else 1024 / 2048 / 4096
    int chunk_mb = 0;             // pipeline chunk: this much bitstream (0 = 8 MB for pinned sources, 24 MB for pageable ones, which are staged by a few threads) ...
    int direct_kb = 128;          // pinned bitstreams of at least this many KiB are copied straight from the caller's buffer
    int lanes = 6;                // pipeline lanes (decoder + stream each) the chunks rotate over
    int chunk_images = 64;        // ... or this many images, whichever comes first
};

Options parse_options(const char* options, const char* decoder_id);

class Decoder {
  public:
    Decoder(int device_id, const Options& opt);
    ~Decoder();

    int prepare(int n, const void* const* data, const size_t* len, const b200jpegOutput_t* outs, cudaStream_t stream);
    int execute(cudaStream_t stream);
    int finish(int* status, cudaStream_t stream);
    int finish_async(cudaStream_t stream);          // queue read-backs / host-output copies
    int finish_wait(int* status, cudaStream_t stream); // wait, fix-up path, statuses

    int64_t stat(int which) const;
    void set_profiling(bool on) { profiling_ = on; }
    int stage_times(float* ms, const char** names, int cap);
    int copy_coefficients(int index, int16_t* host_out, size_t cap_elems);
    int scan_segment(const void* dev_raw, size_t length, uint32_t expected_rst, void* dev_words, uint32_t* dev_bnd, uint32_t bnd_cap,
        void* dev_result, cudaStream_t stream);
    const char* last_error() const { return err_.c_str(); }
    int device() const { return device_; }

  private:
    int fail(int code, const std::string& msg);
    bool cuda_ok(cudaError_t e, const char* what);
    void mark(int stage, cudaStream_t s);

    int device_;
    Options opt_;
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
    };
    std::vector<HostOut> host_outs_;    // per device image; host_ptr == nullptr if output is a device buffer
    bool prepared_ = false, executed_ = false, finish_queued_ = false;

    // device memory
    DeviceBuffer upload_, scratch_, cmp_, coef_, planes_, bounce_, scan_scratch_;
    PinnedBuffer upload_host_, state_host_, scan_host_;
    size_t device_bytes_ = 0;
    size_t coef_bytes_ = 0;
    BatchDev bd_{};

    // huffman set cache (content keyed by hash)
    std::unordered_map<uint64_t, int> set_index_; // per batch
    std::unordered_map<uint64_t, HuffSetDev> set_cache_;

    // stats
    int64_t launches_ = 0, algo_bytes_ = 0, comp_bytes_ = 0, n_coefs_ = 0, out_bytes_ = 0, resync_images_ = 0;

    // profiling
    bool profiling_ = false;
    static constexpr int kStages = 7;
    cudaEvent_t ev_[kStages + 1] = {};
    bool ev_valid_ = false;
    bool ev_recorded_ = false;
};

// Chunked, overlapped decode of a whole batch (what b200jpegDecodeBatch and the extension's decodeBatch use).
class Pipeline {
  public:
    Pipeline(int device_id, const Options& opt, int n_lanes = 6);
    ~Pipeline();
    int decode(int n, const void* const* data, const size_t* len, const b200jpegOutput_t* outs, int* status, cudaStream_t user_stream);
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
    };
    int device_;
    Options opt_;
    std::vector<Lane> lanes_;
    cudaEvent_t user_event_ = nullptr;
    int64_t last_launches_ = 0, last_resync_ = 0;
    std::string err_;
};

} // namespace b200jpeg
