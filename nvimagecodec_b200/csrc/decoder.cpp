// Batch orchestration (see decoder.h).  Host work per batch: marker walk per image, one
// pinned staging arena, ONE host->device copy, a fixed kernel sequence, one status read-back.
#include "decoder.h"

#include <nvtx3/nvtx3.hpp>
#if defined(__SSE2__)
#include <emmintrin.h>
#endif
#include <sched.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdlib>
#include <condition_variable>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <thread>

namespace b200jpeg {

namespace {
inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }
const char* kStageNames[] = {"scan(K0)", "huff_sync(K1)", "verify_scan(K2)", "huff_write(K3)", "dc_scan(K3b)", "idct(K4a)", "color(K4b)", "pixels(K4)"};
} // namespace

bool trace_enabled()
{
    static const bool on = std::getenv("B200JPEG_TRACE") != nullptr;
    return on;
}
double trace_now_us()
{
    return std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

// ---- memory: the host's allocators when it supplied some, else stream-ordered allocation (cudaMallocAsync) where the
// device has memory pools, else cudaMalloc -- the order src/imgproc/device_buffer.cpp:42-69 uses.
namespace {
bool device_has_pools()
{
    static const bool off = std::getenv("B200JPEG_NO_POOL") != nullptr; // experiments: plain cudaMalloc instead of stream-ordered allocation
    if (off)
        return false;
    static int cached[64];
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64)
        return false;
    if (cached[dev] == 0) {
        int v = 0;
        cudaDeviceGetAttribute(&v, cudaDevAttrMemoryPoolsSupported, dev);
        cudaGetLastError();
        cached[dev] = v ? 2 : 1;
    }
    return cached[dev] == 2;
}
void device_block_free(const Services& sv, void* ptr, size_t cap, cudaStream_t s, bool teardown)
{
    if (!ptr)
        return;
    if (sv.device_malloc && sv.device_free)
        sv.device_free(sv.device_ctx, ptr, cap, s);
    else if (!teardown && device_has_pools())
        cudaFreeAsync(ptr, s);
    else
        cudaFree(ptr); // also valid for stream-ordered allocations; synchronises, which is what teardown wants
}
} // namespace

bool DeviceBuffer::ensure(size_t bytes, size_t* total, const Services& sv, cudaStream_t s)
{
    if (bytes <= cap) {
        // an outlier batch must not pin HBM for good: after a run of batches that need less than a quarter of a
        // large block it is given back and re-allocated at the size in use
        idle = (cap > (size_t(256) << 20) && bytes < cap / 4) ? idle + 1 : 0;
        if (idle < 16)
            return true;
    }
    if (ptr) {
        device_block_free(sv, ptr, cap, s, false);
        *total -= cap;
        ptr = nullptr;
        cap = 0;
    }
    idle = 0;
    const size_t want = align_up(bytes + bytes / 8, size_t(2) << 20);
    bool ok;
    if (sv.device_malloc && sv.device_free)
        ok = sv.device_malloc(sv.device_ctx, &ptr, want, s) == 0 && ptr;
    else if (device_has_pools())
        ok = cudaMallocAsync(&ptr, want, s) == cudaSuccess;
    else
        ok = cudaMalloc(&ptr, want) == cudaSuccess;
    if (!ok) {
        ptr = nullptr;
        cudaGetLastError();
        return false;
    }
    cap = want;
    stream = s;
    *total += cap;
    return true;
}
void DeviceBuffer::release(const Services& sv)
{
    device_block_free(sv, ptr, cap, nullptr, true);
    ptr = nullptr;
    cap = 0;
}
bool PinnedBuffer::ensure(size_t bytes, const Services& sv, cudaStream_t s)
{
    if (bytes <= cap)
        return true;
    release(sv);
    const size_t want = align_up(bytes + bytes / 8, size_t(2) << 20);
    bool ok;
    if (sv.pinned_malloc && sv.pinned_free)
        ok = sv.pinned_malloc(sv.pinned_ctx, &ptr, want, s) == 0 && ptr;
    else
        ok = cudaMallocHost(&ptr, want) == cudaSuccess;
    if (!ok) {
        ptr = nullptr;
        cudaGetLastError();
        return false;
    }
    cap = want;
    return true;
}
void PinnedBuffer::release(const Services& sv)
{
    if (ptr) {
        if (sv.pinned_malloc && sv.pinned_free)
            sv.pinned_free(sv.pinned_ctx, ptr, cap, nullptr);
        else
            cudaFreeHost(ptr);
    }
    ptr = nullptr;
    cap = 0;
}

// ---- helper threads
struct HostPool::Impl {
    std::mutex run_mu; // one job at a time (callers on different threads take turns)
    std::mutex mu;
    std::condition_variable cv_work, cv_done;
    void (*fn)(int, void*) = nullptr;
    void* arg = nullptr;
    int n = 0;
    std::atomic<int> next{0};
    int pending = 0;      // helpers still inside the current job
    uint64_t job = 0;
    bool stop = false;
};

HostPool::HostPool(int n_threads)
    : impl_(new Impl)
{
    for (int t = 0; t < n_threads; ++t)
        threads_.emplace_back([this]() {
            Impl& m = *impl_;
            uint64_t seen = 0;
            for (;;) {
                std::unique_lock<std::mutex> lk(m.mu);
                m.cv_work.wait(lk, [&] { return m.stop || m.job != seen; });
                if (m.stop)
                    return;
                seen = m.job;
                auto fn = m.fn;
                void* arg = m.arg;
                const int n = m.n;
                lk.unlock();
                for (int i; (i = m.next.fetch_add(1)) < n;)
                    fn(i, arg);
                lk.lock();
                if (--m.pending == 0)
                    m.cv_done.notify_one();
            }
        });
}

HostPool::~HostPool()
{
    {
        std::lock_guard<std::mutex> lk(impl_->mu);
        impl_->stop = true;
    }
    impl_->cv_work.notify_all();
    for (auto& t : threads_)
        t.join();
    delete impl_;
}

void HostPool::run(int n, void (*fn)(int, void*), void* arg)
{
    Impl& m = *impl_;
    if (n <= 1 || threads_.empty()) {
        for (int i = 0; i < n; ++i)
            fn(i, arg);
        return;
    }
    std::lock_guard<std::mutex> turn(m.run_mu);
    {
        std::lock_guard<std::mutex> lk(m.mu);
        m.fn = fn;
        m.arg = arg;
        m.n = n;
        m.next.store(0);
        m.pending = int(threads_.size());
        ++m.job;
    }
    m.cv_work.notify_all();
    for (int i; (i = m.next.fetch_add(1)) < n;)
        fn(i, arg);
    std::unique_lock<std::mutex> lk(m.mu);
    m.cv_done.wait(lk, [&] { return m.pending == 0; });
}

// "<decoder_id>:<name>=<value>" tokens separated by spaces; an empty id addresses every decoder
// (include/nvimgcodec.h:1967-1971; same idiom as libjpeg_turbo_decoder.cpp:229-248).
Options parse_options(const char* options, const char* decoder_id)
{
    Options o;
    if (!options)
        return o;
    std::string s(options);
    size_t pos = 0;
    while (pos < s.size()) {
        size_t end = s.find(' ', pos);
        if (end == std::string::npos)
            end = s.size();
        std::string tok = s.substr(pos, end - pos);
        pos = end + 1;
        if (tok.empty())
            continue;
        std::string module, name = tok, value;
        const size_t colon = tok.find(':');
        if (colon != std::string::npos) {
            module = tok.substr(0, colon);
            name = tok.substr(colon + 1);
        }
        const size_t eq = name.find('=');
        if (eq == std::string::npos)
            continue;
        value = name.substr(eq + 1);
        name = name.substr(0, eq);
        if (!module.empty() && module != decoder_id)
            continue;
        if (name == "fancy_upsampling")
            o.fancy_upsampling = std::atoi(value.c_str()) != 0;
        else if (name == "debug_no_warm")
            o.debug_no_warm = std::atoi(value.c_str()) != 0;
        else if (name == "chunk_mb")
            o.chunk_mb = std::max(0, std::atoi(value.c_str()));
        else if (name == "direct_kb")
            o.direct_kb = std::max(1, std::atoi(value.c_str()));
        else if (name == "lanes")
            o.lanes = std::min(8, std::max(1, std::atoi(value.c_str())));
        else if (name == "chunk_images")
            o.chunk_images = std::max(1, std::atoi(value.c_str()));
        else if (name == "entry_stream")
            o.entry_stream = std::atoi(value.c_str()) != 0;
        else if (name == "progressive")
            o.progressive = std::atoi(value.c_str()) != 0;
        else if (name == "progressive_min_batch")
            o.progressive_min_batch = std::max(1, std::atoi(value.c_str()));
        else if (name == "fused_pixels")
            o.fused_pixels = std::atoi(value.c_str()) != 0;
        else if (name == "use_executor")
            o.use_executor = std::atoi(value.c_str()) != 0;
        else if (name == "sync_finish")
            o.sync_finish = std::atoi(value.c_str()) != 0;
        else if (name == "host_threads")
            o.host_threads = std::min(64, std::max(1, std::atoi(value.c_str())));
        else if (name == "max_arena_mb")
            o.max_arena_mb = std::max(1, std::atoi(value.c_str()));
        else if (name == "subseq_bits") {
            const int v = std::atoi(value.c_str());
            if (v == 512 || v == 1024 || v == 2048 || v == 4096)
                o.subseq_bits = v;
        } else if (name == "max_fix_rounds")
            o.max_fix_rounds = std::atoi(value.c_str());
    }
    return o;
}

Decoder::Decoder(int device_id, const Options& opt, const Services& sv, HostPool* pool)
    : device_(device_id)
    , opt_(opt)
    , sv_(sv)
    , pool_(pool)
{
}

Decoder::~Decoder()
{
    cudaSetDevice(device_);
    cudaDeviceSynchronize(); // pixel kernels of the last batch may still be running on a stream we no longer know to be alive
    if (ev_valid_)
        for (auto& e : ev_)
            cudaEventDestroy(e);
    if (entropy_event_)
        cudaEventDestroy(entropy_event_);
    if (done_event_)
        cudaEventDestroy(done_event_);
    if (side_stream_)
        cudaStreamDestroy(side_stream_);
    for (DeviceBuffer* b : {&upload_, &scratch_, &cmp_, &coef_, &sym_, &pcoef_, &planes_, &bounce_, &scan_scratch_})
        b->release(sv_);
    for (PinnedBuffer* b : {&scan_host_, &upload_host_, &state_host_})
        b->release(sv_);
}

int Decoder::fail(int code, const std::string& msg)
{
    err_ = msg;
    return code;
}

bool Decoder::cuda_ok(cudaError_t e, const char* what)
{
    if (e == cudaSuccess)
        return true;
    err_ = std::string(what) + ": " + cudaGetErrorString(e);
    return false;
}

void Decoder::mark(int stage, cudaStream_t s)
{
    if (profiling_ && ev_valid_)
        cudaEventRecord(ev_[stage], s);
}

void Decoder::parallel_for(int n, void (*fn)(int, void*), void* arg)
{
    if (sv_.parallel_for)
        sv_.parallel_for(sv_.par_ctx, n, fn, arg);
    else if (pool_)
        pool_->run(n, fn, arg);
    else
        for (int i = 0; i < n; ++i)
            fn(i, arg);
}

// The marker walk of one input (both walks for a stream that comes in three scans).
void parse_input(const void* data, size_t len, ParsedInput* out)
{
    out->scans.reset();
    out->multi_status = kStreamOk;
    out->frame = StreamHeader(); // entries are reused across batches: nothing of the previous stream (restart interval, tables) may survive
    out->status = data ? parse_stream(static_cast<const uint8_t*>(data), len, &out->frame) : kStreamNotJpeg;
    out->prog.reset();
    if (out->status == kStreamMultiScan) {
        out->scans.reset(new StreamHeader[3]);
        out->multi_status = parse_multiscan(static_cast<const uint8_t*>(data), len, &out->frame, out->scans.get());
    } else if (out->status == kStreamProgressive) {
        out->prog.reset(new ProgStream);
        out->multi_status = parse_progressive(static_cast<const uint8_t*>(data), len, &out->frame, out->prog.get());
    }
}

int Decoder::prepare(int n, const void* const* data, const size_t* len, const b200jpegOutput_t* outs, cudaStream_t stream,
    const ParsedInput* const* parsed, int sources_pinned, bool defer_upload)
{
    if (n < 0 || (n > 0 && (!data || !len || !outs)))
        return fail(B200JPEG_ERR_ARG, "null argument");
    if (!cuda_ok(cudaSetDevice(device_), "cudaSetDevice"))
        return B200JPEG_ERR_CUDA;
    // The arenas and the pinned staging block are reused.  A batch that was never waited for may still be uploading out
    // of the staging block; pixel kernels of the last batch may still be running on another stream.
    if (inflight_ && !cuda_ok(cudaStreamSynchronize(last_stream_), "synchronize"))
        return B200JPEG_ERR_CUDA;
    inflight_ = false;
    if (have_done_ && last_stream_ != stream)
        cudaStreamWaitEvent(stream, done_event_, 0);
    if (!entropy_event_) {
        if (!cuda_ok(cudaEventCreateWithFlags(&entropy_event_, cudaEventDisableTiming), "event") ||
            !cuda_ok(cudaEventCreateWithFlags(&done_event_, cudaEventDisableTiming), "event") ||
            !cuda_ok(cudaStreamCreateWithFlags(&side_stream_, cudaStreamNonBlocking), "stream"))
            return B200JPEG_ERR_CUDA;
    }
    nvtx3::scoped_range nvtx_prepare{"b200jpeg prepare (marker walk, layout, staging, H2D)"};
    prepared_ = executed_ = false;
    n_in_ = n;
    host_status_.assign(n, B200JPEG_OK);
    dev_index_.assign(n, -1);
    in_index_.clear();
    gates_.clear();
    imgs_.clear();
    host_outs_.clear();
    set_index_.clear();
    any_host_out_ = false;
    if (set_cache_.size() > kSetCacheMax) // never while a batch is being built: `sets` points into the cache
        set_cache_.clear();
    std::vector<const HuffSetDev*> sets;
    std::vector<const StreamHeader*> hdrs;
    hdrs.reserve(n);
    std::vector<size_t> item_of_dev; // device image -> its entry in hdrs

    // Subsequence length for this batch: longer subsequences cost fewer re-synchronisation rounds
    // (the block-in-MCU phase of a 4:2:0 stream needs a few MCUs to lock), shorter ones give more
    // threads; pick the longest that still fills the GPU.
    {
        uint64_t total_bits = 0;
        for (int i = 0; i < n; ++i)
            total_bits += uint64_t(len[i]) * 8;
        // Measured on B200 (profiles/README.md): full-size batches decode fastest with 2048-bit subsequences,
        // smaller ones with shorter subsequences, i.e. more threads, until the GPU is filled again.
        uint32_t sb = 512;
        if (opt_.subseq_bits)
            sb = uint32_t(opt_.subseq_bits);
        else if (total_bits >= (uint64_t(1600) << 20))
            sb = 4096;
        else if (total_bits >= (uint64_t(400) << 20))
            sb = 2048;
        else if (total_bits >= (uint64_t(48) << 20))
            sb = 1024;
        bd_.sub_bits = sb;
        bd_.seq_warm = std::max<uint32_t>(kSeqWarmMin, kSeqWarmBits / sb);
        bd_.sym_mode = opt_.entry_stream ? 1 : 0;
        bd_.sym_chunks = sym_chunks_for(sb);
        bd_.sub_shift = 0;
        while ((1u << bd_.sub_shift) < sb)
            ++bd_.sub_shift;
    }
    const size_t sub_bytes = bd_.sub_bits / 8;
    size_t raw_bytes = 0, cmp_bytes = 0, coef_elems = 0, plane_bytes = 0, bounce_bytes = 0, sym_chunks_total = 0;
    uint32_t n_tiles = 0, n_sub = 0, n_seq = 0, n_wseq = 0, n_bnd = 0, n_dc = 0, n_idct = 0, n_color = 0;
    comp_bytes_ = n_coefs_ = out_bytes_ = 0;

    const double t0 = trace_enabled() ? trace_now_us() : 0;
    // Marker walks: one per input the caller has not walked already, spread over the helper threads when there are many.
    parsed_own_.resize(size_t(n));
    {
        nvtx3::scoped_range nvtx_walk{"b200jpeg marker walk"};
        struct Ctx {
            const void* const* data;
            const size_t* len;
            const ParsedInput* const* given;
            ParsedInput* own;
        } ctx{data, len, parsed, parsed_own_.data()};
        auto walk = [](int i, void* a) {
            Ctx& c = *static_cast<Ctx*>(a);
            if (!(c.given && c.given[i]))
                parse_input(c.data[i], c.len[i], &c.own[i]);
        };
        int to_walk = 0;
        for (int i = 0; i < n; ++i)
            to_walk += (parsed && parsed[i]) ? 0 : 1;
        if (to_walk >= 16)
            parallel_for(n, walk, &ctx);
        else if (to_walk)
            for (int i = 0; i < n; ++i)
                walk(i, &ctx);
    }

    const double t1 = trace_enabled() ? trace_now_us() : 0;
    // Work list: one entry per device image.  A stream whose three components come in three scans becomes up to three
    // one-component images (kChild; each decodes like a grayscale image of the component's size) plus the image itself
    // (kParent: no entropy data of its own, its colour stage reads the children's planes).
    // A progressive stream becomes one image per scan (kScan: K0 only -- unstuffed words and restart boundaries for the
    // progressive kernels) plus the image itself (kProg: coefficient blocks, planes, pixel stage).
    enum { kNormal = 0, kChild = 1, kParent = 2, kScan = 3, kProg = 4 };
    struct Item {
        int i, role, scan;
    };
    std::vector<Item> items;
    items.reserve(n);
    auto status_of = [](StreamStatus ss) {
        return ss == kStreamNotJpeg ? B200JPEG_ERR_NOT_JPEG : (ss == kStreamUnsupported ? B200JPEG_ERR_UNSUPPORTED : B200JPEG_ERR_CORRUPT);
    };
    b200jpegOutput_t child_out;
    std::memset(&child_out, 0, sizeof(child_out));
    child_out.buffer = &child_out; // never written: a child has no colour stage
    child_out.format = B200JPEG_FMT_Y;
    std::vector<StreamHeader> parent_frames; // frames of multi-scan streams with restart_interval cleared (the parent has no scan)
    parent_frames.reserve(n);
    int n_progressive = 0;
    for (int i = 0; i < n; ++i)
        n_progressive += ((parsed && parsed[i]) ? *parsed[i] : parsed_own_[size_t(i)]).status == kStreamProgressive ? 1 : 0;
    for (int i = 0; i < n; ++i) {
        const ParsedInput& pi = (parsed && parsed[i]) ? *parsed[i] : parsed_own_[size_t(i)];
        if (pi.status == kStreamOk) {
            hdrs.push_back(&pi.frame);
            items.push_back({i, kNormal, 0});
        } else if (pi.status == kStreamMultiScan) {
            if (pi.multi_status != kStreamOk || !pi.scans) {
                host_status_[i] = status_of(pi.multi_status);
                continue;
            }
            const int n_child = outs[i].format == B200JPEG_FMT_Y ? 1 : 3; // Y output consumes luma only
            for (int c = 0; c < n_child; ++c) {
                hdrs.push_back(&pi.scans[c]);
                items.push_back({i, kChild, 0});
            }
            parent_frames.push_back(pi.frame);
            parent_frames.back().restart_interval = 0;
            hdrs.push_back(&parent_frames.back());
            items.push_back({i, kParent, 0});
        } else if (pi.status == kStreamProgressive) {
            if (!opt_.progressive || n_progressive < opt_.progressive_min_batch) {
                host_status_[i] = B200JPEG_ERR_UNSUPPORTED;
                continue;
            }
            if (pi.multi_status != kStreamOk || !pi.prog) {
                host_status_[i] = status_of(pi.multi_status);
                continue;
            }
            for (size_t sidx = 0; sidx < pi.prog->scans.size(); ++sidx) {
                hdrs.push_back(&pi.frame);
                items.push_back({i, kScan, int(sidx)});
            }
            hdrs.push_back(&pi.frame);
            items.push_back({i, kProg, 0});
        } else {
            host_status_[i] = status_of(pi.status);
        }
    }
    int children_ok = 0; // children of the multi-scan stream under way that made it into the batch
    std::vector<ProgScanDev> prog_scans;
    std::vector<ProgHuffDev> prog_tables;
    std::vector<uint32_t> prog_imgs;
    std::vector<uint32_t> prog_chain_scan; // chain -> scan, image by image, scans in file order (predecessors first)
    uint32_t prog_max_blocks = 0;
    size_t pcoef_elems = 0;
    std::vector<size_t> scan_offs; // per device image: where its entropy-coded bytes start in the input
    scan_offs.reserve(items.size());
    for (size_t item = 0; item < items.size(); ++item) {
        const int i = items[item].i, role = items[item].role;
        const StreamHeader& hd = *hdrs[item];
        if ((role == kChild || role == kScan) && (item == 0 || items[item - 1].role != role || items[item - 1].i != i))
            children_ok = 0;
        if (role == kScan) {
            // one scan of a progressive stream: what K0 needs, nothing else
            const ParsedInput& pi = (parsed && parsed[i]) ? *parsed[i] : parsed_own_[size_t(i)];
            const ProgScan& ps = pi.prog->scans[size_t(items[item].scan)];
            if (ps.length >= (size_t(1) << 28)) {
                host_status_[i] = B200JPEG_ERR_TOO_LARGE;
                continue;
            }
            ImageDev im;
            std::memset(&im, 0, sizeof(im));
            im.raw_off = raw_bytes;
            im.raw_len = uint32_t(ps.length);
            raw_bytes += align_up(ps.length, 16) + 16;
            im.cmp_off = cmp_bytes;
            cmp_bytes += align_up(ps.length, 16) + 32;
            im.tile_off = n_tiles;
            im.n_tiles = uint32_t((ps.length + kScanTileBytes - 1) / kScanTileBytes);
            n_tiles += im.n_tiles;
            im.sub_off = n_sub;
            im.seq_off = n_seq;
            im.wseq_off = n_wseq;
            im.restart_interval = uint32_t(ps.restart_interval);
            im.n_rst_expected = ps.restart_interval ? uint32_t((ps.mcus + ps.restart_interval - 1) / ps.restart_interval - 1) : 0u;
            im.bnd_off = n_bnd;
            im.n_bnd_cap = im.n_rst_expected + 1;
            n_bnd += im.n_bnd_cap;
            im.coef_off = coef_elems;
            im.sym_off = sym_chunks_total;
            im.dc_cta_off = n_dc;
            im.idct_cta_off = n_idct;
            im.color_cta_off = n_color;
            im.ncomp = im.ncomp_needed = 1;
            const uint32_t self = uint32_t(imgs_.size());
            const uint32_t gate_self[3] = {self, self, self};
            gates_.insert(gates_.end(), gate_self, gate_self + 3);
            in_index_.push_back(i);
            item_of_dev.push_back(item);
            scan_offs.push_back(ps.offset);
            imgs_.push_back(im);
            host_outs_.push_back(HostOut{nullptr, 0, 0, 0, 0, 0});
            ++children_ok;
            continue;
        }
        const b200jpegOutput_t& o = role == kChild ? child_out : outs[i];
        if (role == kProg) {
            const ParsedInput& pi = (parsed && parsed[i]) ? *parsed[i] : parsed_own_[size_t(i)];
            if (host_status_[i] != B200JPEG_OK || children_ok != int(pi.prog->scans.size())) {
                if (host_status_[i] == B200JPEG_OK)
                    host_status_[i] = B200JPEG_ERR_UNSUPPORTED;
                continue;
            }
        }
        if (role == kParent) {
            const int want = o.format == B200JPEG_FMT_Y ? 1 : 3;
            if (host_status_[i] != B200JPEG_OK || children_ok != want) {
                if (host_status_[i] == B200JPEG_OK)
                    host_status_[i] = B200JPEG_ERR_UNSUPPORTED;
                continue;
            }
        }
        int x0 = o.roi_x0, y0 = o.roi_y0, x1 = o.roi_x1, y1 = o.roi_y1;
        if (x0 == 0 && y0 == 0 && x1 == 0 && y1 == 0) {
            x1 = hd.width;
            y1 = hd.height;
        }
        if (!o.buffer || o.format < 0 || o.format > B200JPEG_FMT_P_YUV || x0 < 0 || y0 < 0 || x1 > hd.width || y1 > hd.height || x0 >= x1 || y0 >= y1) {
            host_status_[i] = B200JPEG_ERR_ARG;
            continue;
        }
        const size_t rw = size_t(x1 - x0), rh = size_t(y1 - y0);
        const bool planar = (o.format == B200JPEG_FMT_P_RGB || o.format == B200JPEG_FMT_P_BGR || o.format == B200JPEG_FMT_P_YUV);
        const size_t nch = o.format == B200JPEG_FMT_Y ? 1 : 3;
        if (o.orientation < 0 || o.orientation > 8) {
            host_status_[i] = B200JPEG_ERR_ARG;
            continue;
        }
        const bool transposed = o.orientation >= 5;       // EXIF 5..8: output rows run along the source columns
        const size_t out_w = transposed ? rh : rw, out_h = transposed ? rw : rh;
        const size_t tight = planar ? out_w : out_w * nch;
        const size_t stride = o.row_stride ? o.row_stride : tight;
        if (stride < tight || stride > 0xFFFFFFFFull) {
            host_status_[i] = B200JPEG_ERR_ARG;
            continue;
        }
        if (hd.scan_length >= (size_t(1) << 28)) { // bit positions (and the bit after the last subsequence) must fit 32 bits
            host_status_[i] = B200JPEG_ERR_TOO_LARGE;
            continue;
        }
        if (role == kProg) {
            // Same reasoning as below: a first DC scan spends at least one bit per block.
            const ParsedInput& pi = (parsed && parsed[i]) ? *parsed[i] : parsed_own_[size_t(i)];
            const uint64_t blocks = uint64_t(hd.mcus_x) * uint64_t(hd.mcus_y) * uint64_t(hd.blocks_per_mcu);
            uint64_t scan_bytes = 0;
            for (const ProgScan& ps : pi.prog->scans)
                scan_bytes += ps.length;
            if (scan_bytes * 8 < blocks) {
                host_status_[i] = B200JPEG_ERR_CORRUPT;
                continue;
            }
            uint64_t chains = 0; // one warp (and a few words of bookkeeping) per restart interval of every scan
            for (const ProgScan& ps : pi.prog->scans)
                chains += ps.restart_interval ? uint64_t((ps.mcus + ps.restart_interval - 1) / ps.restart_interval) : 1u;
            if (blocks * 192 > (uint64_t(opt_.max_arena_mb) << 20) || coef_elems / 64 + blocks >= (uint64_t(1) << 28) || chains > (uint64_t(1) << 20)) {
                host_status_[i] = B200JPEG_ERR_TOO_LARGE;
                continue;
            }
        } else if (role != kParent) {
            // The arenas are sized from the frame header, which costs an attacker nothing to inflate.  Every coded block
            // takes at least two bits (a DC code and an end-of-block code), so a scan shorter than that cannot hold the
            // blocks the header claims: corrupt, before anything is allocated.  What remains is bounded per image
            // (max_arena_mb) and by the 32-bit block / group indices of the kernels.
            const uint64_t blocks = uint64_t(hd.mcus_x) * uint64_t(hd.mcus_y) * uint64_t(hd.blocks_per_mcu);
            if (uint64_t(hd.scan_length) * 8 < 2 * blocks) {
                host_status_[i] = B200JPEG_ERR_CORRUPT;
                continue;
            }
            // coefficients (dense groups, or the entry stream: 36 bytes reserved per byte of scan) + planes, upper bound
            const uint64_t arena = (bd_.sym_mode ? uint64_t(hd.scan_length) * 40 : blocks * 128) + blocks * 64;
            if (arena > (uint64_t(opt_.max_arena_mb) << 20) || coef_elems / 64 + blocks >= (uint64_t(1) << 28)) {
                host_status_[i] = B200JPEG_ERR_TOO_LARGE;
                continue;
            }
        }
        ImageDev im;
        std::memset(&im, 0, sizeof(im));
        uint32_t gate[3] = {0, 0, 0};
        const size_t plane_bytes_before = plane_bytes;
        im.raw_off = raw_bytes;
        im.raw_len = uint32_t(hd.scan_length);
        raw_bytes += align_up(hd.scan_length, 16) + 16;
        im.cmp_off = cmp_bytes;
        cmp_bytes += align_up(hd.scan_length, 16) + 32;
        im.tile_off = n_tiles;
        im.n_tiles = uint32_t((hd.scan_length + kScanTileBytes - 1) / kScanTileBytes);
        n_tiles += im.n_tiles;
        im.sub_off = n_sub;
        im.n_sub_max = uint32_t((hd.scan_length + sub_bytes - 1) / sub_bytes);
        n_sub += im.n_sub_max;
        im.seq_off = n_seq;
        im.n_seq = (im.n_sub_max + (kSeqLogical - bd_.seq_warm) - 1) / (kSeqLogical - bd_.seq_warm);
        n_seq += im.n_seq;
        im.wseq_off = n_wseq;
        im.n_wseq = (im.n_sub_max + kSeqThreads - 1) / kSeqThreads;
        n_wseq += im.n_wseq;
        const uint32_t n_mcu = uint32_t(hd.mcus_x) * uint32_t(hd.mcus_y);
        im.restart_interval = uint32_t(hd.restart_interval);
        im.n_rst_expected = hd.restart_interval ? (n_mcu + hd.restart_interval - 1) / hd.restart_interval - 1 : 0;
        im.bnd_off = n_bnd;
        im.n_bnd_cap = im.n_rst_expected + 1;
        n_bnd += im.n_bnd_cap;
        im.total_blocks = role == kParent ? 0u : n_mcu * uint32_t(hd.blocks_per_mcu);
        im.coef_off = coef_elems;
        coef_elems += size_t(im.total_blocks) * 64;
        im.sym_off = sym_chunks_total; // entry stream: the chunks of 32 subsequences interleave
        if (role != kParent)
            sym_chunks_total += size_t((im.n_sub_max + 31) / 32) * bd_.sym_chunks * 32;
        im.width = uint16_t(hd.width);
        im.height = uint16_t(hd.height);
        im.mcus_x = uint16_t(hd.mcus_x);
        im.mcus_y = uint16_t(hd.mcus_y);
        im.ncomp = uint8_t(hd.ncomp);
        im.hmax = uint8_t(hd.hmax);
        im.vmax = uint8_t(hd.vmax);
        im.bpm = uint8_t(hd.blocks_per_mcu);
        im.fmt = uint8_t(o.format);
        im.fancy = opt_.fancy_upsampling ? 1 : 0;
        im.ncomp_needed = (o.format == B200JPEG_FMT_Y || hd.ncomp == 1) ? 1 : 3;
        // MCU rows the ROI needs: its own rows plus one MCU row of context on each side (the vertical
        // triangle filter looks one chroma row up / down)
        const int mcu_h = 8 * hd.vmax;
        const int mcu_row_lo = std::max(0, y0 / mcu_h - 1), mcu_row_hi = std::min(hd.mcus_y, (y1 + mcu_h - 1) / mcu_h + 1);
        im.mcu_row_lo = uint16_t(mcu_row_lo);
        im.mcu_row_hi = uint16_t(mcu_row_hi);
        im.write_blk_lo = role == kParent ? 0u : uint32_t(mcu_row_lo) * uint32_t(hd.mcus_x) * uint32_t(hd.blocks_per_mcu);
        im.write_blk_hi = role == kParent ? 0u : uint32_t(mcu_row_hi) * uint32_t(hd.mcus_x) * uint32_t(hd.blocks_per_mcu);
        {   // DC differences are needed from the last predictor reset at or before the first wanted block
            const uint32_t ri_blocks = uint32_t(hd.restart_interval) * uint32_t(hd.blocks_per_mcu);
            im.dc_blk_lo = ri_blocks ? im.write_blk_lo / ri_blocks * ri_blocks : 0u;
            im.dc_cta_off = n_dc;
            im.n_dc_cta = role == kProg ? 0u : (im.write_blk_hi - im.dc_blk_lo + kDcChunk - 1) / kDcChunk; // (progressive DC values are final)
            n_dc += im.n_dc_cta;
        }
        int b = 0;
        uint32_t idct_blocks = 0;
        for (int c = 0; c < hd.ncomp; ++c) {
            im.h[c] = uint8_t(hd.h[c]);
            im.v[c] = uint8_t(hd.v[c]);
            im.blk_first[c] = uint8_t(b);
            for (int k = 0; k < hd.h[c] * hd.v[c]; ++k, ++b) {
                im.blk_comp |= uint32_t(c) << (2 * b);
                im.blk_tab |= uint32_t(hd.td[c]) << b;
                im.blk_tab |= uint32_t(hd.ta[c]) << (16 + b);
            }
            std::memcpy(im.qt[c], hd.qt[hd.tq[c]], sizeof(im.qt[c]));
            im.plane_stride_px[c] = uint32_t(hd.mcus_x) * hd.h[c] * 8;
            im.comp_w[c] = uint16_t((hd.width * hd.h[c] + hd.hmax - 1) / hd.hmax);
            im.comp_h[c] = uint16_t((hd.height * hd.v[c] + hd.vmax - 1) / hd.vmax);
            const size_t pw = size_t(hd.mcus_x) * hd.h[c] * 8, ph = size_t(hd.mcus_y) * hd.v[c] * 8;
            if (role == kParent) {
                if (c < im.ncomp_needed) { // the planes are the children's (the last `ncomp_needed` images pushed)
                    const ImageDev& ch = imgs_[imgs_.size() - size_t(im.ncomp_needed) + size_t(c)];
                    im.plane_off[c] = ch.plane_off[0];
                    im.plane_stride_px[c] = ch.plane_stride_px[0];
                    gate[c] = uint32_t(imgs_.size() - size_t(im.ncomp_needed) + size_t(c));
                }
            } else if (c < im.ncomp_needed) {
                im.plane_off[c] = plane_bytes;
                plane_bytes += align_up(pw * ph, 256);
                idct_blocks += uint32_t(pw / 8) * uint32_t(mcu_row_hi - mcu_row_lo) * hd.v[c];
            }
        }
        im.idct_cta_off = n_idct;
        im.n_idct_cta = (idct_blocks + 128 * kIdctPer - 1) / (128 * kIdctPer);
        n_idct += im.n_idct_cta;
        im.roi_x0 = uint16_t(x0);
        im.roi_y0 = uint16_t(y0);
        im.roi_x1 = uint16_t(x1);
        im.roi_y1 = uint16_t(y1);
        im.row_stride = uint32_t(stride);
        im.plane_stride = o.plane_stride ? o.plane_stride : stride * out_h;
        if (planar && o.plane_stride && (o.plane_stride < stride * out_h || o.buffer_is_host)) {
            host_status_[i] = B200JPEG_ERR_ARG; // (host outputs are bounced tightly: no embedding there)
            continue;
        }
        im.orient = uint8_t(o.orientation);
        const size_t obytes = planar ? 3 * stride * out_h : stride * out_h;
        // K4b specialisation (kernels.cu color_kernel): fast paths need full-resolution luma and
        // both chroma components subsampled alike by 1 or 2 in each direction
        im.color_mode = 0;
        im.color_unit_rows = 1;
        const bool luma_full = hd.h[0] == hd.hmax && hd.v[0] == hd.vmax;
        if (im.ncomp_needed == 1) {
            if (luma_full)
                im.color_mode = 1;
        } else if (luma_full && hd.h[1] == hd.h[2] && hd.v[1] == hd.v[2]) {
            const int hx = hd.hmax / hd.h[1], vx = hd.vmax / hd.v[1];
            if ((hx == 1 || hx == 2) && (vx == 1 || vx == 2)) {
                im.color_mode = uint8_t(2 + (hx - 1) + 2 * (vx - 1));
                im.color_unit_rows = uint8_t(vx);
                // libjpeg's choice of the triangle filter (jdsample.c jinit_upsampler): h2v1 and h2v2 need
                // downsampled_width > 2, h1v2 does not
                const int dw = (hd.width * hd.h[1] + hd.hmax - 1) / hd.hmax;
                const bool tri = opt_.fancy_upsampling && ((hx == 2 && dw > 2) || (hx == 1 && vx == 2));
                if (tri && im.color_mode >= 3)
                    im.color_mode = uint8_t(im.color_mode + 3); // 6: h2v1, 7: h1v2, 8: h2v2 with the triangle filter
            }
        }
        {
            const int ur = im.color_unit_rows;
            // units are aligned in image coordinates: 16 pixels wide on the specialised paths, 8 on the generic one
            const size_t uw = (im.color_mode >= 1 && o.orientation <= 1) ? 16 : 8;
            const size_t ax0 = size_t(x0) & ~(uw - 1), ay0 = size_t(y0) / ur * ur;
            im.color_cta_off = n_color;
            im.n_color_cta_x = uint32_t((size_t(x1) - ax0 + 32 * uw - 1) / (32 * uw));
            im.n_color_cta_y = uint32_t((size_t(y1) - ay0 + 8 * ur * kColorUnitsY - 1) / (8 * ur * kColorUnitsY));
            im.color_cta_x_rcp = im.n_color_cta_x > 1 ? uint32_t((uint64_t(1) << 32) / im.n_color_cta_x + 1) : 0u;
            if (role == kChild)
                im.n_color_cta_x = im.n_color_cta_y = 0; // planes only
        }
        // Fused pixel kernel (K4) for the specialised sampling modes: tiles over the MCUs the region touches.  Strip width and
        // band height are set for the whole batch once every image is known (below); here only the MCU window.
        {
            im.q8 = 1;
            for (int c = 0; c < hd.ncomp; ++c)
                for (int k = 0; k < 64; ++k)
                    im.q8 = (im.q8 && hd.qt[hd.tq[c]][k] <= 255) ? 1 : 0;
            // (the fused kernel dequantises with 8-bit quantisers only; every baseline table is 8 bits wide)
            bool fused = opt_.fused_pixels && !bd_.sym_mode && role == kNormal && im.color_mode >= 1 && o.orientation <= 1 && im.q8;
            for (int c = 0; c < hd.ncomp; ++c)
                fused = fused && hd.h[c] <= 2 && hd.v[c] <= 2;
            if (fused) {
                const int mcu_w = 8 * hd.hmax;
                im.px_mcu_col_lo = uint16_t(x0 / mcu_w);
                im.px_mcu_col_hi = uint16_t((x1 + mcu_w - 1) / mcu_w);
                im.px_mcu_row_lo = uint16_t(y0 / mcu_h);
                im.px_mcu_row_hi = uint16_t((y1 + mcu_h - 1) / mcu_h);
                im.px_n_cta = 1; // marks the image; the real count follows
                // the idct / color kernels and the plane arena are not needed
                plane_bytes = plane_bytes_before;
                n_idct -= im.n_idct_cta;
                im.n_idct_cta = 0;
                im.n_color_cta_x = im.n_color_cta_y = 0;
            }
        }
        n_color += im.n_color_cta_x * im.n_color_cta_y;
        HostOut ho{nullptr, 0, obytes, stride, tight, (planar ? 3 : 1) * out_h};
        if (role == kChild) {
            im.out = nullptr;
        } else if (o.buffer_is_host) {
            any_host_out_ = true;
            ho.host_ptr = o.buffer;
            ho.bounce_off = bounce_bytes;
            bounce_bytes += align_up(obytes, 256);
            im.out = reinterpret_cast<uint8_t*>(ho.bounce_off); // rebased once the bounce arena exists
        } else {
            im.out = static_cast<uint8_t*>(o.buffer);
        }
        // huffman set: deduplicated by content (the canonical bytes of the table specs are the key, so two streams share
        // device tables only when their specs are equal).  Cache nodes stay where they are for the whole batch: eviction
        // happens at the top of prepare() only.
        if (role == kParent || role == kProg) {
            im.huff_set = 0;
        } else {
            huff_set_key(hd, &set_key_);
            auto ci = set_cache_.find(set_key_);
            if (ci == set_cache_.end()) {
                HuffSetDev hs;
                build_huff_set(hd, &hs);
                ci = set_cache_.emplace(set_key_, hs).first;
            }
            const HuffSetDev* node = &ci->second;
            auto it = set_index_.find(node);
            if (it == set_index_.end()) {
                it = set_index_.emplace(node, int(sets.size())).first;
                sets.push_back(node);
            }
            im.huff_set = uint32_t(it->second);
        }

        if (role != kChild)
            comp_bytes_ += int64_t(len[i]);
        if (role != kParent) {
            // coefficients are counted for the MCU rows actually decoded and, for Y output
            // of a colour stream, for luma only
            int64_t blocks_per_mcu_needed = 0;
            for (int c = 0; c < im.ncomp_needed; ++c)
                blocks_per_mcu_needed += hd.h[c] * hd.v[c];
            n_coefs_ += int64_t(mcu_row_hi - mcu_row_lo) * hd.mcus_x * blocks_per_mcu_needed * 64;
        }
        if (role != kChild) {
            out_bytes_ += int64_t(rw * rh * nch);
            dev_index_[i] = int(imgs_.size());
        } else {
            ++children_ok;
        }
        if (role != kParent)
            gate[0] = gate[1] = gate[2] = uint32_t(imgs_.size());
        else if (im.ncomp_needed == 1)
            gate[1] = gate[2] = gate[0];
        im.multi = role == kParent ? 1 : 0;
        if (role == kProg) {
            // the scans just pushed are this image's: their chains follow, in file order (a predecessor has the lower index)
            const ParsedInput& pi = (parsed && parsed[i]) ? *parsed[i] : parsed_own_[size_t(i)];
            const ProgStream& pg = *pi.prog;
            im.prog = 1;
            im.pcoef_off = pcoef_elems;
            pcoef_elems += size_t(im.total_blocks) * 64;
            const uint32_t self = uint32_t(imgs_.size()), table0 = uint32_t(prog_tables.size());
            for (const HuffSpec& hs : pg.tables) {
                prog_tables.emplace_back();
                build_prog_table(hs, &prog_tables.back());
            }
            int last_scan[2][kMaxComp]; // per class (DC / AC) and component: the scan that wrote it last
            for (int cls = 0; cls < 2; ++cls)
                for (int c = 0; c < kMaxComp; ++c)
                    last_scan[cls][c] = -1;
            for (size_t sidx = 0; sidx < pg.scans.size(); ++sidx) {
                const ProgScan& ps = pg.scans[sidx];
                ProgScanDev sd;
                std::memset(&sd, 0, sizeof(sd));
                const int cls = ps.ss ? 1 : 0;
                for (int k = 0; k < ps.ncomp; ++k) {
                    const int before = last_scan[cls][ps.comp[k]];
                    bool known = before < 0;
                    for (int q = 0; q < sd.n_pred; ++q)
                        known = known || sd.pred[q] == uint32_t(before);
                    if (!known) {
                        sd.pred[sd.n_pred++] = uint32_t(before);
                        prog_scans[size_t(before)].followed = 1;
                    }
                    last_scan[cls][ps.comp[k]] = int(prog_scans.size());
                }
                if (sd.n_pred == 1) {
                    const ProgScanDev& pd = prog_scans[sd.pred[0]];
                    bool same = pd.ncomp == ps.ncomp && pd.ri == uint32_t(ps.restart_interval) && pd.n_mcu == uint32_t(ps.mcus);
                    for (int k = 0; same && k < ps.ncomp; ++k)
                        same = pd.comp[k] == ps.comp[k];
                    sd.pred_fine = same ? 1 : 0;
                }
                sd.parent = self;
                sd.child = self - uint32_t(pg.scans.size()) + uint32_t(sidx);
                sd.ri = uint32_t(ps.restart_interval);
                sd.n_mcu = uint32_t(ps.mcus);
                sd.n_chain = ps.restart_interval ? uint32_t((ps.mcus + ps.restart_interval - 1) / ps.restart_interval) : 1u;
                sd.wb = uint16_t(ps.wb);
                sd.hb = uint16_t(ps.hb);
                sd.ncomp = uint8_t(ps.ncomp);
                for (int k = 0; k < ps.ncomp; ++k) {
                    sd.comp[k] = uint8_t(ps.comp[k]);
                    sd.table[k] = ps.table[k] >= 0 ? table0 + uint32_t(ps.table[k]) : 0u;
                }
                sd.ss = uint8_t(ps.ss);
                sd.se = uint8_t(ps.se);
                sd.ah = uint8_t(ps.ah);
                sd.al = uint8_t(ps.al);
                sd.chain_off = uint32_t(prog_chain_scan.size());
                prog_chain_scan.insert(prog_chain_scan.end(), sd.n_chain, uint32_t(prog_scans.size()));
                prog_scans.push_back(sd);
            }
            prog_imgs.push_back(self);
            prog_max_blocks = std::max(prog_max_blocks, im.total_blocks);
        }
        gates_.insert(gates_.end(), gate, gate + 3);
        in_index_.push_back(i);
        item_of_dev.push_back(item);
        scan_offs.push_back(hd.scan_offset);
        imgs_.push_back(im);
        host_outs_.push_back(ho);
    }
    algo_bytes_ = comp_bytes_ + 4 * n_coefs_ + out_bytes_;
    const int nd = int(imgs_.size());

    // ---- tiles of the fused pixel kernel: one per warp.  A strip is as wide as gives the warp three full passes of blocks per
    // MCU row (96 blocks, the chroma of the two neighbouring MCU columns included where the horizontal triangle filter
    // needs it), narrower if its component tiles would not fit the warp's share of shared memory; bands are as tall as
    // still leaves every SM several waves of warps (tall bands amortise the extra chroma rows the vertical triangle
    // filter needs at a band's top and bottom).
    uint32_t n_px = 0;
    size_t px_smem = 0;
    {
        long rows_total = 0;
        for (ImageDev& im : imgs_) {
            if (!im.px_n_cta)
                continue;
            const int mode = im.color_mode;
            const bool fancy_h = mode == 6 || mode == 8, fancy_v = mode == 7 || mode == 8;
            const int ncomp = mode == 1 ? 1 : 3;
            int hh[3] = {im.h[0], im.h[1], im.h[2]}, vv[3] = {im.v[0], im.v[1], im.v[2]};
            int bpm_held = 0, halo = 0;
            for (int c = 0; c < ncomp; ++c) {
                bpm_held += hh[c] * vv[c];
                if (c > 0 && fancy_h)
                    halo += 2 * hh[c] * vv[c];
            }
            const int cols = im.px_mcu_col_hi - im.px_mcu_col_lo;
            int per_strip = std::min(cols, pixels_strip_mcus(bpm_held, halo));
            while (per_strip > 1 && pixels_tile_smem(ncomp, hh, vv, im.vmax, per_strip, fancy_h, fancy_v) > pixels_smem_budget())
                --per_strip;
            const int n_strips = (cols + per_strip - 1) / per_strip;
            im.px_n_strips = uint16_t(n_strips);
            im.px_strip_mcus = uint16_t(per_strip); // (not evened out: full strips keep the passes full, the last one takes the rest)
            rows_total += long(n_strips) * (im.px_mcu_row_hi - im.px_mcu_row_lo);
            px_smem = std::max(px_smem, pixels_tile_smem(ncomp, hh, vv, im.vmax, per_strip, fancy_h, fancy_v));
        }
        const long waves = 148L * 24 * 4; // warps an SM holds x a few rounds
        const int band = int(std::max<long>(1, std::min<long>(64, rows_total / waves)));
        for (ImageDev& im : imgs_) {
            if (!im.px_n_cta)
                continue;
            const int rows = im.px_mcu_row_hi - im.px_mcu_row_lo;
            const int n_bands = (rows + band - 1) / band;
            im.px_band_rows = uint16_t((rows + n_bands - 1) / n_bands);
            im.px_n_bands = uint16_t((rows + im.px_band_rows - 1) / im.px_band_rows);
            im.px_cta_off = n_px;
            im.px_n_cta = uint32_t(im.px_n_bands) * im.px_n_strips;
            n_px += im.px_n_cta;
        }
        if (px_smem > pixels_smem_budget())
            return fail(B200JPEG_ERR_STATE, "pixel tile exceeds its shared-memory budget");
    }

    // ---- upload arena: descriptors | cta->image maps | huffman sets | raw scan bytes
    size_t off = 0;
    const size_t off_img = off;
    off = align_up(off + sizeof(ImageDev) * nd, 256);
    const size_t off_tile = off;
    off = align_up(off + 4ull * n_tiles, 256);
    const size_t off_seq = off;
    off = align_up(off + 4ull * n_seq, 256);
    const size_t off_wseq = off;
    off = align_up(off + 4ull * n_wseq, 256);
    const size_t off_dc = off;
    off = align_up(off + 4ull * n_dc, 256);
    const size_t off_idct = off;
    off = align_up(off + 4ull * n_idct, 256);
    const size_t off_color = off;
    off = align_up(off + 4ull * n_color, 256);
    const size_t off_px = off;
    off = align_up(off + 4ull * n_px, 256);
    const size_t off_gate = off;
    off = align_up(off + 12ull * nd, 256);
    const size_t off_huff = off;
    off = align_up(off + sizeof(HuffSetDev) * sets.size(), 256);
    const size_t off_pscan = off;
    off = align_up(off + sizeof(ProgScanDev) * prog_scans.size(), 256);
    const size_t off_pchain = off;
    off = align_up(off + 4ull * prog_chain_scan.size(), 256);
    const size_t off_phuff = off;
    off = align_up(off + sizeof(ProgHuffDev) * prog_tables.size(), 256);
    const size_t off_pimg = off;
    off = align_up(off + 4ull * prog_imgs.size(), 256);
    const size_t off_raw = off;
    off += raw_bytes + 256;
    const size_t upload_bytes = off;

    // ---- scratch arena
    size_t so = 0;
    const size_t so_state = so;
    so = align_up(so + sizeof(ImageState) * nd, 256);
    const size_t so_tinfo = so;
    so = align_up(so + 8ull * n_tiles, 256);
    const size_t so_tbase = so;
    so = align_up(so + 8ull * n_tiles, 256);
    const size_t so_bnd = so;
    so = align_up(so + 4ull * n_bnd, 256);
    const size_t so_in = so;
    so = align_up(so + 4ull * n_sub, 256);
    const size_t so_out = so;
    so = align_up(so + 4ull * n_sub, 256);
    const size_t so_info = so;
    so = align_up(so + 4ull * n_sub, 256);
    const size_t so_start = so;
    so = align_up(so + 4ull * n_sub, 256);
    const size_t so_dcsum = so;
    so = align_up(so + 16ull * n_dc, 256);
    const size_t so_stats = so;
    so += 256;
    const size_t so_mask = so;
    so = align_up(so + 4ull * (coef_elems / 64), 256);
    const size_t so_pos = so;
    so = align_up(so + (bd_.sym_mode ? 4ull * (coef_elems / 64) : 0ull), 256);
    const size_t so_prog = so;
    so = align_up(so + 8ull * prog_chain_scan.size(), 256); // progress words, then elapsed clocks (B200JPEG_TRACE)

    // the coefficients live either as dense 16-byte groups (coef_) or as the entry stream (sym_), never both
    coef_bytes_ = bd_.sym_mode ? 0 : coef_elems * sizeof(int16_t);
    const size_t sym_bytes = bd_.sym_mode ? sym_chunks_total * 16 : 0;
    if (!upload_host_.ensure(upload_bytes, sv_, stream) || !state_host_.ensure(sizeof(ImageState) * std::max(nd, 1), sv_, stream))
        return fail(B200JPEG_ERR_CUDA, "pinned host allocation failed");
    if (!upload_.ensure(upload_bytes, &device_bytes_, sv_, stream) || !scratch_.ensure(so + 256, &device_bytes_, sv_, stream) ||
        !cmp_.ensure(cmp_bytes + 256, &device_bytes_, sv_, stream) || !coef_.ensure(coef_bytes_ + 256, &device_bytes_, sv_, stream) ||
        !sym_.ensure(sym_bytes + 256, &device_bytes_, sv_, stream) || !pcoef_.ensure(pcoef_elems * sizeof(int16_t) + 256, &device_bytes_, sv_, stream) ||
        !planes_.ensure(plane_bytes + 256, &device_bytes_, sv_, stream) || !bounce_.ensure(bounce_bytes + 256, &device_bytes_, sv_, stream))
        return fail(B200JPEG_ERR_CUDA, "device allocation failed");

    uint8_t* up = upload_host_.as<uint8_t>();
    uint32_t* tile_img = reinterpret_cast<uint32_t*>(up + off_tile);
    uint32_t* seq_img = reinterpret_cast<uint32_t*>(up + off_seq);
    uint32_t* wseq_img = reinterpret_cast<uint32_t*>(up + off_wseq);
    uint32_t* dc_img = reinterpret_cast<uint32_t*>(up + off_dc);
    uint32_t* idct_img = reinterpret_cast<uint32_t*>(up + off_idct);
    uint32_t* color_img = reinterpret_cast<uint32_t*>(up + off_color);
    uint32_t* px_img = reinterpret_cast<uint32_t*>(up + off_px);
    // Bitstreams that already sit in pinned / registered host memory are copied from where they lie when they are big
    // enough to be worth a copy of their own; everything else is gathered into the pinned staging block (one big copy).
    // Whether the sources are pinned is asked once per batch (the pipeline asks once per call), not per image.
    if (sources_pinned < 0) {
        cudaPointerAttributes attr{};
        sources_pinned = (nd && cudaPointerGetAttributes(&attr, data[in_index_[0]]) == cudaSuccess && attr.type == cudaMemoryTypeHost) ? 1 : 0;
        cudaGetLastError();
    }
    direct_.assign(size_t(nd), 0);
    src_scan_.assign(size_t(nd), nullptr);
    staged_bytes_ = 0;
    for (int d = 0; d < nd; ++d) {
        ImageDev& im = imgs_[d];
        if (host_outs_[d].host_ptr)
            im.out = bounce_.as<uint8_t>() + host_outs_[d].bounce_off;
        std::fill(tile_img + im.tile_off, tile_img + im.tile_off + im.n_tiles, uint32_t(d));
        std::fill(seq_img + im.seq_off, seq_img + im.seq_off + im.n_seq, uint32_t(d));
        std::fill(wseq_img + im.wseq_off, wseq_img + im.wseq_off + im.n_wseq, uint32_t(d));
        std::fill(dc_img + im.dc_cta_off, dc_img + im.dc_cta_off + im.n_dc_cta, uint32_t(d));
        std::fill(idct_img + im.idct_cta_off, idct_img + im.idct_cta_off + im.n_idct_cta, uint32_t(d));
        std::fill(color_img + im.color_cta_off, color_img + im.color_cta_off + im.n_color_cta_x * im.n_color_cta_y, uint32_t(d));
        std::fill(px_img + im.px_cta_off, px_img + im.px_cta_off + im.px_n_cta, uint32_t(d));
        src_scan_[d] = static_cast<const uint8_t*>(data[in_index_[d]]) + scan_offs[size_t(d)];
        if (sources_pinned && im.raw_len >= (uint32_t(opt_.direct_kb) << 10))
            direct_[d] = 1;
        else
            staged_bytes_ += im.raw_len;
    }
    if (nd) {
        std::memcpy(up + off_img, imgs_.data(), sizeof(ImageDev) * nd);
        std::memcpy(up + off_gate, gates_.data(), 12ull * nd);
    }
    for (size_t si = 0; si < sets.size(); ++si)
        std::memcpy(up + off_huff + si * sizeof(HuffSetDev), sets[si], sizeof(HuffSetDev));
    if (!prog_scans.empty()) {
        std::memcpy(up + off_pscan, prog_scans.data(), sizeof(ProgScanDev) * prog_scans.size());
        std::memcpy(up + off_pchain, prog_chain_scan.data(), 4ull * prog_chain_scan.size());
        std::memcpy(up + off_phuff, prog_tables.data(), sizeof(ProgHuffDev) * prog_tables.size());
        std::memcpy(up + off_pimg, prog_imgs.data(), 4ull * prog_imgs.size());
    }
    pcoef_bytes_ = pcoef_elems * sizeof(int16_t);
    off_raw_ = off_raw;
    upload_bytes_ = upload_bytes;
    const double t2 = trace_enabled() ? trace_now_us() : 0;
    if (trace_enabled()) {
        t_walk_ = t1 - t0;
        t_layout_ = t2 - t1;
        t_stage_ = t_h2d_ = 0;
    }

    uint8_t* dv = upload_.as<uint8_t>();
    uint8_t* sc = scratch_.as<uint8_t>();
    bd_.n_images = nd;
    bd_.img = reinterpret_cast<const ImageDev*>(dv + off_img);
    bd_.tile_img = reinterpret_cast<const uint32_t*>(dv + off_tile);
    bd_.seq_img = reinterpret_cast<const uint32_t*>(dv + off_seq);
    bd_.wseq_img = reinterpret_cast<const uint32_t*>(dv + off_wseq);
    bd_.dc_cta_img = reinterpret_cast<const uint32_t*>(dv + off_dc);
    bd_.idct_cta_img = reinterpret_cast<const uint32_t*>(dv + off_idct);
    bd_.color_cta_img = reinterpret_cast<const uint32_t*>(dv + off_color);
    bd_.px_cta_img = reinterpret_cast<const uint32_t*>(dv + off_px);
    bd_.n_px_task = int(n_px);
    bd_.px_smem = uint32_t(px_smem);
    bd_.gate = reinterpret_cast<const uint32_t*>(dv + off_gate);
    bd_.n_multi = 0;
    for (const ImageDev& im : imgs_)
        bd_.n_multi += im.multi;
    bd_.huff = reinterpret_cast<const HuffSetDev*>(dv + off_huff);
    bd_.prog_scan = reinterpret_cast<const ProgScanDev*>(dv + off_pscan);
    bd_.prog_chain = reinterpret_cast<const uint32_t*>(dv + off_pchain);
    bd_.prog_huff = reinterpret_cast<const ProgHuffDev*>(dv + off_phuff);
    bd_.prog_img = reinterpret_cast<const uint32_t*>(dv + off_pimg);
    bd_.pcoef = pcoef_.as<int16_t>();
    bd_.n_prog_img = int(prog_imgs.size());
    bd_.n_prog_chain = int(prog_chain_scan.size());
    bd_.prog_progress = reinterpret_cast<uint32_t*>(scratch_.as<uint8_t>() + so_prog);
    bd_.prog_max_blocks = prog_max_blocks;
    bd_.raw = dv + off_raw;
    bd_.state = reinterpret_cast<ImageState*>(sc + so_state);
    bd_.tile_info = reinterpret_cast<uint2*>(sc + so_tinfo);
    bd_.tile_base = reinterpret_cast<uint2*>(sc + so_tbase);
    bd_.bnd = reinterpret_cast<uint32_t*>(sc + so_bnd);
    bd_.sub_in = reinterpret_cast<uint32_t*>(sc + so_in);
    bd_.sub_out = reinterpret_cast<uint32_t*>(sc + so_out);
    bd_.sub_info = reinterpret_cast<uint32_t*>(sc + so_info);
    bd_.sub_start = reinterpret_cast<uint32_t*>(sc + so_start);
    bd_.sync_stats = reinterpret_cast<uint32_t*>(sc + so_stats);
    bd_.blkrec = reinterpret_cast<uint32_t*>(sc + so_mask);
    bd_.cmp = cmp_.as<uint32_t>();
    bd_.coef = coef_.as<int16_t>();
    bd_.sym = sym_.as<uint4>();
    bd_.blkpos = reinterpret_cast<uint32_t*>(sc + so_pos);
    bd_.planes = planes_.as<uint8_t>();
    bd_.n_tiles = int(n_tiles);
    bd_.n_seq = int(n_seq);
    bd_.n_wseq = int(n_wseq);
    bd_.n_sub = int(n_sub);
    bd_.n_dc_cta = int(n_dc);
    bd_.dc_sum = reinterpret_cast<int4*>(sc + so_dcsum);
    bd_.n_idct_cta = int(n_idct);
    bd_.n_color_cta = int(n_color);
    states_.assign(nd, ImageState{0, 0, 0, 0});
    if (profiling_ && !ev_valid_) {
        for (auto& e : ev_)
            cudaEventCreate(&e);
        ev_valid_ = true;
    }
    prepared_ = true;
    staged_ = false;
    if (defer_upload)
        return B200JPEG_OK;
    stage_run();
    return upload(stream);
}

// Copy into the pinned staging block with non-temporal stores: the block is only ever read by the copy engine, so pulling its
// lines into the caches first (what a plain store does) is wasted memory traffic.  dst is 16-byte aligned.
static void stream_copy(uint8_t* dst, const uint8_t* src, size_t n)
{
#if defined(__SSE2__)
    size_t i = 0;
    for (; i + 64 <= n; i += 64) {
        const __m128i a = _mm_loadu_si128(reinterpret_cast<const __m128i*>(src + i));
        const __m128i b = _mm_loadu_si128(reinterpret_cast<const __m128i*>(src + i + 16));
        const __m128i c = _mm_loadu_si128(reinterpret_cast<const __m128i*>(src + i + 32));
        const __m128i d = _mm_loadu_si128(reinterpret_cast<const __m128i*>(src + i + 48));
        _mm_stream_si128(reinterpret_cast<__m128i*>(dst + i), a);
        _mm_stream_si128(reinterpret_cast<__m128i*>(dst + i + 16), b);
        _mm_stream_si128(reinterpret_cast<__m128i*>(dst + i + 32), c);
        _mm_stream_si128(reinterpret_cast<__m128i*>(dst + i + 48), d);
    }
    if (i < n)
        std::memcpy(dst + i, src + i, n - i);
#else
    std::memcpy(dst, src, n);
#endif
}

// The staging copies of a prepared batch (bitstreams that are not copied straight from the caller's buffers) into the pinned
// block, cut into jobs of about 1 MB for the helper threads.  May run on any host thread.
void Decoder::stage_run()
{
    const double t0 = trace_enabled() ? trace_now_us() : 0;
    const int nd = int(imgs_.size());
    if (staged_bytes_) {
        nvtx3::scoped_range nvtx_stage{"b200jpeg stage bitstreams"};
        struct Job {
            int d0, d1;
        };
        std::vector<Job> jobs;
        {
            size_t acc = 0;
            int d0 = 0;
            for (int d = 0; d < nd; ++d) {
                acc += direct_[d] ? 0 : imgs_[d].raw_len;
                if (acc >= (size_t(1) << 20) || d + 1 == nd) {
                    jobs.push_back({d0, d + 1});
                    d0 = d + 1;
                    acc = 0;
                }
            }
        }
        struct Ctx {
            const Job* jobs;
            const uint8_t* direct;
            const ImageDev* imgs;
            const uint8_t* const* src;
            uint8_t* dst;
        } ctx{jobs.data(), direct_.data(), imgs_.data(), src_scan_.data(), upload_host_.as<uint8_t>() + off_raw_};
        auto stage = [](int j, void* a) {
            const Ctx& c = *static_cast<const Ctx*>(a);
            for (int d = c.jobs[j].d0; d < c.jobs[j].d1; ++d)
                if (!c.direct[d])
                    stream_copy(c.dst + c.imgs[d].raw_off, c.src[d], c.imgs[d].raw_len);
#if defined(__SSE2__)
            _mm_sfence();
#endif
        };
        if (staged_bytes_ >= (size_t(2) << 20))
            parallel_for(int(jobs.size()), stage, &ctx);
        else
            for (int j = 0; j < int(jobs.size()); ++j)
                stage(j, &ctx);
    }
    staged_ = true;
    if (trace_enabled())
        t_stage_ = trace_now_us() - t0;
}

// Host-to-device copies of a prepared and staged batch: one for descriptors + maps + tables (+ the staged bitstreams), then
// one per bitstream that is copied straight from the caller's pinned buffer.
int Decoder::upload(cudaStream_t stream)
{
    if (!prepared_ || !staged_)
        return fail(B200JPEG_ERR_STATE, "upload without a prepared and staged batch");
    const double t0 = trace_enabled() ? trace_now_us() : 0;
    const int nd = int(imgs_.size());
    const uint8_t* up = upload_host_.as<uint8_t>();
    const size_t first_copy = staged_bytes_ ? upload_bytes_ : off_raw_;
    last_stream_ = stream;
    inflight_ = nd > 0;
    if (nd && !cuda_ok(cudaMemcpyAsync(upload_.ptr, up, first_copy, cudaMemcpyHostToDevice, stream), "upload"))
        return B200JPEG_ERR_CUDA;
    for (int d = 0; d < nd; ++d) {
        if (!direct_[d])
            continue;
        const ImageDev& im = imgs_[d];
        if (!cuda_ok(cudaMemcpyAsync(upload_.as<uint8_t>() + off_raw_ + im.raw_off, src_scan_[d], im.raw_len, cudaMemcpyHostToDevice, stream),
                "bitstream upload"))
            return B200JPEG_ERR_CUDA;
    }
    if (trace_enabled())
        t_h2d_ = trace_now_us() - t0;
    return B200JPEG_OK;
}

int Decoder::execute(cudaStream_t stream)
{
    if (!prepared_)
        return fail(B200JPEG_ERR_STATE, "execute without a prepared batch");
    if (!cuda_ok(cudaSetDevice(device_), "cudaSetDevice"))
        return B200JPEG_ERR_CUDA;
    nvtx3::scoped_range nvtx_execute{"b200jpeg execute (K0-K4 launches)"};
    launches_ = 0;
    executed_ = true;
    ev_recorded_ = false;
    if (bd_.n_images == 0)
        return B200JPEG_OK;
    cudaMemsetAsync(bd_.sync_stats, 0, 16, stream);
    mark(0, stream);
    launches_ += launch_scan(bd_, stream);
    mark(1, stream);
    launches_ += launch_sync(bd_, false, opt_.debug_no_warm, stream);
    mark(2, stream);
    launches_ += launch_verify_scan(bd_, false, stream);
    mark(3, stream);
    launches_ += launch_write(bd_, false, stream); // writes every coefficient exactly once: no clearing pass
    mark(4, stream);
    launches_ += launch_dc_scan(bd_, false, stream);
    if (bd_.n_prog_img) { // progressive images: their scans accumulate into zeroed blocks
        cudaMemsetAsync(bd_.pcoef, 0, pcoef_bytes_, stream);
        launches_ += launch_progressive(bd_, stream);
        if (trace_enabled()) { // how long every chain took (thousands of SM clocks), with its scan's parameters
            std::vector<uint32_t> clk(size_t(bd_.n_prog_chain)), chain_scan(size_t(bd_.n_prog_chain));
            std::vector<ProgScanDev> scans(chain_scan.size());
            cudaStreamSynchronize(stream);
            cudaMemcpy(clk.data(), bd_.prog_progress + bd_.n_prog_chain, 4 * clk.size(), cudaMemcpyDeviceToHost);
            cudaMemcpy(chain_scan.data(), bd_.prog_chain, 4 * clk.size(), cudaMemcpyDeviceToHost);
            for (size_t c = 0; c < clk.size() && c < 24; ++c) {
                ProgScanDev sd;
                cudaMemcpy(&sd, bd_.prog_scan + chain_scan[c], sizeof(sd), cudaMemcpyDeviceToHost);
                std::fprintf(stderr, "[b200jpeg] chain %zu: scan %u ncomp %u Ss %u Se %u Ah %u Al %u: %u kclk\n", c, chain_scan[c], sd.ncomp, sd.ss,
                    sd.se, sd.ah, sd.al, clk[c]);
            }
        }
    }
    // Everything a status depends on is known once the entropy stage is through (the pixel kernels only read the
    // flags).  The read-back goes over a side stream so that it neither waits for the pixel kernels nor holds them up;
    // finish_wait() then only has to wait for it, and the caller gets its statuses while the pixels are still being made.
    cudaEventRecord(entropy_event_, stream);
    cudaStreamWaitEvent(side_stream_, entropy_event_, 0);
    cudaMemcpyAsync(state_host_.ptr, bd_.state, sizeof(ImageState) * bd_.n_images, cudaMemcpyDeviceToHost, side_stream_);
    cudaEventRecord(entropy_event_, side_stream_);
    mark(5, stream);
    launches_ += run_pixel_stage(false, stream);
    mark(8, stream);
    ev_recorded_ = profiling_ && ev_valid_;
    last_stream_ = stream;
    if (!cuda_ok(cudaGetLastError(), "kernel launch"))
        return B200JPEG_ERR_CUDA;
    return B200JPEG_OK;
}

int Decoder::run_pixel_stage(bool redo_only, cudaStream_t stream)
{
    int n = launch_idct(bd_, redo_only, stream);
    if (!redo_only)
        mark(6, stream);
    n += launch_color(bd_, redo_only, stream);
    if (!redo_only)
        mark(7, stream);
    n += launch_pixels(bd_, redo_only, stream);
    return n;
}

// Finish, part 1 (asynchronous): queue the copies of host outputs, mark the end of the batch.
int Decoder::finish_async(cudaStream_t stream)
{
    if (!executed_)
        return fail(B200JPEG_ERR_STATE, "finish without an executed batch");
    if (!cuda_ok(cudaSetDevice(device_), "cudaSetDevice"))
        return B200JPEG_ERR_CUDA;
    const int nd = bd_.n_images;
    for (int d = 0; d < nd; ++d)
        if (!copy_host_output(d, stream))
            return B200JPEG_ERR_CUDA;
    cudaEventRecord(done_event_, stream);
    have_done_ = true;
    last_stream_ = stream;
    finish_queued_ = true;
    return B200JPEG_OK;
}

// Pixels of a host output: only the pixel bytes of every row travel (a caller's row padding is left alone, and a
// buffer that ends with the last pixel of the last row is not overrun).
bool Decoder::copy_host_output(int d, cudaStream_t stream)
{
    const HostOut& ho = host_outs_[d];
    if (!ho.host_ptr)
        return true;
    const uint8_t* src = bounce_.as<uint8_t>() + ho.bounce_off;
    if (ho.stride == ho.tight)
        return cuda_ok(cudaMemcpyAsync(ho.host_ptr, src, ho.tight * ho.rows, cudaMemcpyDeviceToHost, stream), "output copy");
    return cuda_ok(cudaMemcpy2DAsync(ho.host_ptr, ho.stride, src, ho.stride, ho.tight, ho.rows, cudaMemcpyDeviceToHost, stream), "output copy");
}

// Finish, part 2: wait for the statuses (and for the pixels when some of them go to host memory), run the rare
// re-synchronisation path, report.  With device outputs only, the pixel kernels may still be running on `stream` when
// this returns: work queued on that stream afterwards sees the finished pixels, as with any stream-ordered CUDA call.
int Decoder::finish_wait(int* status, cudaStream_t stream)
{
    if (!finish_queued_)
        return fail(B200JPEG_ERR_STATE, "finish_wait without finish_async");
    finish_queued_ = false;
    nvtx3::scoped_range nvtx_finish{"b200jpeg finish (wait for statuses)"};
    if (!cuda_ok(cudaSetDevice(device_), "cudaSetDevice"))
        return B200JPEG_ERR_CUDA;
    const int nd = bd_.n_images;
    resync_images_ = 0;
    if (nd) {
        if (!cuda_ok(cudaEventSynchronize(entropy_event_), "synchronize"))
            return B200JPEG_ERR_CUDA;
        std::memcpy(states_.data(), state_host_.ptr, sizeof(ImageState) * nd);
    }
    inflight_ = false; // the uploads lie before the entropy stage
    if ((any_host_out_ || opt_.sync_finish) && !cuda_ok(cudaStreamSynchronize(stream), "synchronize"))
        return B200JPEG_ERR_CUDA;
    auto read_states = [&]() -> bool {
        if (!cuda_ok(cudaMemcpyAsync(state_host_.ptr, bd_.state, sizeof(ImageState) * nd, cudaMemcpyDeviceToHost, stream), "state read-back"))
            return false;
        if (!cuda_ok(cudaStreamSynchronize(stream), "synchronize"))
            return false;
        std::memcpy(states_.data(), state_host_.ptr, sizeof(ImageState) * nd);
        return true;
    };
    // flags that decide about image d's pixels: its own, or those of the one-component images its planes came from
    auto flags_of = [&](int d) { return states_[gates_[3 * d]].flags | states_[gates_[3 * d + 1]].flags | states_[gates_[3 * d + 2]].flags; };
    auto count_dirty = [&]() {
        int c = 0;
        for (int d = 0; d < nd; ++d) // (an image assembled from one-component images only mirrors their flags)
            c += (!imgs_[d].multi && (states_[d].flags & kFlagDirty)) ? 1 : 0;
        return c;
    };
    int dirty = count_dirty();
    if (dirty) {
        // Rare path: some sequence of some image did not enter with the state its predecessor left.
        // Propagate corrected entry states one sequence per round until K2 finds every image consistent,
        // then produce coefficients and pixels of those images again.
        nvtx3::scoped_range nvtx_fix{"b200jpeg re-synchronisation path"};
        resync_images_ = dirty;
        int rounds = 0;
        while (dirty && rounds < opt_.max_fix_rounds) {
            launches_ += launch_sync(bd_, true, false, stream);
            launches_ += launch_verify_scan(bd_, true, stream);
            if (!read_states())
                return B200JPEG_ERR_CUDA;
            dirty = count_dirty();
            ++rounds;
        }
        launches_ += launch_write(bd_, true, stream);
        launches_ += launch_dc_scan(bd_, true, stream);
        launches_ += run_pixel_stage(true, stream);
        if (!read_states())
            return B200JPEG_ERR_CUDA;
        bool any_host = false;
        for (int d = 0; d < nd; ++d) {
            if (!host_outs_[d].host_ptr || !(flags_of(d) & kFlagRedo))
                continue;
            any_host = true;
            if (!copy_host_output(d, stream))
                return B200JPEG_ERR_CUDA;
        }
        if (any_host && !cuda_ok(cudaStreamSynchronize(stream), "synchronize"))
            return B200JPEG_ERR_CUDA;
        cudaEventRecord(done_event_, stream);
    }
    if (status) {
        for (int i = 0; i < n_in_; ++i) {
            int st = host_status_[i];
            const int d = dev_index_[i];
            if (st == B200JPEG_OK && d >= 0) {
                const uint32_t f = flags_of(d);
                if (f & kFlagDirty)
                    st = B200JPEG_ERR_UNSUPPORTED;
                else if (f & (kFlagBadMarker | kFlagBadHuffman | kFlagOverrun | kFlagPhase))
                    st = B200JPEG_ERR_CORRUPT;
            }
            status[i] = st;
        }
    }
    return B200JPEG_OK;
}

int Decoder::finish(int* status, cudaStream_t stream)
{
    const int rc = finish_async(stream);
    return rc == B200JPEG_OK ? finish_wait(status, stream) : rc;
}

int64_t Decoder::stat(int which) const
{
    switch (which) {
    case B200JPEG_STAT_KERNEL_LAUNCHES: return launches_;
    case B200JPEG_STAT_ALGO_BYTES: return algo_bytes_;
    case B200JPEG_STAT_COMPRESSED_BYTES: return comp_bytes_;
    case B200JPEG_STAT_NUM_COEFS: return n_coefs_;
    case B200JPEG_STAT_OUTPUT_BYTES: return out_bytes_;
    case B200JPEG_STAT_RESYNC_IMAGES: return resync_images_;
    case B200JPEG_STAT_DEVICE_BYTES: return int64_t(device_bytes_);
    case B200JPEG_STAT_NUM_SUBSEQ: return bd_.n_sub;
    case B200JPEG_STAT_SYNC_CTAS:
    case B200JPEG_STAT_SYNC_REPAIR_ROUNDS:
    case B200JPEG_STAT_SYNC_REPAIRED:
    case B200JPEG_STAT_SYNC_MAX_ROUNDS: {
        uint32_t v[4] = {0, 0, 0, 0};
        if (!bd_.sync_stats || cudaMemcpy(v, bd_.sync_stats, sizeof(v), cudaMemcpyDeviceToHost) != cudaSuccess)
            return -1;
        return v[which - B200JPEG_STAT_SYNC_CTAS];
    }
    default: return -1;
    }
}

int Decoder::stage_times(float* ms, const char** names, int cap)
{
    if (!ev_recorded_)
        return 0;
    cudaEventSynchronize(ev_[kStages]);
    int n = 0;
    for (int i = 0; i < kStages && n < cap; ++i, ++n) {
        float t = 0.f;
        cudaEventElapsedTime(&t, ev_[i], ev_[i + 1]);
        ms[n] = t;
        if (names)
            names[n] = kStageNames[i];
    }
    return n;
}

int Decoder::copy_coefficients(int index, int16_t* host_out, size_t cap_elems)
{
    if (!executed_ || index < 0 || index >= n_in_ || dev_index_[index] < 0)
        return fail(B200JPEG_ERR_ARG, "no such decoded image");
    const ImageDev& im = imgs_[dev_index_[index]];
    const size_t elems = size_t(im.total_blocks) * 64;
    if (cap_elems < elems)
        return fail(B200JPEG_ERR_ARG, "buffer too small");
    cudaSetDevice(device_);
    if (im.prog) { // dense blocks of their own; the record pass moved the DC value into the block record
        std::vector<uint32_t> recs(im.total_blocks);
        if (!cuda_ok(cudaMemcpy(host_out, bd_.pcoef + im.pcoef_off, elems * sizeof(int16_t), cudaMemcpyDeviceToHost), "coefficient copy") ||
            !cuda_ok(cudaMemcpy(recs.data(), bd_.blkrec + im.coef_off / 64, recs.size() * sizeof(uint32_t), cudaMemcpyDeviceToHost), "record copy"))
            return B200JPEG_ERR_CUDA;
        for (size_t b = 0; b < recs.size(); ++b)
            host_out[b * 64] = int16_t(recs[b] >> 16);
        return B200JPEG_OK;
    }
    if (bd_.sym_mode) {
        // entry-stream form: the same walk the IDCT stage does (sym_fetch_block in pixels.cuh), on the host
        const int d = dev_index_[index];
        cudaDeviceSynchronize();
        const size_t chunks = size_t((im.n_sub_max + 31) / 32) * bd_.sym_chunks * 32;
        std::vector<uint32_t> sym(chunks * 4);
        std::vector<uint32_t> recs(im.total_blocks), pos(im.total_blocks);
        ImageState st{};
        if (!cuda_ok(cudaMemcpy(sym.data(), bd_.sym + im.sym_off, chunks * 16, cudaMemcpyDeviceToHost), "entry copy") ||
            !cuda_ok(cudaMemcpy(recs.data(), bd_.blkrec + im.coef_off / 64, recs.size() * 4, cudaMemcpyDeviceToHost), "record copy") ||
            !cuda_ok(cudaMemcpy(pos.data(), bd_.blkpos + im.coef_off / 64, pos.size() * 4, cudaMemcpyDeviceToHost), "record copy") ||
            !cuda_ok(cudaMemcpy(&st, bd_.state + d, sizeof(st), cudaMemcpyDeviceToHost), "state copy"))
            return B200JPEG_ERR_CUDA;
        std::memset(host_out, 0, elems * sizeof(int16_t));
        for (size_t b = 0; b < recs.size(); ++b) {
            int16_t* blk = host_out + b * 64;
            uint32_t k = pos[b] >> bd_.sub_shift, slot = (recs[b] & 0xFFFFu) + 1u;
            bool done = k >= st.n_sub;
            for (int guard = 0; !done && guard < 8192; ++guard) {
                if (slot >= bd_.sym_chunks * 4)
                    break;
                const size_t chunk = (size_t(k >> 5) * bd_.sym_chunks + (slot >> 2)) * 32 + (k & 31u);
                const uint32_t e = sym[chunk * 4 + (slot & 3u)];
                ++slot;
                const uint32_t off = (e >> 16) & 0x3E02u; // byte offset in the scatter column -> zig-zag position
                blk[((off / kScatterStride) << 1) | ((off >> 1) & 1u)] = int16_t(e & 0xFFFFu);
                if (e & 0x40000000u) {
                    done = true;
                } else if (e & 0x80000000u) {
                    ++k;
                    slot = 0;
                    done = k >= st.n_sub;
                }
            }
            blk[0] = int16_t(recs[b] >> 16);
        }
        return B200JPEG_OK;
    }
    if (!cuda_ok(cudaMemcpy(host_out, bd_.coef + im.coef_off, elems * sizeof(int16_t), cudaMemcpyDeviceToHost), "coefficient copy"))
        return B200JPEG_ERR_CUDA;
    // groups K3 skipped (all zero) hold stale bytes on the device: blank them using the per-block records,
    // which also carry the DC value (K3 never stores it into the coefficient array)
    std::vector<uint32_t> recs(im.total_blocks);
    if (!cuda_ok(cudaMemcpy(recs.data(), bd_.blkrec + im.coef_off / 64, recs.size() * sizeof(uint32_t), cudaMemcpyDeviceToHost), "record copy"))
        return B200JPEG_ERR_CUDA;
    for (size_t b = 0; b < recs.size(); ++b) {
        for (int g = 0; g < 8; ++g)
            if (!((recs[b] >> g) & 1))
                std::memset(host_out + b * 64 + g * 8, 0, 16);
        host_out[b * 64] = int16_t(recs[b] >> 16);
    }
    return B200JPEG_OK;
}

// Kernel-level K0 entry (b200jpegScanSegment): one segment, descriptors and per-tile counts in
// decoder-owned scratch, everything else supplied by the caller.
int Decoder::scan_segment(const void* dev_raw, size_t length, uint32_t expected_rst, void* dev_words, uint32_t* dev_bnd, uint32_t bnd_cap,
    void* dev_result, cudaStream_t stream)
{
    if (!dev_raw || !dev_words || !dev_bnd || !dev_result || length == 0 || (reinterpret_cast<uintptr_t>(dev_raw) & 15u) ||
        (reinterpret_cast<uintptr_t>(dev_words) & 3u) || bnd_cap < expected_rst + 1)
        return fail(B200JPEG_ERR_ARG, "bad argument");
    if (length >= (size_t(1) << 28))
        return fail(B200JPEG_ERR_TOO_LARGE, "segment too large");
    if (!cuda_ok(cudaSetDevice(device_), "cudaSetDevice"))
        return B200JPEG_ERR_CUDA;
    const uint32_t n_tiles = uint32_t((length + kScanTileBytes - 1) / kScanTileBytes);
    const size_t off_tile = align_up(sizeof(ImageDev), 256);
    const size_t off_info = align_up(off_tile + 4ull * n_tiles, 256);
    const size_t off_base = align_up(off_info + 8ull * n_tiles, 256);
    const size_t total = off_base + 8ull * n_tiles;
    if (!scan_scratch_.ensure(total, &device_bytes_, sv_, stream) || !scan_host_.ensure(off_info, sv_, stream))
        return fail(B200JPEG_ERR_CUDA, "allocation failed");
    // the previous call's upload may still be in flight out of the pinned staging block
    if (!cuda_ok(cudaStreamSynchronize(stream), "synchronize"))
        return B200JPEG_ERR_CUDA;
    uint8_t* hp = scan_host_.as<uint8_t>();
    ImageDev im;
    std::memset(&im, 0, sizeof(im));
    im.raw_len = uint32_t(length);
    im.n_tiles = n_tiles;
    im.n_rst_expected = expected_rst;
    im.n_bnd_cap = bnd_cap;
    im.n_sub_max = 0xFFFFFFFFu; // (no subsequence arrays here: the count is only reported)
    std::memcpy(hp, &im, sizeof(im));
    std::memset(hp + off_tile, 0, 4ull * n_tiles);
    uint8_t* dp = scan_scratch_.as<uint8_t>();
    if (!cuda_ok(cudaMemcpyAsync(dp, hp, off_tile + 4ull * n_tiles, cudaMemcpyHostToDevice, stream), "upload"))
        return B200JPEG_ERR_CUDA;
    BatchDev b{};
    b.n_images = 1;
    b.img = reinterpret_cast<const ImageDev*>(dp);
    b.tile_img = reinterpret_cast<const uint32_t*>(dp + off_tile);
    b.tile_info = reinterpret_cast<uint2*>(dp + off_info);
    b.tile_base = reinterpret_cast<uint2*>(dp + off_base);
    b.state = static_cast<ImageState*>(dev_result);
    b.raw = static_cast<const uint8_t*>(dev_raw);
    b.cmp = static_cast<uint32_t*>(dev_words);
    b.bnd = dev_bnd;
    b.n_tiles = int(n_tiles);
    b.sub_bits = 1024; // only sizes the n_sub field of the result (documented as 1024-bit units)
    launch_scan(b, stream);
    if (!cuda_ok(cudaGetLastError(), "kernel launch"))
        return B200JPEG_ERR_CUDA;
    return B200JPEG_OK;
}

// ------------------------------------------------------------------------------------------------
// Pipeline: the batch is cut into chunks that rotate over a few lanes (decoder + stream each), so
// that the host marker walk / staging of chunk i+1, the H2D of its bitstreams, the kernels of chunk
// i and the D2H of chunk i-1 overlap.  Chunk results are identical to an unchunked decode: images
// are independent.
// ------------------------------------------------------------------------------------------------
Pipeline::Pipeline(int device_id, const Options& opt, int n_lanes, const Services& sv)
    : device_(device_id)
    , opt_(opt)
    , sv_(sv)
{
    if (!sv_.parallel_for) {
        int n = opt_.host_threads;
        if (n <= 0) {
            cpu_set_t set;
            CPU_ZERO(&set);
            const int cpus = sched_getaffinity(0, sizeof(set), &set) == 0 ? CPU_COUNT(&set) : int(std::thread::hardware_concurrency());
            n = std::min(12, std::max(1, cpus - 4));
        }
        if (n > 1)
            pool_ = std::make_unique<HostPool>(n - 1); // the calling thread takes part
    }
    lanes_.resize(size_t(std::max(1, n_lanes)));
    for (auto& l : lanes_)
        l.dec = std::make_unique<Decoder>(device_id, opt, sv_, pool_.get());
}

void Pipeline::stager_main()
{
    for (;;) {
        Lane* l;
        {
            std::unique_lock<std::mutex> lk(st_mu_);
            st_cv_.wait(lk, [&] { return st_stop_ || !st_queue_.empty(); });
            if (st_queue_.empty())
                return;
            l = st_queue_.front();
            st_queue_.pop_front();
        }
        l->dec->stage_run();
        {
            std::lock_guard<std::mutex> lk(st_mu_);
            l->staged = true;
        }
        st_done_cv_.notify_all();
    }
}

void Pipeline::submit_stage(Lane& l)
{
    if (!stager_.joinable())
        stager_ = std::thread([this] { stager_main(); });
    {
        std::lock_guard<std::mutex> lk(st_mu_);
        l.staged = false;
        st_queue_.push_back(&l);
    }
    st_cv_.notify_one();
}

void Pipeline::wait_staged(Lane& l)
{
    std::unique_lock<std::mutex> lk(st_mu_);
    st_done_cv_.wait(lk, [&] { return l.staged; });
}

Pipeline::~Pipeline()
{
    if (stager_.joinable()) {
        {
            std::lock_guard<std::mutex> lk(st_mu_);
            st_stop_ = true;
        }
        st_cv_.notify_all();
        stager_.join();
    }
    cudaSetDevice(device_);
    for (auto& l : lanes_)
        l.dec.reset(); // while the lane streams are still alive
    for (auto& l : lanes_) {
        if (l.stream)
            cudaStreamDestroy(l.stream);
    }
    if (user_event_)
        cudaEventDestroy(user_event_);
}

int Pipeline::decode(int n, const void* const* data, const size_t* len, const b200jpegOutput_t* outs, int* status, cudaStream_t user_stream,
    const ParsedInput* const* parsed)
{
    if (n < 0 || (n > 0 && (!data || !len || !outs)))
        return B200JPEG_ERR_ARG;
    if (cudaSetDevice(device_) != cudaSuccess)
        return B200JPEG_ERR_CUDA;
    last_launches_ = 0;
    last_resync_ = 0;
    // are the bitstreams in pinned memory?  asked once per call
    int pinned = 0;
    for (int i = 0; i < n && !pinned; ++i) {
        if (!data[i])
            continue;
        cudaPointerAttributes attr{};
        pinned = (cudaPointerGetAttributes(&attr, data[i]) == cudaSuccess && attr.type == cudaMemoryTypeHost) ? 1 : 0;
        cudaGetLastError();
        break;
    }
    // chunk boundaries: chunk_mb of bitstream or chunk_images images, whichever comes first; small batches stay whole
    std::vector<int> cuts{0};
    {
        size_t acc = 0;
        int cnt = 0;
        size_t total = 0;
        for (int i = 0; i < n; ++i)
            total += len[i];
        size_t chunk_bytes = size_t(opt_.chunk_mb) << 20;
        if (!chunk_bytes) {
            // Small chunks start the GPU early.  Pinned bitstreams are copied straight from the caller's buffers; pageable ones
            // are gathered into a pinned staging block by the background stager while the previous chunk is launched.
            chunk_bytes = size_t(8) << 20;
        }
        const int chunk_images = opt_.chunk_images;
        // Progressive streams decode one warp per scan, each scan a serial chain: such a batch takes about as long as its
        // longest scan however many images it holds, so cutting it up would only queue those latencies one behind the other.
        bool any_progressive = false;
        if (opt_.progressive)
            for (int i = 0; i < n && !any_progressive; ++i) {
                if (parsed && parsed[i]) {
                    any_progressive = parsed[i]->status == kStreamProgressive;
                } else if (data[i]) {
                    any_progressive = sof_marker_of(static_cast<const uint8_t*>(data[i]), len[i]) == 0xC2;
                }
            }
        const bool split = !any_progressive && lanes_.size() > 1 && (total > chunk_bytes + chunk_bytes / 3 || n > 2 * chunk_images);
        for (int i = 0; i < n; ++i) {
            acc += len[i];
            ++cnt;
            if (split && (acc >= chunk_bytes || cnt >= chunk_images) && i + 1 < n) {
                cuts.push_back(i + 1);
                acc = 0;
                cnt = 0;
            }
        }
        cuts.push_back(n);
    }
    const int n_chunks = int(cuts.size()) - 1;
    if (n_chunks > 1) {
        // lanes run on their own streams, ordered after whatever the caller already queued
        if (!user_event_ && cudaEventCreateWithFlags(&user_event_, cudaEventDisableTiming) != cudaSuccess)
            return B200JPEG_ERR_CUDA;
        cudaEventRecord(user_event_, user_stream);
    }
    int rc_all = B200JPEG_OK;
    const double tr0 = trace_enabled() ? trace_now_us() : 0;
    double tr_walk = 0, tr_layout = 0, tr_stage = 0, tr_h2d = 0, tr_exec = 0, tr_wait = 0;
    auto drain = [&](Lane& l) {
        if (!l.busy)
            return;
        const double w0 = trace_enabled() ? trace_now_us() : 0;
        const int rc = l.dec->finish_wait(status ? status + l.first : nullptr, l.active_stream);
        if (trace_enabled())
            tr_wait += trace_now_us() - w0;
        if (rc != B200JPEG_OK && rc_all == B200JPEG_OK) {
            rc_all = rc;
            err_ = l.dec->last_error();
        }
        last_launches_ += l.dec->stat(B200JPEG_STAT_KERNEL_LAUNCHES);
        last_resync_ += l.dec->stat(B200JPEG_STAT_RESYNC_IMAGES);
        l.busy = false;
    };
    for (auto& l : lanes_)
        l.used = false;
    // Pageable bitstreams of a split batch are staged in the background: chunk c is being gathered into pinned memory while
    // this thread lays out chunk c + 1 and uploads / launches chunk c - 1.
    const bool background = n_chunks > 1 && !pinned;
    auto launch = [&](Lane& l, int first, int cnt, int rc) {
        cudaStream_t s = l.active_stream;
        const double e0 = trace_enabled() ? trace_now_us() : 0;
        if (rc == B200JPEG_OK && background) {
            wait_staged(l);
            rc = l.dec->upload(s);
        }
        if (rc == B200JPEG_OK)
            rc = l.dec->execute(s);
        if (rc == B200JPEG_OK)
            rc = l.dec->finish_async(s);
        if (trace_enabled()) {
            double a, b, c2, d;
            l.dec->trace_times(&a, &b, &c2, &d);
            tr_walk += a;
            tr_layout += b;
            tr_stage += c2;
            tr_h2d += d;
            tr_exec += trace_now_us() - e0;
        }
        if (rc != B200JPEG_OK) {
            if (rc_all == B200JPEG_OK) {
                rc_all = rc;
                err_ = l.dec->last_error();
            }
            if (status)
                for (int i = 0; i < cnt; ++i)
                    status[first + i] = rc;
            return;
        }
        l.first = first;
        l.busy = true;
        l.used = true;
    };
    Lane* pending = nullptr; // laid out and handed to the stager, not launched yet
    int pending_first = 0, pending_cnt = 0, pending_rc = B200JPEG_OK;
    for (int c = 0; c < n_chunks; ++c) {
        Lane& l = lanes_[size_t(c) % lanes_.size()];
        if (pending == &l) { // (a single lane: nothing to overlap with)
            launch(*pending, pending_first, pending_cnt, pending_rc);
            pending = nullptr;
        }
        drain(l);
        const int first = cuts[c], cnt = cuts[c + 1] - cuts[c];
        cudaStream_t s = user_stream;
        if (n_chunks > 1) {
            if (!l.stream && cudaStreamCreateWithFlags(&l.stream, cudaStreamNonBlocking) != cudaSuccess)
                return B200JPEG_ERR_CUDA;
            s = l.stream;
            cudaStreamWaitEvent(s, user_event_, 0);
        }
        l.active_stream = s;
        const int rc = l.dec->prepare(cnt, data + first, len + first, outs + first, s, parsed ? parsed + first : nullptr, pinned, background);
        if (background) {
            if (rc == B200JPEG_OK)
                submit_stage(l);
            if (pending)
                launch(*pending, pending_first, pending_cnt, pending_rc);
            pending = &l;
            pending_first = first;
            pending_cnt = cnt;
            pending_rc = rc;
        } else {
            launch(l, first, cnt, rc);
        }
    }
    if (pending)
        launch(*pending, pending_first, pending_cnt, pending_rc);
    for (auto& l : lanes_)
        drain(l);
    // The caller's stream sees the pixels: it waits (on the device, not on the host) for the last work of every lane.
    if (n_chunks > 1)
        for (auto& l : lanes_)
            if (l.used)
                cudaStreamWaitEvent(user_stream, l.dec->done_event(), 0);
    if (trace_enabled())
        std::fprintf(stderr, "[b200jpeg] decode n=%d chunks=%d pinned=%d total=%.0f us: walk %.0f layout %.0f stage %.0f h2d-issue %.0f launch %.0f wait %.0f\n", n,
            n_chunks, pinned, trace_now_us() - tr0, tr_walk, tr_layout, tr_stage, tr_h2d, tr_exec, tr_wait);
    return rc_all;
}

} // namespace b200jpeg
