// extern "C" surface declared in include/b200jpeg.h.
#include <algorithm>
#include <cstring>
#include <memory>
#include <new>
#include <queue>
#include <thread>
#include <vector>

#include "b200jpeg.h"
#include "decoder.h"

using b200jpeg::Decoder;
using b200jpeg::Pipeline;

struct b200jpegDecoder_st {
    Pipeline* pipe;
    Decoder* impl; // lane 0 of the pipeline: Prepare / Execute / Finish, kernel-level entry points, introspection
    bool last_was_pipeline;
};

namespace {
const char* kDecoderId = "b200_jpeg_decoder";
thread_local char g_err[256] = "";
} // namespace

extern "C" {

int b200jpegProbe(const void* data, size_t length, b200jpegStreamInfo_t* info)
{
    if (!data || !info)
        return B200JPEG_ERR_ARG;
    std::memset(info, 0, sizeof(*info));
    b200jpeg::StreamHeader hd;
    b200jpeg::StreamStatus st = b200jpeg::parse_stream(static_cast<const uint8_t*>(data), length, &hd);
    if (st == b200jpeg::kStreamMultiScan) { // three scans of one component each: decoded as three one-component images
        b200jpeg::StreamHeader sc[3];
        st = b200jpeg::parse_multiscan(static_cast<const uint8_t*>(data), length, &hd, sc);
        if (st == b200jpeg::kStreamOk) {
            hd.scan_offset = sc[0].scan_offset;
            for (int c = 0; c < 3; ++c)
                if (sc[c].scan_length >= (size_t(1) << 28))
                    hd.scan_length = sc[c].scan_length;
        }
    }
    if (st == b200jpeg::kStreamProgressive) { // SOF2: the scans are decoded one after the other into dense coefficient blocks
        b200jpeg::ProgStream pg;
        st = b200jpeg::parse_progressive(static_cast<const uint8_t*>(data), length, &hd, &pg);
        if (st == b200jpeg::kStreamOk) {
            hd.scan_offset = pg.scans[0].offset;
            for (const b200jpeg::ProgScan& ps : pg.scans)
                hd.scan_length = std::max(hd.scan_length, ps.length);
        }
    }
    info->width = hd.width;
    info->height = hd.height;
    info->num_components = hd.ncomp;
    for (int c = 0; c < 4; ++c) {
        info->h[c] = hd.h[c];
        info->v[c] = hd.v[c];
    }
    info->sof_marker = hd.sof;
    info->precision = hd.precision;
    info->restart_interval = hd.restart_interval;
    info->num_scans_seen = hd.nscans;
    info->scan_offset = hd.scan_offset;
    info->scan_length = hd.scan_length;
    info->mcus_x = hd.mcus_x;
    info->mcus_y = hd.mcus_y;
    info->blocks_per_mcu = hd.blocks_per_mcu;
    info->supported = (st == b200jpeg::kStreamOk) ? 1 : 0;
    switch (st) {
    case b200jpeg::kStreamOk: info->reason = B200JPEG_OK; break;
    case b200jpeg::kStreamNotJpeg: info->reason = B200JPEG_ERR_NOT_JPEG; break;
    case b200jpeg::kStreamUnsupported: info->reason = B200JPEG_ERR_UNSUPPORTED; break;
    default: info->reason = B200JPEG_ERR_CORRUPT; break;
    }
    if (st == b200jpeg::kStreamOk && hd.scan_length >= (size_t(1) << 28)) {
        info->supported = 0;
        info->reason = B200JPEG_ERR_TOO_LARGE;
    }
    return info->reason;
}

int b200jpegCreateEx(b200jpegDecoder_t* decoder, int device_id, const char* options, const b200jpegAllocators_t* allocators)
{
    if (!decoder)
        return B200JPEG_ERR_ARG;
    *decoder = nullptr;
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0) {
        cudaGetLastError();
        std::strncpy(g_err, "no CUDA device: this decoder has no CPU path", sizeof(g_err) - 1);
        return B200JPEG_ERR_CUDA;
    }
    if (device_id < 0) {
        if (cudaGetDevice(&device_id) != cudaSuccess)
            return B200JPEG_ERR_CUDA;
    }
    if (device_id >= count)
        return B200JPEG_ERR_ARG;
    b200jpeg::Services sv;
    if (allocators) {
        if (allocators->device_malloc && allocators->device_free) {
            sv.device_ctx = allocators->device_ctx;
            sv.device_malloc = reinterpret_cast<int (*)(void*, void**, size_t, cudaStream_t)>(allocators->device_malloc);
            sv.device_free = reinterpret_cast<int (*)(void*, void*, size_t, cudaStream_t)>(allocators->device_free);
        }
        if (allocators->pinned_malloc && allocators->pinned_free) {
            sv.pinned_ctx = allocators->pinned_ctx;
            sv.pinned_malloc = reinterpret_cast<int (*)(void*, void**, size_t, cudaStream_t)>(allocators->pinned_malloc);
            sv.pinned_free = reinterpret_cast<int (*)(void*, void*, size_t, cudaStream_t)>(allocators->pinned_free);
        }
    }
    auto* h = new (std::nothrow) b200jpegDecoder_st;
    if (!h)
        return B200JPEG_ERR_ARG;
    try {
        const b200jpeg::Options opt = b200jpeg::parse_options(options, kDecoderId);
        h->pipe = new Pipeline(device_id, opt, opt.lanes, sv);
    } catch (...) {
        delete h;
        return B200JPEG_ERR_STATE;
    }
    h->impl = &h->pipe->lane0();
    h->last_was_pipeline = false;
    *decoder = h;
    return B200JPEG_OK;
}

int b200jpegCreate(b200jpegDecoder_t* decoder, int device_id, const char* options)
{
    return b200jpegCreateEx(decoder, device_id, options, nullptr);
}

int b200jpegDestroy(b200jpegDecoder_t decoder)
{
    if (!decoder)
        return B200JPEG_OK;
    delete decoder->pipe;
    delete decoder;
    return B200JPEG_OK;
}

int b200jpegPrepareBatch(b200jpegDecoder_t decoder, int batch_size, const void* const* data, const size_t* lengths,
    const b200jpegOutput_t* outputs, void* cuda_stream)
{
    if (!decoder)
        return B200JPEG_ERR_ARG;
    try {
        decoder->last_was_pipeline = false;
        return decoder->impl->prepare(batch_size, data, lengths, outputs, static_cast<cudaStream_t>(cuda_stream));
    } catch (...) {
        return B200JPEG_ERR_STATE;
    }
}

int b200jpegExecuteBatch(b200jpegDecoder_t decoder, void* cuda_stream)
{
    if (!decoder)
        return B200JPEG_ERR_ARG;
    try {
        return decoder->impl->execute(static_cast<cudaStream_t>(cuda_stream));
    } catch (...) {
        return B200JPEG_ERR_STATE;
    }
}

int b200jpegFinishBatch(b200jpegDecoder_t decoder, int* status, void* cuda_stream)
{
    if (!decoder)
        return B200JPEG_ERR_ARG;
    try {
        return decoder->impl->finish(status, static_cast<cudaStream_t>(cuda_stream));
    } catch (...) {
        return B200JPEG_ERR_STATE;
    }
}

int b200jpegDecodeBatch(b200jpegDecoder_t decoder, int batch_size, const void* const* data, const size_t* lengths,
    const b200jpegOutput_t* outputs, int* status, void* cuda_stream)
{
    if (!decoder)
        return B200JPEG_ERR_ARG;
    try {
        decoder->last_was_pipeline = true;
        return decoder->pipe->decode(batch_size, data, lengths, outputs, status, static_cast<cudaStream_t>(cuda_stream));
    } catch (...) {
        return B200JPEG_ERR_STATE;
    }
}

int b200jpegScanSegment(b200jpegDecoder_t decoder, const void* dev_raw, size_t length, uint32_t expected_restart_markers, void* dev_out_words,
    uint32_t* dev_out_boundaries, uint32_t boundary_capacity, b200jpegScanResult_t* dev_result, void* cuda_stream)
{
    if (!decoder)
        return B200JPEG_ERR_ARG;
    static_assert(sizeof(b200jpegScanResult_t) == sizeof(b200jpeg::ImageState), "result layout");
    try {
        return decoder->impl->scan_segment(dev_raw, length, expected_restart_markers, dev_out_words, dev_out_boundaries, boundary_capacity,
            dev_result, static_cast<cudaStream_t>(cuda_stream));
    } catch (...) {
        return B200JPEG_ERR_STATE;
    }
}

// ------------------------------------------------------------------------------------------------
// multi-GPU driver: LPT partition by compressed bytes, one pipeline + stream + host thread per device
// ------------------------------------------------------------------------------------------------
struct b200jpegMultiDecoder_st {
    std::vector<int> devices;
    std::vector<std::unique_ptr<Pipeline>> pipes;
    std::vector<cudaStream_t> streams;
};

int b200jpegMultiCreate(b200jpegMultiDecoder_t* multi, const int* device_ids, int num_devices, const char* options)
{
    if (!multi || !device_ids || num_devices <= 0)
        return B200JPEG_ERR_ARG;
    *multi = nullptr;
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0) {
        cudaGetLastError();
        return B200JPEG_ERR_CUDA;
    }
    auto m = std::make_unique<b200jpegMultiDecoder_st>();
    const b200jpeg::Options opt = b200jpeg::parse_options(options, kDecoderId);
    for (int i = 0; i < num_devices; ++i) {
        if (device_ids[i] < 0 || device_ids[i] >= count)
            return B200JPEG_ERR_ARG;
        m->devices.push_back(device_ids[i]);
        m->pipes.push_back(std::make_unique<Pipeline>(device_ids[i], opt, opt.lanes));
        cudaStream_t s = nullptr;
        if (cudaSetDevice(device_ids[i]) != cudaSuccess || cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking) != cudaSuccess)
            return B200JPEG_ERR_CUDA;
        m->streams.push_back(s);
    }
    *multi = m.release();
    return B200JPEG_OK;
}

int b200jpegMultiDestroy(b200jpegMultiDecoder_t multi)
{
    if (!multi)
        return B200JPEG_OK;
    for (size_t i = 0; i < multi->streams.size(); ++i) {
        cudaSetDevice(multi->devices[i]);
        cudaStreamDestroy(multi->streams[i]);
    }
    delete multi;
    return B200JPEG_OK;
}

int b200jpegMultiPlan(b200jpegMultiDecoder_t multi, int batch_size, const size_t* lengths, int* device_of_image)
{
    if (!multi || batch_size < 0 || (batch_size > 0 && (!lengths || !device_of_image)))
        return B200JPEG_ERR_ARG;
    // longest-processing-time first: biggest stream to the least loaded device; ties by index -> deterministic
    std::vector<int> order(static_cast<size_t>(batch_size));
    for (int i = 0; i < batch_size; ++i)
        order[size_t(i)] = i;
    std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return lengths[a] > lengths[b]; });
    using Load = std::pair<uint64_t, int>;
    std::priority_queue<Load, std::vector<Load>, std::greater<Load>> heap;
    for (int d = 0; d < int(multi->devices.size()); ++d)
        heap.push({0, d});
    for (int i : order) {
        Load l = heap.top();
        heap.pop();
        device_of_image[i] = l.second;
        heap.push({l.first + lengths[i], l.second});
    }
    return B200JPEG_OK;
}

int b200jpegMultiDecodeBatch(b200jpegMultiDecoder_t multi, int batch_size, const void* const* data, const size_t* lengths,
    const b200jpegOutput_t* outputs, const int* device_of_image, int* status)
{
    if (!multi || batch_size < 0 || (batch_size > 0 && (!data || !lengths || !outputs || !device_of_image)))
        return B200JPEG_ERR_ARG;
    const int nd = int(multi->devices.size());
    std::vector<std::vector<int>> share(static_cast<size_t>(nd));
    for (int i = 0; i < batch_size; ++i) {
        if (device_of_image[i] < 0 || device_of_image[i] >= nd)
            return B200JPEG_ERR_ARG;
        share[size_t(device_of_image[i])].push_back(i);
    }
    std::vector<int> rc(static_cast<size_t>(nd), B200JPEG_OK);
    auto work = [&](int d) {
        const std::vector<int>& idx = share[size_t(d)];
        const int n = int(idx.size());
        if (n == 0)
            return;
        std::vector<const void*> dd(static_cast<size_t>(n));
        std::vector<size_t> ll(static_cast<size_t>(n));
        std::vector<b200jpegOutput_t> oo(static_cast<size_t>(n));
        std::vector<int> st(static_cast<size_t>(n), B200JPEG_ERR_STATE);
        for (int j = 0; j < n; ++j) {
            dd[size_t(j)] = data[idx[size_t(j)]];
            ll[size_t(j)] = lengths[idx[size_t(j)]];
            oo[size_t(j)] = outputs[idx[size_t(j)]];
        }
        try {
            rc[size_t(d)] = multi->pipes[size_t(d)]->decode(n, dd.data(), ll.data(), oo.data(), st.data(), multi->streams[size_t(d)]);
        } catch (...) {
            rc[size_t(d)] = B200JPEG_ERR_STATE;
        }
        if (status)
            for (int j = 0; j < n; ++j)
                status[idx[size_t(j)]] = st[size_t(j)];
    };
    std::vector<std::thread> threads;
    for (int d = 1; d < nd; ++d)
        threads.emplace_back(work, d);
    work(0);
    for (auto& t : threads)
        t.join();
    for (int d = 0; d < nd; ++d)
        if (rc[size_t(d)] != B200JPEG_OK)
            return rc[size_t(d)];
    return B200JPEG_OK;
}

int64_t b200jpegGetStat(b200jpegDecoder_t decoder, int which)
{
    if (!decoder)
        return -1;
    if (decoder->last_was_pipeline && which == B200JPEG_STAT_KERNEL_LAUNCHES)
        return decoder->pipe->last_launches();
    if (decoder->last_was_pipeline && which == B200JPEG_STAT_RESYNC_IMAGES)
        return decoder->pipe->last_resync();
    return decoder->impl->stat(which);
}

int b200jpegSetProfiling(b200jpegDecoder_t decoder, int enable)
{
    if (!decoder)
        return B200JPEG_ERR_ARG;
    decoder->impl->set_profiling(enable != 0);
    return B200JPEG_OK;
}

int b200jpegGetStageTimes(b200jpegDecoder_t decoder, float* ms, const char** names, int capacity)
{
    if (!decoder || !ms)
        return 0;
    return decoder->impl->stage_times(ms, names, capacity);
}

int b200jpegDebugCopyCoefficients(b200jpegDecoder_t decoder, int index, int16_t* host_out, size_t capacity_elems)
{
    if (!decoder || !host_out)
        return B200JPEG_ERR_ARG;
    return decoder->impl->copy_coefficients(index, host_out, capacity_elems);
}

const char* b200jpegGetLastError(b200jpegDecoder_t decoder)
{
    if (!decoder)
        return g_err;
    return decoder->last_was_pipeline ? decoder->pipe->last_error() : decoder->impl->last_error();
}

const char* b200jpegVersion(void) { return "b200jpeg 0.1.0 (sm_100a)"; }

} // extern "C"
