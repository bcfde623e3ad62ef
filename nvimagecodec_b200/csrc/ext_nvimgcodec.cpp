// nvImageCodec extension adapter: registers the b200 JPEG decoder through the reference's
// C extension ABI so that nvimgcodecDecoderDecode / nvimgcodec.Decoder / nvimtrans pick it up
// by codec priority with no caller changes.
//
//   entry point        nvimgcodecExtensionModuleEntry            include/nvimgcodec.h:1696
//   extension desc     nvimgcodecExtensionDesc_t                 include/nvimgcodec.h:1652-1680
//   decoder desc       nvimgcodecDecoderDesc_t                   include/nvimgcodec.h:1438-1542
//   behaviour mirrored extensions/libjpeg_turbo/libjpeg_turbo_decoder.cpp:110-212 (canDecode),
//                      :295-420 (decode); batched shape as extensions/nvjpeg/hw_decoder.cpp:636-888
//
// The adapter only translates descriptors; all work happens in b200jpeg::Decoder (decoder.cpp),
// the same object the plain C ABI (include/b200jpeg.h) drives.
#include <nvimgcodec.h>

#include <algorithm>
#include <atomic>
#include <cstdio>
#include <cstring>
#include <deque>
#include <memory>
#include <mutex>
#include <sstream>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

#include "b200jpeg.h"
#include "decoder.h"

namespace {

constexpr const char* kExtensionId = "b200_jpeg_ext";
constexpr const char* kDecoderId = "b200_jpeg_decoder";
constexpr float kPriority = 50.0f;      // ahead of nvjpeg HW (100), nvjpeg CUDA (200), libjpeg_turbo (300)
constexpr int kMiniBatch = 65536;       // keeps the framework's first-batch share cap from ever biting

void log_msg(const nvimgcodecFrameworkDesc_t* fw, nvimgcodecDebugMessageSeverity_t sev, const std::string& msg)
{
    if (!fw || !fw->log)
        return;
    nvimgcodecDebugMessageData_t data{};
    data.struct_type = NVIMGCODEC_STRUCTURE_TYPE_DEBUG_MESSAGE_DATA;
    data.struct_size = sizeof(data);
    data.message = msg.c_str();
    data.codec_id = kDecoderId;
    data.codec_version = NVIMGCODEC_VER;
    fw->log(fw->instance, sev, NVIMGCODEC_DEBUG_MESSAGE_CATEGORY_GENERAL, &data);
}

struct MappedStream {
    const uint8_t* data = nullptr;
    size_t size = 0;
    std::vector<uint8_t> owned; // used when the io stream cannot be mapped
};

bool map_stream(const nvimgcodecCodeStreamDesc_t* cs, MappedStream* out)
{
    nvimgcodecIoStreamDesc_t* io = cs ? cs->io_stream : nullptr;
    if (!io)
        return false;
    size_t size = 0;
    if (io->size(io->instance, &size) != NVIMGCODEC_STATUS_SUCCESS || size == 0)
        return false;
    void* ptr = nullptr;
    if (io->map && io->map(io->instance, &ptr, 0, size) == NVIMGCODEC_STATUS_SUCCESS && ptr) {
        out->data = static_cast<const uint8_t*>(ptr);
        out->size = size;
        return true;
    }
    out->owned.resize(size);
    size_t got = 0;
    if (io->seek(io->instance, 0, SEEK_SET) != NVIMGCODEC_STATUS_SUCCESS)
        return false;
    if (io->read(io->instance, &got, out->owned.data(), size) != NVIMGCODEC_STATUS_SUCCESS || got != size)
        return false;
    out->data = out->owned.data();
    out->size = size;
    return true;
}

int to_exif_code(const nvimgcodecOrientation_t& o);

// A region that reaches outside the image (what nvjpeg_ext accepts and fills, extensions/nvjpeg/cuda_decoder.cpp:763-766 with
// src/imgproc/out_of_bound_roi_fill.cpp:90-199): the decoder writes the part inside the image at its place in the caller's
// buffer, the border is set to the region's constant samples -- one value per plane -- with 2-D memsets on the batch's stream.
struct BorderFill {
    bool needed = false;
    uint8_t* base = nullptr;   // the caller's buffer
    size_t row_stride = 0, plane_bytes = 0;
    int planes = 1, bytes_per_px = 1;
    int width = 0, height = 0; // the whole region, in pixels
    int in_x0 = 0, in_y0 = 0, in_x1 = 0, in_y1 = 0; // the part inside the image, in region coordinates
    int value[3] = {0, 0, 0};
};

uint32_t fill_bits(const nvimgcodecRegion_t& region, uint32_t comp) // get_fill_bits, out_of_bound_roi_fill.cpp:48-88, for 8-bit samples
{
    const auto& sm = region.out_of_bounds_samples[std::min<uint32_t>(comp, NVIMGCODEC_MAX_ROI_FILL_CHANNELS - 1)];
    switch (sm.type) {
    case NVIMGCODEC_SAMPLE_DATA_TYPE_UINT8:
    case NVIMGCODEC_SAMPLE_DATA_TYPE_UINT16:
    case NVIMGCODEC_SAMPLE_DATA_TYPE_UINT32: return sm.value.as_uint;
    case NVIMGCODEC_SAMPLE_DATA_TYPE_INT8:
    case NVIMGCODEC_SAMPLE_DATA_TYPE_INT16:
    case NVIMGCODEC_SAMPLE_DATA_TYPE_INT32: return uint32_t(sm.value.as_int);
    default: return 0;
    }
}

// Translate the requested nvimgcodecImageInfo_t into a b200jpegOutput_t.  Returns a processing
// status: SUCCESS or an OR of the *_UNSUPPORTED bits (include/nvimgcodec.h:735-768).
nvimgcodecProcessingStatus_t translate_output(const nvimgcodecImageInfo_t& info, const nvimgcodecRegion_t* region, int img_w, int img_h,
    bool apply_orientation, b200jpegOutput_t* out, BorderFill* border = nullptr)
{
    nvimgcodecProcessingStatus_t st = NVIMGCODEC_PROCESSING_STATUS_SUCCESS;
    std::memset(out, 0, sizeof(*out));
    int want_planes = 1, want_channels = 3;
    switch (info.sample_format) {
    case NVIMGCODEC_SAMPLEFORMAT_I_RGB: out->format = B200JPEG_FMT_I_RGB; break;
    case NVIMGCODEC_SAMPLEFORMAT_I_BGR: out->format = B200JPEG_FMT_I_BGR; break;
    case NVIMGCODEC_SAMPLEFORMAT_P_RGB:
        out->format = B200JPEG_FMT_P_RGB;
        want_planes = 3;
        want_channels = 1;
        break;
    case NVIMGCODEC_SAMPLEFORMAT_P_BGR:
        out->format = B200JPEG_FMT_P_BGR;
        want_planes = 3;
        want_channels = 1;
        break;
    case NVIMGCODEC_SAMPLEFORMAT_P_Y:
    case NVIMGCODEC_SAMPLEFORMAT_I_Y:
        out->format = B200JPEG_FMT_Y;
        want_channels = 1;
        break;
    case NVIMGCODEC_SAMPLEFORMAT_I_YUV: out->format = B200JPEG_FMT_I_YUV; break;
    case NVIMGCODEC_SAMPLEFORMAT_P_YUV:
        out->format = B200JPEG_FMT_P_YUV;
        want_planes = 3;
        want_channels = 1;
        break;
    case NVIMGCODEC_SAMPLEFORMAT_I_UNCHANGED: // "as stored": RGB for colour streams, Y for gray ones
        out->format = info.plane_info[0].num_channels == 1 ? B200JPEG_FMT_Y : B200JPEG_FMT_I_RGB;
        want_channels = info.plane_info[0].num_channels == 1 ? 1 : 3;
        break;
    case NVIMGCODEC_SAMPLEFORMAT_P_UNCHANGED:
        out->format = info.num_planes == 1 ? B200JPEG_FMT_Y : B200JPEG_FMT_P_RGB;
        want_planes = info.num_planes == 1 ? 1 : 3;
        want_channels = 1;
        break;
    default: st |= NVIMGCODEC_PROCESSING_STATUS_SAMPLE_FORMAT_UNSUPPORTED; break;
    }
    if (info.num_planes != uint32_t(want_planes))
        st |= NVIMGCODEC_PROCESSING_STATUS_NUM_PLANES_UNSUPPORTED;
    if (info.plane_info[0].num_channels != uint32_t(want_channels))
        st |= NVIMGCODEC_PROCESSING_STATUS_NUM_CHANNELS_UNSUPPORTED;
    if (info.plane_info[0].sample_type != NVIMGCODEC_SAMPLE_DATA_TYPE_UINT8)
        st |= NVIMGCODEC_PROCESSING_STATUS_SAMPLE_TYPE_UNSUPPORTED;
    int rw = img_w, rh = img_h;
    if (region && region->ndim != 0) {
        if (region->ndim != 2) {
            st |= NVIMGCODEC_PROCESSING_STATUS_ROI_UNSUPPORTED;
        } else {
            // start/end are [y, x] (extensions/libjpeg_turbo/libjpeg_turbo_decoder.cpp:359-362)
            const int y0 = region->start[0], x0 = region->start[1], y1 = region->end[0], x1 = region->end[1];
            const bool oob = y0 < 0 || x0 < 0 || y1 > img_h || x1 > img_w;
            const int cy0 = std::max(y0, 0), cx0 = std::max(x0, 0), cy1 = std::min(y1, img_h), cx1 = std::min(x1, img_w);
            if (y0 >= y1 || x0 >= x1) {
                st |= NVIMGCODEC_PROCESSING_STATUS_ROI_UNSUPPORTED;
            } else if (oob) {
                // Out of bounds: filled like nvjpeg_ext does where that is well defined -- constant policy, device buffer, some of
                // the region inside the image, one fill value per plane (out_of_bound_roi_fill.cpp:104-107, 190-192)
                bool ok = cy0 < cy1 && cx0 < cx1 && region->out_of_bounds_policy == NVIMGCODEC_OUT_OF_BOUNDS_POLICY_CONSTANT &&
                          info.buffer_kind == NVIMGCODEC_IMAGE_BUFFER_KIND_STRIDED_DEVICE;
                if (want_planes == 1 && want_channels == 3)
                    ok = ok && (fill_bits(*region, 0) & 0xFFu) == (fill_bits(*region, 1) & 0xFFu) && (fill_bits(*region, 0) & 0xFFu) == (fill_bits(*region, 2) & 0xFFu);
                if (!ok) {
                    st |= NVIMGCODEC_PROCESSING_STATUS_ROI_UNSUPPORTED;
                } else {
                    out->roi_x0 = cx0;
                    out->roi_y0 = cy0;
                    out->roi_x1 = cx1;
                    out->roi_y1 = cy1;
                    rw = x1 - x0; // the image the caller holds has the size of the whole region
                    rh = y1 - y0;
                    if (border) {
                        border->needed = true;
                        border->width = rw;
                        border->height = rh;
                        border->in_x0 = cx0 - x0;
                        border->in_y0 = cy0 - y0;
                        border->in_x1 = cx1 - x0;
                        border->in_y1 = cy1 - y0;
                        for (int p = 0; p < 3; ++p)
                            border->value[p] = int(fill_bits(*region, want_planes == 3 ? uint32_t(p) : 0u) & 0xFFu);
                    }
                }
            } else {
                out->roi_x0 = x0;
                out->roi_y0 = y0;
                out->roi_x1 = x1;
                out->roi_y1 = y1;
                rw = x1 - x0;
                rh = y1 - y0;
            }
        }
    }
    // EXIF orientation (nvimgcodecDecodeParams_t::apply_exif_orientation): applied while storing; the caller's image
    // already has the oriented size (python/decoder.cpp and example/nvimtrans/main.cpp:487-490 swap it for 90/270)
    int exif = 1;
    if (apply_orientation) {
        exif = to_exif_code(info.orientation);
        if (exif == 0)
            st |= NVIMGCODEC_PROCESSING_STATUS_ORIENTATION_UNSUPPORTED;
        else if (exif > 1 && region && region->ndim != 0)
            st |= NVIMGCODEC_PROCESSING_STATUS_ROI_UNSUPPORTED; // which frame a region refers to is not defined by the API
    }
    out->orientation = exif;
    if (exif >= 5)
        std::swap(rw, rh);
    if (int(info.plane_info[0].width) < rw || int(info.plane_info[0].height) < rh)
        st |= NVIMGCODEC_PROCESSING_STATUS_RESOLUTION_UNSUPPORTED;
    out->buffer = info.buffer;
    out->buffer_is_host = (info.buffer_kind == NVIMGCODEC_IMAGE_BUFFER_KIND_STRIDED_HOST) ? 1 : 0;
    out->row_stride = info.plane_info[0].row_stride;
    if (border && border->needed) {
        if (exif > 1)
            st |= NVIMGCODEC_PROCESSING_STATUS_ROI_UNSUPPORTED;
        border->base = static_cast<uint8_t*>(info.buffer);
        border->row_stride = info.plane_info[0].row_stride;
        border->planes = want_planes;
        border->bytes_per_px = want_planes == 1 ? want_channels : 1;
        border->plane_bytes = border->row_stride * size_t(info.plane_info[0].height);
        // the decoder writes the inside part where it belongs in the caller's image
        out->buffer = border->base + size_t(border->in_y0) * border->row_stride + size_t(border->in_x0) * border->bytes_per_px;
        out->plane_stride = border->plane_bytes;
    }
    // planes of a planar image lie back to back at row_stride*height (extensions/nvjpeg/cuda_decoder.cpp:716-722);
    // the kernels assume exactly that with height == ROI height
    if (want_planes == 3 && int(info.plane_info[0].height) != rh)
        st |= NVIMGCODEC_PROCESSING_STATUS_RESOLUTION_UNSUPPORTED;
    return st;
}

// nvimgcodecOrientation_t (rotation counter-clockwise in degrees + flips, as src/parsers/exif_orientation.h builds it from
// the EXIF tag) back to the EXIF code the decoder applies.  Returns 0 for combinations the parser never produces.
int to_exif_code(const nvimgcodecOrientation_t& o)
{
    const int rot = ((o.rotated % 360) + 360) % 360;
    if (rot == 0 && !o.flip_x && !o.flip_y)
        return 1;
    if (rot == 0 && o.flip_x && !o.flip_y)
        return 2;
    if (rot == 180 && !o.flip_x && !o.flip_y)
        return 3;
    if (rot == 0 && !o.flip_x && o.flip_y)
        return 4;
    if (rot == 90 && !o.flip_x && o.flip_y)
        return 5;
    if (rot == 270 && !o.flip_x && !o.flip_y)
        return 6;
    if (rot == 270 && !o.flip_x && o.flip_y)
        return 7;
    if (rot == 90 && !o.flip_x && !o.flip_y)
        return 8;
    return 0;
}

nvimgcodecProcessingStatus_t to_processing_status(int b200_status)
{
    switch (b200_status) {
    case B200JPEG_OK: return NVIMGCODEC_PROCESSING_STATUS_SUCCESS;
    case B200JPEG_ERR_UNSUPPORTED:
    case B200JPEG_ERR_TOO_LARGE: return NVIMGCODEC_PROCESSING_STATUS_CODESTREAM_UNSUPPORTED;
    case B200JPEG_ERR_CORRUPT:
    case B200JPEG_ERR_NOT_JPEG: return NVIMGCODEC_PROCESSING_STATUS_IMAGE_CORRUPTED;
    default: return NVIMGCODEC_PROCESSING_STATUS_FAIL;
    }
}

// ---- what the host framework lends a plugin: its allocators and its executor (SURVEY.md 8b "Memory / ownership")
// One job handed to the framework's executor.  The helpers pull indices from a shared counter; the caller takes part
// and only waits until every index is done, so helpers that the executor starts late (its threads may be busy with
// the framework's own per-sample work) find nothing to do and drop their reference.
struct ParJob {
    std::atomic<int> next{0}, done{0}, refs{1};
    int n = 0;
    void (*fn)(int, void*) = nullptr;
    void* arg = nullptr;
    static void drop(ParJob* j)
    {
        if (j->refs.fetch_sub(1) == 1)
            delete j;
    }
    static void worker(int, int, void* ctx)
    {
        ParJob* j = static_cast<ParJob*>(ctx);
        for (int i; (i = j->next.fetch_add(1)) < j->n;) {
            j->fn(i, j->arg);
            j->done.fetch_add(1, std::memory_order_release);
        }
        drop(j);
    }
};

struct ExecutorCtx {
    nvimgcodecExecutorDesc_t* executor = nullptr;
    int device_id = 0;
};

void executor_parallel_for(void* ctx, int n, void (*fn)(int, void*), void* arg)
{
    auto* ec = static_cast<ExecutorCtx*>(ctx);
    nvimgcodecExecutorDesc_t* ex = ec->executor;
    const int helpers = std::min(n - 1, std::min(8, ex->getNumThreads(ex->instance)));
    ParJob* job = new ParJob;
    job->n = n;
    job->fn = fn;
    job->arg = arg;
    int scheduled = 0;
    for (int h = 0; h < helpers; ++h) {
        job->refs.fetch_add(1);
        if (ex->schedule(ex->instance, ec->device_id, h, job, &ParJob::worker) == NVIMGCODEC_STATUS_SUCCESS)
            ++scheduled;
        else
            job->refs.fetch_sub(1);
    }
    if (scheduled)
        ex->run(ex->instance, ec->device_id);
    for (int i; (i = job->next.fetch_add(1)) < n;) {
        fn(i, arg);
        job->done.fetch_add(1, std::memory_order_release);
    }
    while (job->done.load(std::memory_order_acquire) < n)
        std::this_thread::yield();
    ParJob::drop(job);
}

class DecoderImpl {
  public:
    DecoderImpl(const nvimgcodecFrameworkDesc_t* fw, const nvimgcodecExecutionParams_t* ep, const char* options)
        : fw_(fw)
        , ep_(ep)
        , opt_(b200jpeg::parse_options(options, kDecoderId))
    {
        int n = 1;
        if (ep_->executor && ep_->executor->getNumThreads)
            n = ep_->executor->getNumThreads(ep_->executor->instance) + 1;
        walked_.resize(size_t(n) + 1);
        // memory and helper threads come from the host when it offers them (include/nvimgcodec.h:246-281, 1045-1084),
        // as src/imgproc/device_buffer.cpp:42-69 and the reference's plugins do
        if (ep_->device_allocator && ep_->device_allocator->device_malloc && ep_->device_allocator->device_free) {
            sv_.device_ctx = ep_->device_allocator->device_ctx;
            sv_.device_malloc = ep_->device_allocator->device_malloc;
            sv_.device_free = ep_->device_allocator->device_free;
        }
        if (ep_->pinned_allocator && ep_->pinned_allocator->pinned_malloc && ep_->pinned_allocator->pinned_free) {
            sv_.pinned_ctx = ep_->pinned_allocator->pinned_ctx;
            sv_.pinned_malloc = ep_->pinned_allocator->pinned_malloc;
            sv_.pinned_free = ep_->pinned_allocator->pinned_free;
        }
        if (opt_.use_executor && ep_->executor && ep_->executor->schedule && ep_->executor->run && ep_->executor->getNumThreads &&
            ep_->executor->getNumThreads(ep_->executor->instance) > 0) {
            exec_ctx_.executor = ep_->executor;
            exec_ctx_.device_id = ep_->device_id;
            sv_.par_ctx = &exec_ctx_;
            sv_.parallel_for = &executor_parallel_for;
        }
    }

    nvimgcodecProcessingStatus_t can_decode(const nvimgcodecImageDesc_t* image, const nvimgcodecCodeStreamDesc_t* cs,
        const nvimgcodecDecodeParams_t* params, int thread_idx)
    {
        if (!image || !cs || !params)
            return NVIMGCODEC_PROCESSING_STATUS_FAIL;
        struct Timer {
            std::atomic<long>& acc;
            double t0;
            explicit Timer(std::atomic<long>& a) : acc(a), t0(b200jpeg::trace_enabled() ? b200jpeg::trace_now_us() : 0) {}
            ~Timer()
            {
                if (b200jpeg::trace_enabled())
                    acc.fetch_add(long(b200jpeg::trace_now_us() - t0));
            }
        } timer(can_decode_us_);
        nvimgcodecJpegImageInfo_t jpeg_info{NVIMGCODEC_STRUCTURE_TYPE_JPEG_IMAGE_INFO, sizeof(nvimgcodecJpegImageInfo_t), nullptr};
        nvimgcodecImageInfo_t cs_info{NVIMGCODEC_STRUCTURE_TYPE_IMAGE_INFO, sizeof(nvimgcodecImageInfo_t), &jpeg_info};
        if (cs->getImageInfo(cs->instance, &cs_info) != NVIMGCODEC_STATUS_SUCCESS)
            return NVIMGCODEC_PROCESSING_STATUS_FAIL;
        if (std::strcmp(cs_info.codec_name, "jpeg") != 0)
            return NVIMGCODEC_PROCESSING_STATUS_CODEC_UNSUPPORTED;
        if (jpeg_info.encoding != NVIMGCODEC_JPEG_ENCODING_BASELINE_DCT &&
            jpeg_info.encoding != NVIMGCODEC_JPEG_ENCODING_EXTENDED_SEQUENTIAL_DCT_HUFFMAN &&
            !(jpeg_info.encoding == NVIMGCODEC_JPEG_ENCODING_PROGRESSIVE_DCT_HUFFMAN && opt_.progressive))
            return NVIMGCODEC_PROCESSING_STATUS_ENCODING_UNSUPPORTED;
        nvimgcodecImageInfo_t info{NVIMGCODEC_STRUCTURE_TYPE_IMAGE_INFO, sizeof(nvimgcodecImageInfo_t), nullptr};
        if (image->getImageInfo(image->instance, &info) != NVIMGCODEC_STATUS_SUCCESS)
            return NVIMGCODEC_PROCESSING_STATUS_FAIL;
        nvimgcodecCodeStreamInfo_t csi{NVIMGCODEC_STRUCTURE_TYPE_CODE_STREAM_INFO, sizeof(nvimgcodecCodeStreamInfo_t), nullptr};
        if (cs->getCodeStreamInfo(cs->instance, &csi) != NVIMGCODEC_STATUS_SUCCESS)
            return NVIMGCODEC_PROCESSING_STATUS_FAIL;
        nvimgcodecProcessingStatus_t st = NVIMGCODEC_PROCESSING_STATUS_SUCCESS;
        const nvimgcodecRegion_t* region = nullptr;
        if (csi.code_stream_view) {
            if (csi.code_stream_view->image_idx != 0)
                st |= NVIMGCODEC_PROCESSING_STATUS_NUM_IMAGES_UNSUPPORTED;
            region = &csi.code_stream_view->region;
        }
        b200jpegOutput_t out;
        BorderFill border;
        st |= translate_output(info, region, int(cs_info.plane_info[0].width), int(cs_info.plane_info[0].height),
            params->apply_exif_orientation != 0, &out, &border);
        // YCbCr output only for YCbCr streams without subsampling, as libjpeg_turbo_decoder.cpp:180-189
        if ((info.sample_format == NVIMGCODEC_SAMPLEFORMAT_I_YUV || info.sample_format == NVIMGCODEC_SAMPLEFORMAT_P_YUV) &&
            (cs_info.color_spec != NVIMGCODEC_COLORSPEC_SYCC || cs_info.chroma_subsampling != NVIMGCODEC_SAMPLING_444))
            st |= NVIMGCODEC_PROCESSING_STATUS_SAMPLE_FORMAT_UNSUPPORTED;
        if (st != NVIMGCODEC_PROCESSING_STATUS_SUCCESS)
            return st;
        // stream-level acceptance: our own marker walk (the built-in parser stops before DQT/DHT/DRI/SOS).  Its result
        // is kept for decodeBatch, which would otherwise walk the same bytes again.
        MappedStream ms;
        if (!map_stream(cs, &ms))
            return NVIMGCODEC_PROCESSING_STATUS_FAIL;
        auto walked = std::make_unique<Walked>();
        walked->data = ms.owned.empty() ? ms.data : nullptr; // an unmapped stream was read into a temporary: nothing to key on
        walked->size = ms.size;
        b200jpeg::parse_input(ms.data, ms.size, &walked->parsed);
        const b200jpeg::ParsedInput& pi = walked->parsed;
        b200jpeg::StreamStatus ss = pi.status;
        size_t longest_scan = pi.frame.scan_length;
        if (ss == b200jpeg::kStreamMultiScan) {
            ss = pi.multi_status;
            for (int c = 0; c < 3 && pi.scans; ++c)
                longest_scan = std::max(longest_scan, pi.scans[c].scan_length);
        } else if (ss == b200jpeg::kStreamProgressive) {
            ss = pi.multi_status;
            for (size_t k = 0; pi.prog && k < pi.prog->scans.size(); ++k)
                longest_scan = std::max(longest_scan, pi.prog->scans[k].length);
        }
        if (ss == b200jpeg::kStreamOk && longest_scan >= (size_t(1) << 28))
            return NVIMGCODEC_PROCESSING_STATUS_CODESTREAM_UNSUPPORTED;
        if (ss == b200jpeg::kStreamOk) {
            if (walked->data)
                remember(thread_idx, cs, std::move(walked));
            return NVIMGCODEC_PROCESSING_STATUS_SUCCESS;
        }
        if (ss == b200jpeg::kStreamCorrupt || ss == b200jpeg::kStreamNotJpeg)
            return NVIMGCODEC_PROCESSING_STATUS_IMAGE_CORRUPTED;
        return NVIMGCODEC_PROCESSING_STATUS_CODESTREAM_UNSUPPORTED;
    }

    // Shared by decodeBatch (the whole batch on one framework stream, the chunked pipeline) and the per-sample
    // decode safety net (one extra single-lane pipeline, serialised, released by the next batch).
    nvimgcodecStatus_t decode_batch(const nvimgcodecImageDesc_t** images, const nvimgcodecCodeStreamDesc_t** css, int n, bool batched,
        bool apply_orientation)
    {
        if (n <= 0)
            return NVIMGCODEC_STATUS_SUCCESS;
        const double tr0 = b200jpeg::trace_enabled() ? b200jpeg::trace_now_us() : 0;
        std::unique_lock<std::mutex> fallback_lock(fallback_mu_, std::defer_lock);
        b200jpeg::Pipeline* core = nullptr;
        try {
            if (batched) {
                {
                    std::lock_guard<std::mutex> lk(fallback_mu_);
                    fallback_.reset(); // what a per-sample retry allocated is not kept across batches
                }
                if (!batch_)
                    batch_ = std::make_unique<b200jpeg::Pipeline>(ep_->device_id, opt_, opt_.lanes, sv_);
                core = batch_.get();
            } else {
                fallback_lock.lock();
                if (!fallback_)
                    fallback_ = std::make_unique<b200jpeg::Pipeline>(ep_->device_id, opt_, 1, sv_);
                core = fallback_.get();
            }
        } catch (...) {
            core = nullptr;
        }
        if (!core) {
            for (int i = 0; i < n; ++i)
                images[i]->imageReady(images[i]->instance, NVIMGCODEC_PROCESSING_STATUS_FAIL);
            return NVIMGCODEC_STATUS_EXTENSION_INTERNAL_ERROR;
        }
        std::vector<MappedStream> streams(n);
        std::vector<const void*> data(n, nullptr);
        std::vector<size_t> len(n, 0);
        std::vector<b200jpegOutput_t> outs(n);
        std::vector<nvimgcodecProcessingStatus_t> pre(n, NVIMGCODEC_PROCESSING_STATUS_SUCCESS);
        std::vector<int> status(n, B200JPEG_ERR_STATE);
        std::vector<std::unique_ptr<Walked>> walked(n);
        std::vector<BorderFill> borders; // samples whose region reaches outside the image
        std::vector<const b200jpeg::ParsedInput*> parsed(n, nullptr);
        cudaStream_t stream = nullptr;
        bool have_stream = false;
        for (int i = 0; i < n; ++i) {
            nvimgcodecImageInfo_t info{NVIMGCODEC_STRUCTURE_TYPE_IMAGE_INFO, sizeof(nvimgcodecImageInfo_t), nullptr};
            nvimgcodecImageInfo_t cs_info{NVIMGCODEC_STRUCTURE_TYPE_IMAGE_INFO, sizeof(nvimgcodecImageInfo_t), nullptr};
            nvimgcodecCodeStreamInfo_t csi{NVIMGCODEC_STRUCTURE_TYPE_CODE_STREAM_INFO, sizeof(nvimgcodecCodeStreamInfo_t), nullptr};
            if (images[i]->getImageInfo(images[i]->instance, &info) != NVIMGCODEC_STATUS_SUCCESS ||
                css[i]->getImageInfo(css[i]->instance, &cs_info) != NVIMGCODEC_STATUS_SUCCESS ||
                css[i]->getCodeStreamInfo(css[i]->instance, &csi) != NVIMGCODEC_STATUS_SUCCESS || !map_stream(css[i], &streams[i])) {
                pre[i] = NVIMGCODEC_PROCESSING_STATUS_FAIL;
                std::memset(&outs[i], 0, sizeof(outs[i]));
                continue;
            }
            if (!have_stream) {
                stream = info.cuda_stream; // every sample of a batch shares the framework's stream (hw_decoder.cpp:687-697)
                have_stream = true;
            }
            const nvimgcodecRegion_t* region = csi.code_stream_view ? &csi.code_stream_view->region : nullptr;
            BorderFill border;
            pre[i] = translate_output(info, region, int(cs_info.plane_info[0].width), int(cs_info.plane_info[0].height), apply_orientation,
                &outs[i], &border);
            if (pre[i] == NVIMGCODEC_PROCESSING_STATUS_SUCCESS && border.needed)
                borders.push_back(border);
            if (pre[i] == NVIMGCODEC_PROCESSING_STATUS_SUCCESS) {
                data[i] = streams[i].data;
                len[i] = streams[i].size;
                walked[i] = recall(css[i], streams[i].data, streams[i].size);
                if (walked[i])
                    parsed[i] = &walked[i]->parsed;
            }
        }
        for (const BorderFill& b : borders) { // four strips around the part inside the image, per plane
            for (int p = 0; p < b.planes; ++p) {
                uint8_t* plane = b.base + size_t(p) * b.plane_bytes;
                const size_t bpp = size_t(b.bytes_per_px);
                auto fill = [&](int x, int y, int w, int h) {
                    if (w > 0 && h > 0)
                        cudaMemset2DAsync(plane + size_t(y) * b.row_stride + size_t(x) * bpp, b.row_stride, b.value[p], size_t(w) * bpp, size_t(h), stream);
                };
                fill(0, 0, b.width, b.in_y0);
                fill(0, b.in_y1, b.width, b.height - b.in_y1);
                fill(0, b.in_y0, b.in_x0, b.in_y1 - b.in_y0);
                fill(b.in_x1, b.in_y0, b.width - b.in_x1, b.in_y1 - b.in_y0);
            }
        }
        const double tr1 = b200jpeg::trace_enabled() ? b200jpeg::trace_now_us() : 0;
        const int rc = core->decode(n, data.data(), len.data(), outs.data(), status.data(), stream, parsed.data());
        if (b200jpeg::trace_enabled()) {
            int reused = 0;
            for (int i = 0; i < n; ++i)
                reused += parsed[i] != nullptr;
            std::fprintf(stderr, "[b200jpeg] decodeBatch n=%d: descriptors %.0f us, decode %.0f us, canDecode since last batch %ld us (summed over threads), walks reused %d, since last decodeBatch return %.0f us\n",
                n, tr1 - tr0, b200jpeg::trace_now_us() - tr1, can_decode_us_.exchange(0), reused, last_return_us_ ? tr0 - last_return_us_ : 0.0);
        }
        if (rc != B200JPEG_OK)
            log_msg(fw_, NVIMGCODEC_DEBUG_MESSAGE_SEVERITY_ERROR, std::string("b200 jpeg batch decode failed: ") + core->last_error());
        for (int i = 0; i < n; ++i) {
            nvimgcodecProcessingStatus_t ps = pre[i];
            if (ps == NVIMGCODEC_PROCESSING_STATUS_SUCCESS)
                ps = (rc == B200JPEG_OK) ? to_processing_status(status[i]) : NVIMGCODEC_PROCESSING_STATUS_FAIL;
            images[i]->imageReady(images[i]->instance, ps);
        }
        if (b200jpeg::trace_enabled())
            last_return_us_ = b200jpeg::trace_now_us();
        return rc == B200JPEG_OK ? NVIMGCODEC_STATUS_SUCCESS : NVIMGCODEC_STATUS_EXTENSION_EXECUTION_FAILED;
    }

  private:
    // Marker walks done by canDecode, waiting for the decode call that follows it in the same nvimgcodecDecoderDecode.
    // canDecode runs on executor threads (thread_idx in [0, N]): one table per thread, each with its own lock.
    struct Walked {
        const void* data = nullptr;
        size_t size = 0;
        b200jpeg::ParsedInput parsed;
    };
    struct WalkTable {
        std::mutex mu;
        std::unordered_map<const void*, std::unique_ptr<Walked>> map;
    };
    void remember(int thread_idx, const nvimgcodecCodeStreamDesc_t* cs, std::unique_ptr<Walked> w)
    {
        const size_t slot = (thread_idx < 0 || size_t(thread_idx) >= walked_.size()) ? walked_.size() - 1 : size_t(thread_idx);
        WalkTable& t = walked_[slot];
        std::lock_guard<std::mutex> lk(t.mu);
        if (t.map.size() > 16384) // walked but never decoded by us (another decoder took them): do not let them pile up
            t.map.clear();
        t.map[cs->instance] = std::move(w);
    }
    std::unique_ptr<Walked> recall(const nvimgcodecCodeStreamDesc_t* cs, const void* data, size_t size)
    {
        for (WalkTable& t : walked_) {
            std::lock_guard<std::mutex> lk(t.mu);
            auto it = t.map.find(cs->instance);
            if (it == t.map.end())
                continue;
            std::unique_ptr<Walked> w = std::move(it->second);
            t.map.erase(it);
            if (w && w->data == data && w->size == size) // the same mapped bytes canDecode looked at
                return w;
            return nullptr;
        }
        return nullptr;
    }

    const nvimgcodecFrameworkDesc_t* fw_;
    const nvimgcodecExecutionParams_t* ep_;
    b200jpeg::Options opt_;
    b200jpeg::Services sv_;
    ExecutorCtx exec_ctx_;
    std::deque<WalkTable> walked_;
    std::unique_ptr<b200jpeg::Pipeline> batch_;
    std::mutex fallback_mu_;
    std::unique_ptr<b200jpeg::Pipeline> fallback_;
    std::atomic<long> can_decode_us_{0};
    double last_return_us_ = 0;
};

// ---- C trampolines of nvimgcodecDecoderDesc_t (never let an exception cross the ABI)
struct PluginCtx {
    const nvimgcodecFrameworkDesc_t* framework = nullptr;
    nvimgcodecDecoderDesc_t desc{};
};

nvimgcodecStatus_t dec_create(void* instance, nvimgcodecDecoder_t* decoder, const nvimgcodecExecutionParams_t* ep, const char* options)
{
    auto* ctx = static_cast<PluginCtx*>(instance);
    if (!ctx || !decoder || !ep)
        return NVIMGCODEC_STATUS_INVALID_PARAMETER;
    if (ep->device_id == NVIMGCODEC_DEVICE_CPU_ONLY)
        return NVIMGCODEC_STATUS_INVALID_PARAMETER; // a GPU-only decoder (same refusal as cuda_decoder.cpp:333-334)
    try {
        *decoder = reinterpret_cast<nvimgcodecDecoder_t>(new DecoderImpl(ctx->framework, ep, options));
    } catch (...) {
        return NVIMGCODEC_STATUS_EXTENSION_INTERNAL_ERROR;
    }
    return NVIMGCODEC_STATUS_SUCCESS;
}

nvimgcodecStatus_t dec_destroy(nvimgcodecDecoder_t decoder)
{
    delete reinterpret_cast<DecoderImpl*>(decoder);
    return NVIMGCODEC_STATUS_SUCCESS;
}

nvimgcodecStatus_t dec_get_metadata(nvimgcodecDecoder_t, const nvimgcodecCodeStreamDesc_t*, nvimgcodecMetadata_t**, int*)
{
    return NVIMGCODEC_STATUS_IMPLEMENTATION_UNSUPPORTED; // as libjpeg_turbo_decoder.cpp:105-108
}

nvimgcodecProcessingStatus_t dec_can_decode(nvimgcodecDecoder_t decoder, const nvimgcodecImageDesc_t* image,
    const nvimgcodecCodeStreamDesc_t* cs, const nvimgcodecDecodeParams_t* params, int thread_idx)
{
    try {
        return reinterpret_cast<DecoderImpl*>(decoder)->can_decode(image, cs, params, thread_idx);
    } catch (...) {
        return NVIMGCODEC_PROCESSING_STATUS_FAIL;
    }
}

nvimgcodecStatus_t dec_decode(nvimgcodecDecoder_t decoder, const nvimgcodecImageDesc_t* image, const nvimgcodecCodeStreamDesc_t* cs,
    const nvimgcodecDecodeParams_t* params, int)
{
    try {
        return reinterpret_cast<DecoderImpl*>(decoder)->decode_batch(&image, &cs, 1, false, params && params->apply_exif_orientation);
    } catch (...) {
        if (image)
            image->imageReady(image->instance, NVIMGCODEC_PROCESSING_STATUS_FAIL);
        return NVIMGCODEC_STATUS_EXTENSION_INTERNAL_ERROR;
    }
}

nvimgcodecStatus_t dec_decode_batch(nvimgcodecDecoder_t decoder, const nvimgcodecImageDesc_t** images,
    const nvimgcodecCodeStreamDesc_t** css, int batch_size, const nvimgcodecDecodeParams_t* params, int)
{
    try {
        return reinterpret_cast<DecoderImpl*>(decoder)->decode_batch(images, css, batch_size, true, params && params->apply_exif_orientation);
    } catch (...) {
        for (int i = 0; i < batch_size; ++i)
            images[i]->imageReady(images[i]->instance, NVIMGCODEC_PROCESSING_STATUS_FAIL);
        return NVIMGCODEC_STATUS_EXTENSION_INTERNAL_ERROR;
    }
}

nvimgcodecStatus_t dec_mini_batch(nvimgcodecDecoder_t, int* batch_size)
{
    if (batch_size)
        *batch_size = kMiniBatch;
    return NVIMGCODEC_STATUS_SUCCESS;
}

nvimgcodecStatus_t ext_create(void*, nvimgcodecExtension_t* extension, const nvimgcodecFrameworkDesc_t* framework)
{
    if (!extension || !framework)
        return NVIMGCODEC_STATUS_INVALID_PARAMETER;
    auto* ctx = new (std::nothrow) PluginCtx;
    if (!ctx)
        return NVIMGCODEC_STATUS_ALLOCATOR_FAILURE;
    ctx->framework = framework;
    ctx->desc = nvimgcodecDecoderDesc_t{NVIMGCODEC_STRUCTURE_TYPE_DECODER_DESC, sizeof(nvimgcodecDecoderDesc_t), nullptr, ctx, kDecoderId, "jpeg",
        NVIMGCODEC_BACKEND_KIND_GPU_ONLY, dec_create, dec_destroy, dec_get_metadata, dec_can_decode, dec_decode, dec_decode_batch,
        dec_mini_batch};
    const nvimgcodecStatus_t st = framework->registerDecoder(framework->instance, &ctx->desc, kPriority);
    if (st != NVIMGCODEC_STATUS_SUCCESS) {
        delete ctx;
        return st;
    }
    log_msg(framework, NVIMGCODEC_DEBUG_MESSAGE_SEVERITY_INFO, "b200 jpeg decoder registered");
    *extension = reinterpret_cast<nvimgcodecExtension_t>(ctx);
    return NVIMGCODEC_STATUS_SUCCESS;
}

nvimgcodecStatus_t ext_destroy(nvimgcodecExtension_t extension)
{
    auto* ctx = reinterpret_cast<PluginCtx*>(extension);
    if (!ctx)
        return NVIMGCODEC_STATUS_INVALID_PARAMETER;
    ctx->framework->unregisterDecoder(ctx->framework->instance, &ctx->desc);
    delete ctx;
    return NVIMGCODEC_STATUS_SUCCESS;
}

const nvimgcodecExtensionDesc_t kExtensionDesc = {NVIMGCODEC_STRUCTURE_TYPE_EXTENSION_DESC, sizeof(nvimgcodecExtensionDesc_t), nullptr, nullptr,
    kExtensionId, 100 /* 0.1.0 */, NVIMGCODEC_VER, ext_create, ext_destroy};

} // namespace

extern "C" {

// For in-process registration through nvimgcodecExtensionCreate, as the reference's own tests do
// (test/extensions/libjpeg_turbo_ext_decoder_test.cpp:49-52).
B200JPEG_API nvimgcodecStatus_t get_b200_jpeg_extension_desc(nvimgcodecExtensionDesc_t* ext_desc)
{
    if (!ext_desc)
        return NVIMGCODEC_STATUS_INVALID_PARAMETER;
    if (ext_desc->struct_type != NVIMGCODEC_STRUCTURE_TYPE_EXTENSION_DESC)
        return NVIMGCODEC_STATUS_INVALID_PARAMETER;
    *ext_desc = kExtensionDesc;
    return NVIMGCODEC_STATUS_SUCCESS;
}

// Found by dlsym when the framework scans its extensions directory (src/plugin_framework.cpp:338-372).
B200JPEG_API nvimgcodecStatus_t nvimgcodecExtensionModuleEntry(nvimgcodecExtensionDesc_t* ext_desc)
{
    return get_b200_jpeg_extension_desc(ext_desc);
}

} // extern "C"
