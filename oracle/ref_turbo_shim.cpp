// ref_turbo_shim -- C entry points over the REFERENCE's own CPU JPEG path
// (libjpeg_turbo::Uncompress, extensions/libjpeg_turbo/jpeg_mem.cpp:475, compiled from where it
// lies by oracle/build_ref.sh against oracle/turbo_abi/jpeglib.h and linked to the libjpeg-turbo
// binary Pillow ships).  TEST INFRASTRUCTURE ONLY: it pins the oracle (both upsampling modes, ROI)
// and is the CPU baseline bench.py times.  The flag set-up below follows
// DecoderImpl::decode (extensions/libjpeg_turbo/libjpeg_turbo_decoder.cpp:327-372).
#include <atomic>
#include <chrono>
#include <cstring>
#include <thread>
#include <vector>

#include "jpeg_mem.h"

extern "C" {

// fmt: 0 = I_RGB, 1 = I_BGR, 4 = Y (grayscale), 5 = I_YUV (YCbCr samples, no colour transform).  crop: x, y, w, h (w == 0 -> whole image).
// out must hold w*h*components bytes.  Returns 0 on success.
__attribute__((visibility("default"))) int ref_turbo_decode(const unsigned char* data, int len, int fancy, int fmt, int crop_x, int crop_y,
    int crop_w, int crop_h, unsigned char* out, int* out_w, int* out_h)
{
    int w = 0, h = 0, c = 0;
    if (!libjpeg_turbo::GetImageInfo(data, len, &w, &h, &c))
        return 1;
    libjpeg_turbo::UncompressFlags flags;
    flags.components = (fmt == 4) ? 1 : 3;
    flags.sample_format = fmt == 0 ? NVIMGCODEC_SAMPLEFORMAT_I_RGB
                          : fmt == 1 ? NVIMGCODEC_SAMPLEFORMAT_I_BGR
                          : fmt == 5 ? NVIMGCODEC_SAMPLEFORMAT_I_YUV
                                     : NVIMGCODEC_SAMPLEFORMAT_I_Y;
    flags.dct_method = JDCT_DEFAULT;
    flags.fancy_upscaling = fancy != 0;
    if (crop_w > 0) {
        flags.crop = true;
        flags.crop_x = crop_x;
        flags.crop_y = crop_y;
        flags.crop_width = crop_w;
        flags.crop_height = crop_h;
        w = crop_w;
        h = crop_h;
    }
    auto px = libjpeg_turbo::Uncompress(data, len, flags);
    if (!px)
        return 2;
    if (out)
        std::memcpy(out, px.get(), size_t(w) * h * flags.components);
    if (out_w)
        *out_w = w;
    if (out_h)
        *out_h = h;
    return 0;
}

// Decodes every stream `passes` times on `threads` host threads (dynamic work queue); returns seconds.
__attribute__((visibility("default"))) double ref_turbo_time_batch(int n, const unsigned char* const* data, const int* len, int fancy, int fmt,
    int threads, int passes)
{
    std::atomic<int> next{0};
    std::atomic<int> failed{0};
    const int total = n * passes;
    auto worker = [&]() {
        for (;;) {
            const int j = next.fetch_add(1);
            if (j >= total)
                return;
            const int i = j % n;
            libjpeg_turbo::UncompressFlags flags;
            flags.components = (fmt == 4) ? 1 : 3;
            flags.sample_format = fmt == 0 ? NVIMGCODEC_SAMPLEFORMAT_I_RGB : (fmt == 1 ? NVIMGCODEC_SAMPLEFORMAT_I_BGR : NVIMGCODEC_SAMPLEFORMAT_I_Y);
            flags.fancy_upscaling = fancy != 0;
            auto px = libjpeg_turbo::Uncompress(data[i], len[i], flags);
            if (!px)
                failed.fetch_add(1);
        }
    };
    const auto t0 = std::chrono::steady_clock::now();
    std::vector<std::thread> pool;
    for (int t = 0; t < threads; ++t)
        pool.emplace_back(worker);
    for (auto& th : pool)
        th.join();
    const double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    return failed.load() ? -1.0 : dt;
}
}
