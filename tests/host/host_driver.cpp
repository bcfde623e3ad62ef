// host_driver -- drives the UNMODIFIED nvImageCodec core (built by oracle/build_ref.sh from the
// reference sources) through its public C API, with our extension discovered from a directory,
// exactly as nvimgcodec.Decoder / nvimtrans would (include/nvimgcodec.h:1733, 1973, 2030).
// Test infrastructure: reads a manifest, decodes the whole list as ONE batch, dumps raw pixels.
//
//   host_driver <extension dir> <manifest> <out dir> [decoder options] [repeat] [backend: any|gpu|hybrid|hw|cpu] [exif]
//   exif: decode with apply_exif_orientation = 1; images take the oriented size, as python/decoder.cpp does
//   repeat > 0: after the first (checked, dumped) decode, time `repeat` more decodes of the same batch
//   through nvimgcodecDecoderDecode + future wait + device sync and print images/s (used to compare
//   decoders -- ours vs the reference's nvjpeg_ext -- through the very same framework path).
//   manifest line: <jpeg path> <I_RGB|I_BGR|P_RGB|P_BGR|I_Y> <y0> <x0> <y1> <x1> <device|host>
//                  (all-zero ROI = whole image)
//   environment:
//     HOST_DRIVER_CRC=<k>        write the CRC-32 (zlib polynomial) of every k-th output into status.txt (0 for the others)
//                                instead of dumping the raw pixels (large batches)
//     HOST_DRIVER_COUNT_ALLOC=1  hand the decoder counting device / pinned allocators through nvimgcodecExecutionParams_t
//                                (include/nvimgcodec.h:1095-1096) and print how often each was called
//     HOST_DRIVER_PINNED=1       host output buffers are pinned (cudaMallocHost) instead of malloc'ed, allocated on the memory node
//                                with the faster device-to-host copies (probed: on a two-node box the far node halves the rate)
//     HOST_DRIVER_PINNED_INPUT=1 the bitstreams are handed to the framework in pinned host memory (what a loader that feeds a GPU decoder
//                                reads files into) instead of pageable std::vector storage
//     HOST_DRIVER_THREADS=<n>    nvimgcodecExecutionParams_t::max_num_cpu_threads for timing runs (default 0 = all cores)
//     HOST_DRIVER_FILL=<0..255>  constant sample for the parts of a region that lie outside the image (include/nvimgcodec.h:495-519)
#include <sched.h>

#include <cuda_runtime_api.h>
#include <nvimgcodec.h>

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>
#include <string>
#include <vector>

#define CHECK(x)                                                                  \
    do {                                                                          \
        auto _s = (x);                                                            \
        if (_s != NVIMGCODEC_STATUS_SUCCESS) {                                    \
            std::fprintf(stderr, "%s failed: %d (line %d)\n", #x, int(_s), __LINE__); \
            return 2;                                                             \
        }                                                                         \
    } while (0)

static uint32_t crc32_of(const unsigned char* p, size_t n)
{
    static uint32_t table[256];
    static bool init = false;
    if (!init) {
        for (uint32_t i = 0; i < 256; ++i) {
            uint32_t c = i;
            for (int k = 0; k < 8; ++k)
                c = (c & 1) ? 0xEDB88320u ^ (c >> 1) : c >> 1;
            table[i] = c;
        }
        init = true;
    }
    uint32_t c = 0xFFFFFFFFu;
    for (size_t i = 0; i < n; ++i)
        c = table[(c ^ p[i]) & 0xFF] ^ (c >> 8);
    return c ^ 0xFFFFFFFFu;
}

// counting allocators (signatures: include/nvimgcodec.h:246-281)
static long g_dev_malloc = 0, g_dev_free = 0, g_pin_malloc = 0, g_pin_free = 0;
static long long g_dev_bytes = 0, g_dev_freed = 0;
static int count_dev_malloc(void*, void** ptr, size_t size, cudaStream_t)
{
    ++g_dev_malloc;
    g_dev_bytes += (long long)size;
    return cudaMalloc(ptr, size) == cudaSuccess ? 0 : 1;
}
static int count_dev_free(void*, void* ptr, size_t size, cudaStream_t)
{
    ++g_dev_free;
    g_dev_freed += (long long)size;
    return cudaFree(ptr) == cudaSuccess ? 0 : 1;
}
static int count_pin_malloc(void*, void** ptr, size_t size, cudaStream_t)
{
    ++g_pin_malloc;
    return cudaMallocHost(ptr, size) == cudaSuccess ? 0 : 1;
}
static int count_pin_free(void*, void* ptr, size_t, cudaStream_t)
{
    ++g_pin_free;
    return cudaFreeHost(ptr) == cudaSuccess ? 0 : 1;
}

struct Item {
    std::string path, fmt, where;
    int y0, x0, y1, x1;
    std::vector<unsigned char> bytes;
    void* buffer = nullptr;
    size_t out_bytes = 0;
    int w = 0, h = 0;
};

int main(int argc, char** argv)
{
    if (argc < 4) {
        std::fprintf(stderr, "usage: host_driver <ext dir> <manifest> <out dir> [options]\n");
        return 1;
    }
    const std::string ext_dir = argv[1], manifest = argv[2], out_dir = argv[3];
    const char* options = argc > 4 ? argv[4] : "";
    const int repeat = argc > 5 ? std::atoi(argv[5]) : 0;
    const std::string backend = argc > 6 ? argv[6] : "any";
    const bool apply_exif = argc > 7 && std::string(argv[7]) == "exif";
    const bool crc_mode = std::getenv("HOST_DRIVER_CRC") != nullptr;
    const int crc_every = crc_mode ? std::max(1, std::atoi(std::getenv("HOST_DRIVER_CRC"))) : 1;
    const bool count_alloc = std::getenv("HOST_DRIVER_COUNT_ALLOC") != nullptr;
    const bool pinned_out = std::getenv("HOST_DRIVER_PINNED") != nullptr;
    std::vector<Item> items;
    {
        std::ifstream f(manifest);
        std::string line;
        while (std::getline(f, line)) {
            if (line.empty())
                continue;
            std::istringstream ss(line);
            Item it;
            ss >> it.path >> it.fmt >> it.y0 >> it.x0 >> it.y1 >> it.x1 >> it.where;
            std::ifstream j(it.path, std::ios::binary);
            it.bytes.assign(std::istreambuf_iterator<char>(j), std::istreambuf_iterator<char>());
            items.push_back(std::move(it));
        }
    }
    const int n = int(items.size());

    nvimgcodecInstance_t instance = nullptr;
    nvimgcodecInstanceCreateInfo_t ci{NVIMGCODEC_STRUCTURE_TYPE_INSTANCE_CREATE_INFO, sizeof(nvimgcodecInstanceCreateInfo_t), nullptr};
    ci.load_builtin_modules = 1;
    ci.load_extension_modules = 1;
    ci.extension_modules_path = ext_dir.c_str();
    ci.create_debug_messenger = 1;
    ci.message_severity = NVIMGCODEC_DEBUG_MESSAGE_SEVERITY_INFO | NVIMGCODEC_DEBUG_MESSAGE_SEVERITY_WARNING |
                          NVIMGCODEC_DEBUG_MESSAGE_SEVERITY_ERROR | NVIMGCODEC_DEBUG_MESSAGE_SEVERITY_FATAL;
    ci.message_category = NVIMGCODEC_DEBUG_MESSAGE_CATEGORY_ALL;
    CHECK(nvimgcodecInstanceCreate(&instance, &ci));

    nvimgcodecExecutionParams_t ep{NVIMGCODEC_STRUCTURE_TYPE_EXECUTION_PARAMS, sizeof(nvimgcodecExecutionParams_t), nullptr};
    ep.device_id = 0;
    ep.max_num_cpu_threads = repeat > 0 ? (std::getenv("HOST_DRIVER_THREADS") ? std::atoi(std::getenv("HOST_DRIVER_THREADS")) : 0) : 4; // timing runs: all host cores by default, like the default executor
    nvimgcodecBackend_t be{NVIMGCODEC_STRUCTURE_TYPE_BACKEND, sizeof(nvimgcodecBackend_t), nullptr};
    be.params = {NVIMGCODEC_STRUCTURE_TYPE_BACKEND_PARAMS, sizeof(nvimgcodecBackendParams_t), nullptr, 1.0f, NVIMGCODEC_LOAD_HINT_POLICY_FIXED};
    if (backend != "any") {
        be.kind = backend == "gpu"      ? NVIMGCODEC_BACKEND_KIND_GPU_ONLY
                  : backend == "hybrid" ? NVIMGCODEC_BACKEND_KIND_HYBRID_CPU_GPU
                  : backend == "hw"     ? NVIMGCODEC_BACKEND_KIND_HW_GPU_ONLY
                                        : NVIMGCODEC_BACKEND_KIND_CPU_ONLY;
        ep.num_backends = 1;
        ep.backends = &be;
    }
    nvimgcodecDeviceAllocator_t dev_alloc{NVIMGCODEC_STRUCTURE_TYPE_DEVICE_ALLOCATOR, sizeof(nvimgcodecDeviceAllocator_t), nullptr,
        count_dev_malloc, count_dev_free, nullptr, 0};
    nvimgcodecPinnedAllocator_t pin_alloc{NVIMGCODEC_STRUCTURE_TYPE_PINNED_ALLOCATOR, sizeof(nvimgcodecPinnedAllocator_t), nullptr,
        count_pin_malloc, count_pin_free, nullptr, 0};
    if (count_alloc) {
        ep.device_allocator = &dev_alloc;
        ep.pinned_allocator = &pin_alloc;
    }
    nvimgcodecDecoder_t decoder = nullptr;
    CHECK(nvimgcodecDecoderCreate(instance, &decoder, &ep, options));

    cudaStream_t stream;
    cudaStreamCreate(&stream);
    // pinned outputs: choose the CPU set (memory node) under which a pinned buffer copies fastest from the device
    cpu_set_t all_cpus, best_cpus;
    CPU_ZERO(&all_cpus);
    sched_getaffinity(0, sizeof(all_cpus), &all_cpus);
    best_cpus = all_cpus;
    if (pinned_out) {
        std::vector<int> cpus;
        for (int c = 0; c < CPU_SETSIZE; ++c)
            if (CPU_ISSET(c, &all_cpus))
                cpus.push_back(c);
        if (cpus.size() >= 4) {
            const size_t probe = size_t(256) << 20;
            void* dsrc = nullptr;
            cudaMalloc(&dsrc, probe);
            double best = 0;
            for (int half = 0; half < 2; ++half) {
                cpu_set_t cs;
                CPU_ZERO(&cs);
                for (size_t i = half ? cpus.size() / 2 : 0; i < (half ? cpus.size() : cpus.size() / 2); ++i)
                    CPU_SET(cpus[i], &cs);
                sched_setaffinity(0, sizeof(cs), &cs);
                void* h = nullptr;
                cudaMallocHost(&h, probe);
                cudaMemcpy(h, dsrc, probe, cudaMemcpyDeviceToHost);
                const auto t0 = std::chrono::steady_clock::now();
                for (int r = 0; r < 3; ++r)
                    cudaMemcpy(h, dsrc, probe, cudaMemcpyDeviceToHost);
                const double gbs = 3.0 * probe / std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() / 1e9;
                std::printf("PLACEMENT cpu half %d: D2H %.1f GB/s\n", half, gbs);
                if (gbs > best) {
                    best = gbs;
                    best_cpus = cs;
                }
                cudaFreeHost(h);
            }
            cudaFree(dsrc);
        }
        sched_setaffinity(0, sizeof(best_cpus), &best_cpus); // ... for the allocations below; the full set comes back after them
    }
    std::vector<nvimgcodecCodeStream_t> parents(n, nullptr), streams(n, nullptr);
    std::vector<nvimgcodecImage_t> images(n, nullptr);
    for (int i = 0; i < n; ++i) {
        Item& it = items[i];
        const unsigned char* src_bytes = it.bytes.data();
        if (std::getenv("HOST_DRIVER_PINNED_INPUT")) {
            void* pin = nullptr;
            if (cudaMallocHost(&pin, it.bytes.size()) == cudaSuccess) { // (lives until the process ends)
                std::memcpy(pin, it.bytes.data(), it.bytes.size());
                src_bytes = static_cast<const unsigned char*>(pin);
            }
        }
        CHECK(nvimgcodecCodeStreamCreateFromHostMem(instance, &parents[i], src_bytes, it.bytes.size(), nullptr));
        nvimgcodecImageInfo_t info{NVIMGCODEC_STRUCTURE_TYPE_IMAGE_INFO, sizeof(nvimgcodecImageInfo_t), nullptr};
        CHECK(nvimgcodecCodeStreamGetImageInfo(parents[i], &info));
        it.w = int(info.plane_info[0].width);
        it.h = int(info.plane_info[0].height);
        if (apply_exif && (info.orientation.rotated / 90) % 2)
            std::swap(it.w, it.h);
        streams[i] = parents[i];
        const bool roi = it.y1 > it.y0 && it.x1 > it.x0;
        if (roi) {
            nvimgcodecCodeStreamView_t view{NVIMGCODEC_STRUCTURE_TYPE_CODE_STREAM_VIEW, sizeof(nvimgcodecCodeStreamView_t), nullptr};
            view.image_idx = 0;
            view.region.struct_type = NVIMGCODEC_STRUCTURE_TYPE_REGION;
            view.region.struct_size = sizeof(nvimgcodecRegion_t);
            view.region.ndim = 2;
            view.region.start[0] = it.y0;
            view.region.start[1] = it.x0;
            view.region.end[0] = it.y1;
            view.region.end[1] = it.x1;
            if (const char* fv = std::getenv("HOST_DRIVER_FILL")) {
                view.region.out_of_bounds_policy = NVIMGCODEC_OUT_OF_BOUNDS_POLICY_CONSTANT;
                for (int k = 0; k < NVIMGCODEC_MAX_ROI_FILL_CHANNELS; ++k) {
                    view.region.out_of_bounds_samples[k].type = NVIMGCODEC_SAMPLE_DATA_TYPE_UINT8;
                    view.region.out_of_bounds_samples[k].value.as_uint = unsigned(std::atoi(fv));
                }
            }
            CHECK(nvimgcodecCodeStreamGetSubCodeStream(parents[i], &streams[i], &view));
            it.w = it.x1 - it.x0;
            it.h = it.y1 - it.y0;
        }
        const bool planar = it.fmt == "P_RGB" || it.fmt == "P_BGR";
        const bool gray = it.fmt == "I_Y";
        info.sample_format = it.fmt == "I_RGB" ? NVIMGCODEC_SAMPLEFORMAT_I_RGB
                             : it.fmt == "I_BGR" ? NVIMGCODEC_SAMPLEFORMAT_I_BGR
                             : it.fmt == "P_RGB" ? NVIMGCODEC_SAMPLEFORMAT_P_RGB
                             : it.fmt == "P_BGR" ? NVIMGCODEC_SAMPLEFORMAT_P_BGR
                                                 : NVIMGCODEC_SAMPLEFORMAT_I_Y;
        info.color_spec = gray ? NVIMGCODEC_COLORSPEC_GRAY : NVIMGCODEC_COLORSPEC_SRGB;
        info.chroma_subsampling = NVIMGCODEC_SAMPLING_NONE;
        info.num_planes = planar ? 3 : 1;
        const int channels = planar || gray ? 1 : 3;
        it.out_bytes = 0;
        for (uint32_t p = 0; p < info.num_planes; ++p) {
            info.plane_info[p].struct_type = NVIMGCODEC_STRUCTURE_TYPE_IMAGE_PLANE_INFO;
            info.plane_info[p].struct_size = sizeof(nvimgcodecImagePlaneInfo_t);
            info.plane_info[p].width = uint32_t(it.w);
            info.plane_info[p].height = uint32_t(it.h);
            info.plane_info[p].num_channels = uint32_t(channels);
            info.plane_info[p].row_stride = size_t(it.w) * channels;
            info.plane_info[p].sample_type = NVIMGCODEC_SAMPLE_DATA_TYPE_UINT8;
            info.plane_info[p].precision = 8;
            it.out_bytes += size_t(it.w) * channels * it.h;
        }
        if (it.where == "host") {
            if (pinned_out)
                cudaMallocHost(&it.buffer, it.out_bytes);
            else
                it.buffer = std::malloc(it.out_bytes);
            info.buffer_kind = NVIMGCODEC_IMAGE_BUFFER_KIND_STRIDED_HOST;
        } else {
            cudaMalloc(&it.buffer, it.out_bytes);
            info.buffer_kind = NVIMGCODEC_IMAGE_BUFFER_KIND_STRIDED_DEVICE;
        }
        info.buffer = it.buffer;
        info.cuda_stream = stream;
        CHECK(nvimgcodecImageCreate(instance, &images[i], &info));
    }

    sched_setaffinity(0, sizeof(all_cpus), &all_cpus);
    nvimgcodecDecodeParams_t dp{NVIMGCODEC_STRUCTURE_TYPE_DECODE_PARAMS, sizeof(nvimgcodecDecodeParams_t), nullptr};
    dp.apply_exif_orientation = apply_exif ? 1 : 0;
    nvimgcodecFuture_t future = nullptr;
    CHECK(nvimgcodecDecoderDecode(decoder, streams.data(), images.data(), n, &dp, &future));
    CHECK(nvimgcodecFutureWaitForAll(future));
    std::vector<nvimgcodecProcessingStatus_t> st(n, NVIMGCODEC_PROCESSING_STATUS_UNKNOWN);
    size_t st_n = 0;
    CHECK(nvimgcodecFutureGetProcessingStatus(future, st.data(), &st_n));
    cudaStreamSynchronize(stream);
    cudaDeviceSynchronize();

    if (repeat > 0) {
        int ok = 0;
        for (int i = 0; i < n; ++i)
            ok += st[i] == NVIMGCODEC_PROCESSING_STATUS_SUCCESS;
        cudaEvent_t e0, e1;
        cudaEventCreate(&e0);
        cudaEventCreate(&e1);
        auto run_once = [&]() -> int {
            nvimgcodecFuture_t f = nullptr;
            if (nvimgcodecDecoderDecode(decoder, streams.data(), images.data(), n, &dp, &f) != NVIMGCODEC_STATUS_SUCCESS)
                return 1;
            nvimgcodecFutureWaitForAll(f);
            nvimgcodecFutureDestroy(f);
            cudaStreamSynchronize(stream);
            return 0;
        };
        for (int w = 0; w < 2; ++w)
            run_once();
        cudaDeviceSynchronize();
        const auto t0 = std::chrono::steady_clock::now();
        for (int r = 0; r < repeat; ++r)
            if (run_once())
                return 3;
        cudaDeviceSynchronize();
        const double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        double mpix = 0;
        for (int i = 0; i < n; ++i)
            mpix += double(items[i].w) * items[i].h / 1e6;
        std::printf("BENCH images=%d ok=%d repeat=%d seconds=%.6f images_per_s=%.1f mpix_per_s=%.1f ms_per_batch=%.3f\n", n, ok, repeat, dt,
            n * repeat / dt, mpix * repeat / dt, dt / repeat * 1e3);
    }

    std::ofstream sf(out_dir + "/status.txt");
    for (int i = 0; i < n; ++i) {
        Item& it = items[i];
        sf << int(st[i]) << " " << it.w << " " << it.h;
        if (crc_mode && i % crc_every != 0) {
            sf << " 0\n";
            continue;
        }
        std::vector<unsigned char> host(it.out_bytes);
        if (it.where == "host")
            std::memcpy(host.data(), it.buffer, it.out_bytes);
        else
            cudaMemcpy(host.data(), it.buffer, it.out_bytes, cudaMemcpyDeviceToHost);
        if (crc_mode) {
            sf << " " << crc32_of(host.data(), host.size()) << "\n";
            continue;
        }
        sf << "\n";
        std::ofstream o(out_dir + "/" + std::to_string(i) + ".raw", std::ios::binary);
        o.write(reinterpret_cast<const char*>(host.data()), std::streamsize(host.size()));
    }
    nvimgcodecFutureDestroy(future);
    for (int i = 0; i < n; ++i) {
        nvimgcodecImageDestroy(images[i]);
        if (streams[i] != parents[i])
            nvimgcodecCodeStreamDestroy(streams[i]);
        nvimgcodecCodeStreamDestroy(parents[i]);
        if (items[i].where == "host") {
            if (pinned_out)
                cudaFreeHost(items[i].buffer);
            else
                std::free(items[i].buffer);
        } else
            cudaFree(items[i].buffer);
    }
    nvimgcodecDecoderDestroy(decoder);
    nvimgcodecInstanceDestroy(instance);
    if (count_alloc)
        std::printf("ALLOC device_malloc=%ld device_free=%ld device_bytes=%lld device_bytes_freed=%lld pinned_malloc=%ld pinned_free=%ld\n", g_dev_malloc,
            g_dev_free, g_dev_bytes, g_dev_freed, g_pin_malloc, g_pin_free);
    std::printf("decoded %d images through nvimgcodecDecoderDecode\n", n);
    return 0;
}
