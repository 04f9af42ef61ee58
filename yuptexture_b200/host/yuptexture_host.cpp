// Host side of the drop-in API (include/yuptexture.h): the reference's three classes re-created over
// the C ABI of include/yupvq.h.  Reference files mirrored: src/vq_bcn_compressor.cpp (container,
// orchestration, zstd), src/vq_encoder.cpp:12-30, :248-258, :432-440, :614-622 (config, centroid
// encode), src/bcn_compressor.cpp (codec wrapper).  All VQ arithmetic happens on the GPU.
#include "../../include/yuptexture.h"

#include <dlfcn.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <mutex>
#include <stdexcept>
#include <string>

#include "../../include/yupcodec.h"
#include "../../include/yupvq.h"

namespace {

// ---- zstd through the system libzstd.so.1 (the reference vendors zstd as a submodule that is not in
// its tree; these are the 17 entry points src/vq_bcn_compressor.cpp uses, declared by hand) ----------
struct Zstd {
    void* h = nullptr;
    void* (*createCCtx)() = nullptr;
    void* (*createDCtx)() = nullptr;
    size_t (*freeCCtx)(void*) = nullptr;
    size_t (*freeDCtx)(void*) = nullptr;
    size_t (*compressBound)(size_t) = nullptr;
    size_t (*CCtx_setParameter)(void*, int, int) = nullptr;
    size_t (*compress2)(void*, void*, size_t, const void*, size_t) = nullptr;
    size_t (*compress_usingCDict)(void*, void*, size_t, const void*, size_t, const void*) = nullptr;
    unsigned (*isError)(size_t) = nullptr;
    const char* (*getErrorName)(size_t) = nullptr;
    void* (*createCDict)(const void*, size_t, int) = nullptr;
    void* (*createDDict)(const void*, size_t) = nullptr;
    size_t (*freeCDict)(void*) = nullptr;
    size_t (*freeDDict)(void*) = nullptr;
    unsigned long long (*getFrameContentSize)(const void*, size_t) = nullptr;
    size_t (*decompressDCtx)(void*, void*, size_t, const void*, size_t) = nullptr;
    size_t (*decompress_usingDDict)(void*, void*, size_t, const void*, size_t, const void*) = nullptr;
    // ZSTD_cParameter values (zstd.h, stable API): compressionLevel 100, enableLongDistanceMatching 160, nbWorkers 400
    enum { c_compressionLevel = 100, c_enableLongDistanceMatching = 160, c_nbWorkers = 400 };
    static constexpr unsigned long long CONTENTSIZE_UNKNOWN = 0ULL - 1, CONTENTSIZE_ERROR = 0ULL - 2;

    static Zstd& get() {
        static Zstd z;
        static std::once_flag once;
        std::call_once(once, [] {
            for (const char* n : {"libzstd.so.1", "libzstd.so"}) {
                z.h = dlopen(n, RTLD_NOW);
                if (z.h) break;
            }
            if (!z.h) return;
            auto s = [&](const char* n) { return dlsym(z.h, n); };
#define YT_SYM(field, name) z.field = reinterpret_cast<decltype(z.field)>(s(name))
            YT_SYM(createCCtx, "ZSTD_createCCtx"); YT_SYM(createDCtx, "ZSTD_createDCtx");
            YT_SYM(freeCCtx, "ZSTD_freeCCtx"); YT_SYM(freeDCtx, "ZSTD_freeDCtx");
            YT_SYM(compressBound, "ZSTD_compressBound"); YT_SYM(CCtx_setParameter, "ZSTD_CCtx_setParameter");
            YT_SYM(compress2, "ZSTD_compress2"); YT_SYM(compress_usingCDict, "ZSTD_compress_usingCDict");
            YT_SYM(isError, "ZSTD_isError"); YT_SYM(getErrorName, "ZSTD_getErrorName");
            YT_SYM(createCDict, "ZSTD_createCDict"); YT_SYM(createDDict, "ZSTD_createDDict");
            YT_SYM(freeCDict, "ZSTD_freeCDict"); YT_SYM(freeDDict, "ZSTD_freeDDict");
            YT_SYM(getFrameContentSize, "ZSTD_getFrameContentSize"); YT_SYM(decompressDCtx, "ZSTD_decompressDCtx");
            YT_SYM(decompress_usingDDict, "ZSTD_decompress_usingDDict");
#undef YT_SYM
        });
        if (!z.h || !z.createCCtx || !z.compress2 || !z.decompressDCtx)
            throw std::runtime_error("libzstd.so.1 not available (needed unless params.useZstd is false)");
        return z;
    }
};

// ---- small built-in BCn encoders (stand-ins for the Compressonator encoders the integrator links) -----
inline uint16_t pack565(const int* c) {
    return (uint16_t)((((c[0] * 31 + 127) / 255) << 11) | (((c[1] * 63 + 127) / 255) << 5) | ((c[2] * 31 + 127) / 255));
}
inline void unpack565(uint16_t v, int* c) {
    const int r = (v >> 11) & 31, g = (v >> 5) & 63, b = v & 31;
    c[0] = (r << 3) | (r >> 2);
    c[1] = (g << 2) | (g >> 4);
    c[2] = (b << 3) | (b >> 2);
}

// px: 16 texels x 4 (RGBA).  force_four: BC3 colour blocks are always 4-colour; punch: BC1 with alpha
void encode_bc1_colour(const uint8_t* px, bool force_four, bool punch, uint8_t alpha_threshold, uint8_t* out) {
    int hi[3] = {0, 0, 0}, lo[3] = {255, 255, 255};
    bool any_transparent = false;
    for (int p = 0; p < 16; ++p) {
        if (punch && px[p * 4 + 3] < alpha_threshold) { any_transparent = true; continue; }
        for (int c = 0; c < 3; ++c) {
            hi[c] = std::max(hi[c], (int)px[p * 4 + c]);
            lo[c] = std::min(lo[c], (int)px[p * 4 + c]);
        }
    }
    if (lo[0] > hi[0]) { for (int c = 0; c < 3; ++c) lo[c] = hi[c] = 0; }
    uint16_t c0 = pack565(hi), c1 = pack565(lo);
    if (c0 < c1) std::swap(c0, c1);
    bool four = force_four || (c0 > c1 && !any_transparent);
    if (!four && c0 > c1) std::swap(c0, c1);  // c0 <= c1 selects the 3-colour + transparent mode
    int e0[3], e1[3], pal[4][3];
    unpack565(c0, e0);
    unpack565(c1, e1);
    for (int c = 0; c < 3; ++c) {
        pal[0][c] = e0[c];
        pal[1][c] = e1[c];
        pal[2][c] = four ? (2 * e0[c] + e1[c]) / 3 : (e0[c] + e1[c]) / 2;
        pal[3][c] = four ? (e0[c] + 2 * e1[c]) / 3 : 0;
    }
    uint32_t bits = 0;
    for (int p = 0; p < 16; ++p) {
        int best = 0, bd = 1 << 30;
        if (!four && punch && px[p * 4 + 3] < alpha_threshold) {
            best = 3;
        } else {
            for (int k = 0; k < (four ? 4 : 3); ++k) {
                int d = 0;
                for (int c = 0; c < 3; ++c) d += (px[p * 4 + c] - pal[k][c]) * (px[p * 4 + c] - pal[k][c]);
                if (d < bd) { bd = d; best = k; }
            }
        }
        bits |= (uint32_t)best << (2 * p);
    }
    out[0] = c0 & 255; out[1] = c0 >> 8; out[2] = c1 & 255; out[3] = c1 >> 8;
    std::memcpy(out + 4, &bits, 4);
}

void encode_bc4(const uint8_t* v, int stride, uint8_t* out) {  // 16 values, 8-value mode (or flat)
    int a0 = 0, a1 = 255;
    for (int p = 0; p < 16; ++p) { a0 = std::max(a0, (int)v[p * stride]); a1 = std::min(a1, (int)v[p * stride]); }
    int pal[8] = {a0, a1};
    for (int i = 2; i < 8; ++i)
        pal[i] = a0 > a1 ? ((8 - i) * a0 + (i - 1) * a1 + 3) / 7 : (i < 6 ? ((6 - i) * a0 + (i - 1) * a1 + 2) / 5 : (i == 6 ? 0 : 255));
    uint64_t bits = 0;
    for (int p = 0; p < 16; ++p) {
        int best = 0, bd = 1 << 30;
        for (int k = 0; k < 8; ++k) {
            const int d = std::abs((int)v[p * stride] - pal[k]);
            if (d < bd) { bd = d; best = k; }
        }
        bits |= (uint64_t)best << (3 * p);
    }
    out[0] = (uint8_t)a0; out[1] = (uint8_t)a1;
    for (int i = 0; i < 6; ++i) out[2 + i] = (uint8_t)(bits >> (8 * i));
}

void encode_bc7_mode6(const uint8_t* px, uint8_t* out) {  // one subset, 7777.1 endpoints, 4-bit indices
    static const int w[16] = {0, 4, 9, 13, 17, 21, 26, 30, 34, 38, 43, 47, 51, 55, 60, 64};
    int e[2][4];
    for (int c = 0; c < 4; ++c) {
        e[0][c] = 255; e[1][c] = 0;
        for (int p = 0; p < 16; ++p) { e[0][c] = std::min(e[0][c], (int)px[p * 4 + c]); e[1][c] = std::max(e[1][c], (int)px[p * 4 + c]); }
    }
    int pb[2] = {e[0][0] & 1, e[1][0] & 1}, q[2][4], full[2][4];
    for (int s = 0; s < 2; ++s)
        for (int c = 0; c < 4; ++c) {
            q[s][c] = std::min(127, std::max(0, (e[s][c] - pb[s] + 1) >> 1));
            full[s][c] = (q[s][c] << 1) | pb[s];
        }
    int idx[16];
    for (int p = 0; p < 16; ++p) {
        int best = 0, bd = 1 << 30;
        for (int k = 0; k < 16; ++k) {
            int d = 0;
            for (int c = 0; c < 4; ++c) {
                const int v = (full[0][c] * (64 - w[k]) + full[1][c] * w[k] + 32) >> 6;
                d += (px[p * 4 + c] - v) * (px[p * 4 + c] - v);
            }
            if (d < bd) { bd = d; best = k; }
        }
        idx[p] = best;
    }
    if (idx[0] >= 8) {  // the anchor texel's index MSB is implicit 0: swap the endpoints
        for (int p = 0; p < 16; ++p) idx[p] = 15 - idx[p];
        for (int c = 0; c < 4; ++c) std::swap(q[0][c], q[1][c]);
        std::swap(pb[0], pb[1]);
    }
    unsigned __int128 bits = (unsigned __int128)1 << 6;
    int pos = 7;
    auto put = [&](unsigned v, int n) { bits |= (unsigned __int128)v << pos; pos += n; };
    for (int c = 0; c < 4; ++c) { put(q[0][c], 7); put(q[1][c], 7); }
    put(pb[0], 1); put(pb[1], 1);
    put(idx[0], 3);
    for (int p = 1; p < 16; ++p) put(idx[p], 4);
    std::memcpy(out, &bits, 16);
}

BCnCompressor::Codec g_codec;
std::mutex g_codec_mutex;
std::string g_codec_name = "built-in";

double now_ms() {
    return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

// ---- block-codec plugin (include/yupcodec.h) -------------------------------------------------------------------
bool install_plugin(const char* path) {
    void* h = dlopen(path, RTLD_NOW | RTLD_LOCAL);
    if (!h) return false;
    auto enc = reinterpret_cast<decltype(&yupcodec_compress)>(dlsym(h, "yupcodec_compress"));
    auto enc_hdr = reinterpret_cast<decltype(&yupcodec_compress_hdr)>(dlsym(h, "yupcodec_compress_hdr"));
    auto dec_hdr = reinterpret_cast<decltype(&yupcodec_decompress_hdr)>(dlsym(h, "yupcodec_decompress_hdr"));
    auto name = reinterpret_cast<decltype(&yupcodec_name)>(dlsym(h, "yupcodec_name"));
    if (!enc || !enc_hdr || !dec_hdr) {
        dlclose(h);
        return false;
    }
    BCnCompressor::Codec c;
    c.compress = [enc](const uint8_t* img, uint32_t w, uint32_t hgt, uint32_t ch, BCFormat f, int threads, float q, uint8_t at, bool flip) {
        std::vector<uint8_t> out((size_t)((w + 3) / 4) * ((hgt + 3) / 4) * BCBlockSize::GetSize(f));
        if (enc(img, w, hgt, ch, (int)f, threads, q, at, flip ? 1 : 0, out.data(), out.size())) out.clear();  // empty = failure (:111)
        return out;
    };
    c.compressHDR = [enc_hdr](const float* img, uint32_t w, uint32_t hgt, uint32_t ch, BCFormat f, int threads, float q) {
        std::vector<uint8_t> out((size_t)((w + 3) / 4) * ((hgt + 3) / 4) * BCBlockSize::GetSize(f));
        if (enc_hdr(img, w, hgt, ch, (int)f, threads, q, out.data(), out.size())) out.clear();
        return out;
    };
    c.decompressHDR = [dec_hdr](const uint8_t* bc, uint32_t w, uint32_t hgt, uint32_t ch, BCFormat f) {
        std::vector<float> out((size_t)w * hgt * ch);
        if (dec_hdr(bc, w, hgt, ch, (int)f, out.data())) out.clear();
        return out;
    };
    std::lock_guard<std::mutex> lock(g_codec_mutex);
    g_codec = c;
    g_codec_name = name ? name() : path;
    return true;   // the handle stays open for the life of the process
}

// first use: $YUPTEXTURE_CODEC, else libyupcodec_cmp.so next to this library
void autoload_codec() {
    static std::once_flag once;
    std::call_once(once, [] {
        {
            std::lock_guard<std::mutex> lock(g_codec_mutex);
            if (g_codec.compress || g_codec.compressHDR || g_codec.decompressHDR) return;  // SetCodec was called first
        }
        if (const char* e = std::getenv("YUPTEXTURE_CODEC")) {
            if (*e && install_plugin(e)) return;
            if (!*e) return;   // YUPTEXTURE_CODEC= (empty) keeps the built-in encoders
        }
        Dl_info info;
        if (dladdr(reinterpret_cast<void*>(&autoload_codec), &info) && info.dli_fname) {
            std::string dir(info.dli_fname);
            const size_t slash = dir.find_last_of('/');
            dir = slash == std::string::npos ? std::string(".") : dir.substr(0, slash);
            install_plugin((dir + "/libyupcodec_cmp.so").c_str());
        }
    });
}

BCnCompressor::Codec current_codec() {
    autoload_codec();
    std::lock_guard<std::mutex> lock(g_codec_mutex);
    return g_codec;
}

// fetch a 4x4 RGBA tile (edge clamped), expanding 1/2/3-channel sources the way the VQ path reads them
void fetch_tile(const uint8_t* img, uint32_t w, uint32_t h, uint32_t ch, uint32_t bx, uint32_t by, bool flip, uint8_t* px) {
    for (uint32_t y = 0; y < 4; ++y)
        for (uint32_t x = 0; x < 4; ++x) {
            const uint8_t* s = img + ((size_t)std::min(by * 4 + y, h - 1) * w + std::min(bx * 4 + x, w - 1)) * ch;
            uint8_t* d = px + (y * 4 + x) * 4;
            if (ch >= 3) { d[0] = s[flip ? 2 : 0]; d[1] = s[1]; d[2] = s[flip ? 0 : 2]; d[3] = ch == 4 ? s[3] : 255; }
            else { d[0] = d[1] = d[2] = s[0]; d[3] = ch == 2 ? s[1] : 255; }
        }
}

}  // namespace

// ---- BCnCompressor ---------------------------------------------------------------------------------------
void BCnCompressor::SetCodec(const Codec& codec) {
    std::lock_guard<std::mutex> lock(g_codec_mutex);
    g_codec = codec;
    g_codec_name = (codec.compress || codec.compressHDR || codec.decompressHDR) ? "user (SetCodec)" : "built-in";
}

bool BCnCompressor::LoadCodecLibrary(const char* path) { return path && install_plugin(path); }

const char* BCnCompressor::CodecName() {
    autoload_codec();
    std::lock_guard<std::mutex> lock(g_codec_mutex);
    static thread_local std::string name;
    name = g_codec_name;
    return name.c_str();
}

namespace {
std::atomic<int> g_gpu_encode{BCnCompressor::kGpuAuto};
thread_local const char* t_last_encoder = "none";

bool gpu_encode_wanted(bool have_codec) {
    int mode = g_gpu_encode.load(std::memory_order_relaxed);
    if (mode == BCnCompressor::kGpuAuto) {
        static const int env_mode = [] {
            const char* e = std::getenv("YUPTEXTURE_GPU_BCN");
            return e ? (std::atoi(e) != 0 ? (int)BCnCompressor::kGpuOn : (int)BCnCompressor::kGpuOff) : (int)BCnCompressor::kGpuAuto;
        }();
        mode = env_mode;
    }
    return mode == BCnCompressor::kGpuOn || (mode == BCnCompressor::kGpuAuto && !have_codec);
}

yupvq_ctx* codec_ctx();
extern std::mutex g_codec_ctx_mutex;
}  // namespace

void BCnCompressor::SetGpuEncode(GpuEncode mode) { g_gpu_encode.store((int)mode, std::memory_order_relaxed); }
const char* BCnCompressor::LastEncoderName() { return t_last_encoder; }

std::vector<uint8_t> BCnCompressor::Compress(const uint8_t* inData, uint32_t width, uint32_t height, uint32_t channelCount,
                                             BCFormat format, int numThreads, float quality, uint8_t alphaThreshold, bool flipRGB) {
    const Codec c = current_codec();
    // N4: the pre-pass on the GPU.  Compressonator reads 3-channel sources red/blue-swapped (the reference compensates
    // only on its VQ-bypass path, flipRGB == true: src/bcn_compressor.cpp:61-76), so the same view is encoded here.
    if (inData && width && height && channelCount >= 1 && channelCount <= 4 && yupvq_bcn_encode_supported((int)format) &&
        gpu_encode_wanted((bool)c.compress)) {
        const size_t bytes = (size_t)((width + 3) / 4) * ((height + 3) / 4) * BCBlockSize::GetSize(format);
        std::vector<uint8_t> out(bytes);
        std::lock_guard<std::mutex> lock(g_codec_ctx_mutex);
        if (yupvq_bcn_encode_image(codec_ctx(), inData, width, height, (int)channelCount, (int)format, channelCount == 3 && !flipRGB ? 1 : 0,
                                   alphaThreshold, out.data(), nullptr) == 0) {
            t_last_encoder = "gpu (yupvq_bcn_encode_image)";
            return out;
        }
        return {};   // the reference returns an empty vector on failure (src/bcn_compressor.cpp:111)
    }
    if (c.compress) {
        t_last_encoder = "codec";
        return c.compress(inData, width, height, channelCount, format, numThreads, quality, alphaThreshold, flipRGB);
    }
    t_last_encoder = "built-in (cpu)";
    {
        // Not silent: these stand-in encoders (bounding-box BC1/3/4/5, mode-6-only BC7) are below Compressonator's quality.
        // One line per process the first time they write blocks; YUPTEXTURE_QUIET=1 turns it off.
        static std::once_flag warned;
        std::call_once(warned, [format] {
            if (getenv("YUPTEXTURE_QUIET")) return;
            fprintf(stderr,
                    "yuptexture: no BCn codec installed -- %s blocks come from the built-in stand-in encoder (lower quality than the "
                    "reference's Compressonator). Put libyupcodec_cmp.so next to libyuptexture.so, set YUPTEXTURE_CODEC, or call "
                    "BCnCompressor::SetCodec().\n",
                    format == BCFormat::BC7 ? "BC7 (mode 6 only)" : "BCn");
        });
    }
    if (!inData || width == 0 || height == 0 || channelCount < 1 || channelCount > 4) return {};
    if (format != BCFormat::BC1 && format != BCFormat::BC3 && format != BCFormat::BC4 && format != BCFormat::BC5 && format != BCFormat::BC7)
        return {};  // the reference returns an empty vector on failure (src/bcn_compressor.cpp:111)
    const uint32_t bx = (width + 3) / 4, by = (height + 3) / 4;
    const size_t bs = BCBlockSize::GetSize(format);
    std::vector<uint8_t> out((size_t)bx * by * bs);
#pragma omp parallel for num_threads(numThreads > 0 ? numThreads : 1)
    for (int64_t i = 0; i < (int64_t)bx * by; ++i) {
        uint8_t px[64];
        fetch_tile(inData, width, height, channelCount, (uint32_t)(i % bx), (uint32_t)(i / bx), flipRGB, px);
        uint8_t* o = out.data() + (size_t)i * bs;
        switch (format) {
            case BCFormat::BC1: encode_bc1_colour(px, false, channelCount == 4, alphaThreshold, o); break;
            case BCFormat::BC3: encode_bc4(px + 3, 4, o); encode_bc1_colour(px, true, false, 0, o + 8); break;
            case BCFormat::BC4: encode_bc4(px, 4, o); break;
            case BCFormat::BC5: encode_bc4(px, 4, o); encode_bc4(px + (channelCount == 2 ? 3 : 1), 4, o + 8); break;
            default: encode_bc7_mode6(px, o); break;
        }
    }
    return out;
}

std::vector<uint8_t> BCnCompressor::CompressHDR(const float* inData, uint32_t width, uint32_t height, uint32_t channelCount,
                                                BCFormat format, int numThreads, float quality) {
    if (const Codec c = current_codec(); c.compressHDR) return c.compressHDR(inData, width, height, channelCount, format, numThreads, quality);
    return {};  // no built-in BC6H encoder: a codec plugin (include/yupcodec.h) or SetCodec supplies one
}

std::vector<float> BCnCompressor::DecompressHDR(const uint8_t* bcData, uint32_t width, uint32_t height, uint32_t channelCount,
                                                BCFormat format) {
    if (const Codec c = current_codec(); c.decompressHDR) return c.decompressHDR(bcData, width, height, channelCount, format);
    return {};
}

// ---- device context helpers -----------------------------------------------------------------------------------
namespace {

int default_device() {
    const char* e = std::getenv("YUPVQ_DEVICE");
    return e ? std::atoi(e) : 0;
}

uint64_t default_seed() {
    const char* e = std::getenv("YUPVQ_SEED");
    return e ? std::strtoull(e, nullptr, 10) : 0;
}

yupvq_ctx* create_ctx(int device) {
    yupvq_ctx* c = nullptr;
    if (yupvq_create(device, &c)) throw std::runtime_error(std::string("yupvq: ") + yupvq_last_error(nullptr));
    return c;
}

[[noreturn]] void fail(yupvq_ctx* c, const char* what) {
    throw std::runtime_error(std::string(what) + ": " + yupvq_last_error(c));
}

// one process-wide context for the stateless BCnCompressor::Compress (GPU pre-pass) / Decompress
yupvq_ctx* codec_ctx() {
    static yupvq_ctx* c = create_ctx(default_device());
    return c;
}
std::mutex g_codec_ctx_mutex;

yupvq_config make_cfg(const CompressionConfig& c, uint64_t seed) {
    yupvq_config cfg;
    yupvq_config_init(&cfg, c.quality, c.maxIterations,
                      c.metric == DistanceMetric::PERCEPTUAL_OKLAB ? YUPVQ_METRIC_OKLAB : YUPVQ_METRIC_SAD,
                      c.fastModeSampleRatio, seed);
    cfg.codebook_size = c.codebookSize;
    return cfg;
}

template <typename T>
std::vector<T> flatten(const std::vector<std::vector<T>>& blocks, size_t per) {
    std::vector<T> flat(blocks.size() * per);
    for (size_t i = 0; i < blocks.size(); ++i) {
        if (blocks[i].size() < per) throw std::runtime_error("block has fewer than 16 * channels elements");
        std::memcpy(flat.data() + i * per, blocks[i].data(), per * sizeof(T));
    }
    return flat;
}

template <typename T>
std::vector<std::vector<T>> unflatten(const std::vector<T>& flat, size_t per) {
    std::vector<std::vector<T>> out(flat.size() / per);
    for (size_t i = 0; i < out.size(); ++i) out[i].assign(flat.begin() + i * per, flat.begin() + (i + 1) * per);
    return out;
}

}  // namespace

// BCnCompressor::Decompress, reference src/bcn_compressor.cpp:191-227: whole texture (or one 4x4 block) to
// width x height x channelCount bytes.  Blocks are decoded on the GPU and de-tiled here.
std::vector<uint8_t> BCnCompressor::Decompress(const uint8_t* bcData, uint32_t width, uint32_t height, uint32_t channelCount,
                                               BCFormat format) {
    if (!bcData || width == 0 || height == 0 || channelCount < 1 || channelCount > 4) return {};
    const uint32_t bx = (width + 3) / 4, by = (height + 3) / 4;
    const size_t n = (size_t)bx * by, per = 16 * channelCount;
    std::vector<uint8_t> blocks(n * per);
    {
        std::lock_guard<std::mutex> lock(g_codec_ctx_mutex);
        if (yupvq_decode_bcn(codec_ctx(), bcData, n, (int)format, (int)channelCount, blocks.data())) return {};
    }
    std::vector<uint8_t> img((size_t)width * height * channelCount);
    for (uint32_t y = 0; y < height; ++y)
        for (uint32_t x = 0; x < width; ++x)
            std::memcpy(&img[((size_t)y * width + x) * channelCount],
                        &blocks[((size_t)(y / 4) * bx + x / 4) * per + ((y & 3) * 4 + (x & 3)) * channelCount], channelCount);
    return img;
}

// ---- CompressionConfig (reference src/vq_encoder.cpp:8-18) ---------------------------------------------------
CompressionConfig::CompressionConfig(float quality_level) { SetQuality(quality_level); }

void CompressionConfig::SetQuality(float quality_level) {
    quality = std::max(0.0f, std::min(1.0f, quality_level));
    const uint32_t power = min_cb_power + (uint32_t)std::round(quality * (float)(max_cb_power - min_cb_power));
    codebookSize = 1u << power;
}

// ---- VQEncoder ---------------------------------------------------------------------------------------------------
VQEncoder::VQEncoder(const CompressionConfig& cfg) : config(cfg), seed(default_seed()), device(default_device()) {}

VQEncoder::~VQEncoder() {
    if (ctx && ownsCtx) yupvq_destroy(ctx);
}

void VQEncoder::AttachContext(yupvq_ctx* shared) {
    if (ctx && ownsCtx) yupvq_destroy(ctx);
    ctx = shared;
    ownsCtx = false;
}

void VQEncoder::SetFormat(BCFormat format) { bcFormat = format; }
void VQEncoder::SetSeed(uint64_t s) { seed = s; }

void VQEncoder::SetDevice(int d) {
    if (ctx) throw std::runtime_error("VQEncoder::SetDevice must be called before the first use");
    device = d;
}

yupvq_ctx* VQEncoder::Context() {
    if (!ctx) ctx = create_ctx(device);
    return ctx;
}

// K calls of the 4x4 block encoder, reference src/vq_encoder.cpp:432-440 (CompressSingleBlock :248-250)
VQCodebook VQEncoder::EncodeCentroids(const std::vector<std::vector<uint8_t>>& pixelCentroids, uint8_t channels,
                                      const CompressionParams& params) {
    const double t0 = now_ms();
    VQCodebook cb((uint32_t)BCBlockSize::GetSize(bcFormat), config.codebookSize);
    cb.entries.resize(pixelCentroids.size());
    // GPU encoders active (BCnCompressor::SetGpuEncode): the K blocks stacked into one 4 x 4K image are one kernel launch
    // instead of K single-block calls -- the same encoder, the same bytes as K calls of BCnCompressor::Compress(4, 4).
    if (!pixelCentroids.empty() && yupvq_bcn_encode_supported((int)bcFormat) && gpu_encode_wanted((bool)current_codec().compress)) {
        const size_t per = 16 * (size_t)channels, bs = BCBlockSize::GetSize(bcFormat), K = pixelCentroids.size();
        const std::vector<uint8_t> flat = flatten(pixelCentroids, per);
        std::vector<uint8_t> out(K * bs);
        int rc;
        {
            std::lock_guard<std::mutex> lock(g_codec_ctx_mutex);
            rc = yupvq_bcn_encode_image(codec_ctx(), flat.data(), 4, (uint32_t)(4 * K), (int)channels, (int)bcFormat, channels == 3 ? 1 : 0, 128,
                                        out.data(), nullptr);
        }
        if (rc) throw std::runtime_error("GPU centroid encode failed");
        for (size_t i = 0; i < K; ++i) cb.entries[i].assign(out.begin() + i * bs, out.begin() + (i + 1) * bs);
        lastCentroidMs = now_ms() - t0;
        return cb;
    }
    // CompressSingleBlock passes (numThreads, bcQuality) only, so the default alphaThreshold (128) applies, not params'.
    // One thread per 4x4 call (the result of a single-block encode does not depend on it), the K calls
    // spread over params.numThreads as in the reference's loop.
#pragma omp parallel for num_threads(params.numThreads > 0 ? params.numThreads : 1) schedule(dynamic, 8)
    for (int64_t i = 0; i < (int64_t)pixelCentroids.size(); ++i)
        cb.entries[i] = bcnCompressor.Compress(pixelCentroids[i].data(), 4, 4, channels, bcFormat, 1, params.bcQuality);
    lastCentroidMs = now_ms() - t0;
    return cb;
}

VQCodebook VQEncoder::EncodeCentroidsHDR(const std::vector<std::vector<float>>& centroids, uint8_t channels,
                                         const CompressionParams& params) {
    const double t0 = now_ms();
    VQCodebook cb((uint32_t)BCBlockSize::GetSize(bcFormat), config.codebookSize);
    cb.entries.resize(centroids.size());
#pragma omp parallel for num_threads(params.numThreads > 0 ? params.numThreads : 1) schedule(dynamic, 8)
    for (int64_t i = 0; i < (int64_t)centroids.size(); ++i)
        cb.entries[i] = bcnCompressor.CompressHDR(centroids[i].data(), 4, 4, channels, bcFormat, 1, params.bcQuality);
    for (const auto& e : cb.entries)
        if (e.empty())
            throw std::runtime_error("no HDR block encoder installed (codec plugin or BCnCompressor::SetCodec): cannot build BC6H codebook entries");
    lastCentroidMs = now_ms() - t0;
    return cb;
}

VQCodebook VQEncoder::BuildCodebook(const std::vector<std::vector<uint8_t>>& allPixelBlocks, uint8_t channels,
                                    std::vector<std::vector<uint8_t>>& outPixelCentroids, const CompressionParams& params) {
    const size_t per = 16 * (size_t)channels;
    const std::vector<uint8_t> flat = flatten(allPixelBlocks, per);
    const yupvq_config cfg = make_cfg(config, seed);
    std::vector<uint8_t> cent((size_t)cfg.codebook_size * per);
    const double t_vq = now_ms();
    if (yupvq_build_codebook(Context(), flat.data(), allPixelBlocks.size(), channels, &cfg, cent.data(), nullptr))
        fail(ctx, "BuildCodebook");
    lastVqMs = now_ms() - t_vq;
    outPixelCentroids = unflatten(cent, per);
    return EncodeCentroids(outPixelCentroids, channels, params);
}

std::vector<uint32_t> VQEncoder::QuantizeBlocks(const std::vector<std::vector<uint8_t>>& pixelBlocks, uint8_t channels,
                                                const std::vector<std::vector<uint8_t>>& pixelCentroids, const CompressionParams&) {
    if (pixelBlocks.empty()) return {};
    const size_t per = 16 * (size_t)channels;
    const std::vector<uint8_t> flat = flatten(pixelBlocks, per), cent = flatten(pixelCentroids, per);
    std::vector<uint32_t> idx(pixelBlocks.size());
    if (yupvq_quantize_blocks(Context(), flat.data(), pixelBlocks.size(), channels, cent.data(), (uint32_t)pixelCentroids.size(),
                              config.metric == DistanceMetric::PERCEPTUAL_OKLAB ? YUPVQ_METRIC_OKLAB : YUPVQ_METRIC_SAD,
                              YUPVQ_ASSIGN_AUTO, idx.data(), nullptr))
        fail(ctx, "QuantizeBlocks");
    return idx;
}

VQCodebook VQEncoder::BuildCodebookHDR(const std::vector<std::vector<float>>& allPixelFloatBlocks, uint8_t channels,
                                       std::vector<std::vector<float>>& outPixelFloatCentroids, const CompressionParams& params) {
    if (allPixelFloatBlocks.empty()) return {};
    const size_t per = 16 * (size_t)channels;
    const std::vector<float> flat = flatten(allPixelFloatBlocks, per);
    const yupvq_config cfg = make_cfg(config, seed);
    std::vector<float> cent((size_t)cfg.codebook_size * per);
    const double t_vq = now_ms();
    if (yupvq_build_codebook_hdr(Context(), flat.data(), allPixelFloatBlocks.size(), channels, &cfg, cent.data(), nullptr))
        fail(ctx, "BuildCodebookHDR");
    lastVqMs = now_ms() - t_vq;
    outPixelFloatCentroids = unflatten(cent, per);
    return EncodeCentroidsHDR(outPixelFloatCentroids, channels, params);
}

std::vector<uint32_t> VQEncoder::QuantizeBlocksHDR(const std::vector<std::vector<float>>& pixelFloatBlocks, uint8_t channels,
                                                   const std::vector<std::vector<float>>& pixelFloatCentroids, const CompressionParams&) {
    if (pixelFloatBlocks.empty()) return {};
    const size_t per = 16 * (size_t)channels;
    const std::vector<float> flat = flatten(pixelFloatBlocks, per), cent = flatten(pixelFloatCentroids, per);
    std::vector<uint32_t> idx(pixelFloatBlocks.size());
    if (yupvq_quantize_blocks_hdr(Context(), flat.data(), pixelFloatBlocks.size(), channels, cent.data(),
                                  (uint32_t)pixelFloatCentroids.size(), YUPVQ_ASSIGN_AUTO, idx.data(), nullptr))
        fail(ctx, "QuantizeBlocksHDR");
    return idx;
}

VQCodebook VQEncoder::EncodeBCn(const uint8_t* bcData, size_t numBlocks, uint8_t channels,
                                std::vector<std::vector<uint8_t>>& outPixelCentroids, std::vector<uint32_t>& outIndices,
                                const CompressionParams& params) {
    const size_t per = 16 * (size_t)channels;
    const yupvq_config cfg = make_cfg(config, seed);
    std::vector<uint8_t> cent((size_t)cfg.codebook_size * per);
    outIndices.resize(numBlocks);
    const double t_vq = now_ms();
    if (yupvq_encode_bcn(Context(), bcData, numBlocks, (int)bcFormat, channels, &cfg, cent.data(), outIndices.data(), nullptr))
        fail(ctx, "EncodeBCn");
    lastVqMs = now_ms() - t_vq;
    outPixelCentroids = unflatten(cent, per);
    return EncodeCentroids(outPixelCentroids, channels, params);
}

VQCodebook VQEncoder::EncodeHDR(const float* image, uint32_t width, uint32_t height, uint8_t channels,
                                std::vector<std::vector<float>>& outPixelFloatCentroids, std::vector<uint32_t>& outIndices,
                                const CompressionParams& params) {
    const size_t per = 16 * (size_t)channels;
    const yupvq_config cfg = make_cfg(config, seed);
    std::vector<float> cent((size_t)cfg.codebook_size * per);
    outIndices.resize((size_t)((width + 3) / 4) * ((height + 3) / 4));
    const double t_vq = now_ms();
    if (yupvq_encode_hdr(Context(), image, width, height, channels, &cfg, cent.data(), outIndices.data(), nullptr))
        fail(ctx, "EncodeHDR");
    lastVqMs = now_ms() - t_vq;
    outPixelFloatCentroids = unflatten(cent, per);
    return EncodeCentroidsHDR(outPixelFloatCentroids, channels, params);
}

// ---- VQBCnCompressor (reference src/vq_bcn_compressor.cpp) --------------------------------------------------------
struct VQBCnCompressor::ZstdContext {
    void* cctx = nullptr;
    void* dctx = nullptr;
    void* cdict = nullptr;
    void* ddict = nullptr;
    void ensure() {  // created on first use so that useZstd = false needs no libzstd at all
        if (cctx) return;
        Zstd& z = Zstd::get();
        cctx = z.createCCtx();
        dctx = z.createDCtx();
        if (!cctx || !dctx) throw std::runtime_error("Failed to create ZSTD context");
    }
    ~ZstdContext() {
        if (!cctx && !dctx && !cdict && !ddict) return;
        Zstd& z = Zstd::get();
        if (cctx) z.freeCCtx(cctx);
        if (dctx) z.freeDCtx(dctx);
        if (cdict) z.freeCDict(cdict);
        if (ddict) z.freeDDict(ddict);
    }
};

VQBCnCompressor::VQBCnCompressor() : zstdCtx(std::make_unique<ZstdContext>()), seed_(default_seed()) {}
VQBCnCompressor::~VQBCnCompressor() = default;

yupvq_ctx* VQBCnCompressor::SharedContext() {
    if (!gather_) gather_ = std::make_unique<VQEncoder>();
    return gather_->Context();
}

void VQBCnCompressor::LoadDictionary(const uint8_t* dictData, size_t dictSize) {  // :63-78
    Zstd& z = Zstd::get();
    if (zstdCtx->cdict) z.freeCDict(zstdCtx->cdict);
    if (zstdCtx->ddict) z.freeDDict(zstdCtx->ddict);
    zstdCtx->cdict = z.createCDict(dictData, dictSize, 1);
    if (!zstdCtx->cdict) throw std::runtime_error("Failed to create ZSTD compression dictionary");
    zstdCtx->ddict = z.createDDict(dictData, dictSize);
    if (!zstdCtx->ddict) throw std::runtime_error("Failed to create ZSTD decompression dictionary");
}

std::vector<uint8_t> VQBCnCompressor::compressWithZstd(const std::vector<uint8_t>& payload, int level, int numThreads, bool enableLdm) {  // :20-53
    Zstd& z = Zstd::get();
    zstdCtx->ensure();
    std::vector<uint8_t> out(z.compressBound(payload.size()));
    z.CCtx_setParameter(zstdCtx->cctx, Zstd::c_nbWorkers, numThreads);
    z.CCtx_setParameter(zstdCtx->cctx, Zstd::c_compressionLevel, level);
    z.CCtx_setParameter(zstdCtx->cctx, Zstd::c_enableLongDistanceMatching, enableLdm ? 1 : 0);
    const size_t sz = zstdCtx->cdict
                          ? z.compress_usingCDict(zstdCtx->cctx, out.data(), out.size(), payload.data(), payload.size(), zstdCtx->cdict)
                          : z.compress2(zstdCtx->cctx, out.data(), out.size(), payload.data(), payload.size());
    if (z.isError(sz)) throw std::runtime_error(std::string("Zstd compression failed: ") + z.getErrorName(sz));
    out.resize(sz);
    return out;
}

namespace {

TextureInfo make_info(uint32_t w, uint32_t h, uint8_t ch, const CompressionParams& p, uint32_t flags) {
    TextureInfo info;
    std::memset(static_cast<void*>(&info), 0, sizeof(info));  // the reference streams three uninitialised padding bytes; zero them
    info.width = w;
    info.height = h;
    info.format = p.bcFormat;
    info.originalChannelCount = ch;
    info.compressionFlags = flags;
    if (!p.useVQ) info.compressionFlags |= COMPRESSION_FLAGS_VQ_BYPASSED;
    if (!p.useZstd) info.compressionFlags |= COMPRESSION_FLAGS_ZSTD_BYPASSED;
    if (w < 32 || h < 32) info.compressionFlags |= COMPRESSION_FLAGS_VQ_BYPASSED | COMPRESSION_FLAGS_ZSTD_BYPASSED;  // :102-106
    return info;
}

CompressionConfig vq_config(const CompressionParams& p, bool hdr) {  // :128-134 / :220-225
    CompressionConfig c;
    if (!hdr) c.metric = p.vq_Metric;
    c.fastModeSampleRatio = p.vq_FastModeSampleRatio;
    c.maxIterations = p.vq_maxIterations;
    c.SetQuality(p.quality);                // before the powers are copied: they never take effect
    c.min_cb_power = p.vq_min_cb_power;
    c.max_cb_power = p.vq_max_cb_power;
    return c;
}

std::vector<uint8_t> payload_of(const VQCodebook& cb, const std::vector<uint32_t>& idx, size_t block_size) {  // :152-162
    std::vector<uint8_t> payload(cb.entries.size() * block_size + idx.size() * sizeof(uint32_t));
    size_t off = 0;
    for (const auto& e : cb.entries) {
        if (e.size() < block_size) throw std::runtime_error("codebook entry encode failed");
        std::memcpy(payload.data() + off, e.data(), block_size);
        off += block_size;
    }
    std::memcpy(payload.data() + off, idx.data(), idx.size() * sizeof(uint32_t));
    return payload;
}

std::vector<uint8_t> serialise(const TextureInfo& info, const std::vector<uint8_t>& payload) {  // :173-177
    std::vector<uint8_t> out(sizeof(TextureInfo) + payload.size());
    std::memcpy(out.data(), &info, sizeof(TextureInfo));
    std::memcpy(out.data() + sizeof(TextureInfo), payload.data(), payload.size());
    return out;
}

}  // namespace

std::vector<uint8_t> VQBCnCompressor::Compress(const uint8_t* inData, uint32_t width, uint32_t height, uint8_t channels,
                                               const CompressionParams& params) {
    TextureInfo info = make_info(width, height, channels, params, COMPRESSION_FLAGS_DEFAULT);
    const bool enableLdm = width >= 4000 || height >= 4000;
    timings_ = Timings();
    const double t0 = now_ms();
    auto bcData = bcnCompressor.Compress(inData, width, height, channels, params.bcFormat, params.numThreads, params.bcQuality,
                                         params.alphaThreshold, /*flipRGB=*/!params.useVQ);
    if (bcData.empty()) throw std::runtime_error("BCn compression failed");
    timings_.bcnEncodeMs = now_ms() - t0;
    std::vector<uint8_t> payload;
    if (info.compressionFlags & COMPRESSION_FLAGS_VQ_BYPASSED) {
        info.storedCodebookEntries = 0;
        payload = std::move(bcData);
    } else {
        VQEncoder enc(vq_config(params, false));
        enc.AttachContext(SharedContext());
        enc.SetFormat(params.bcFormat);
        enc.SetSeed(seed_);
        const size_t bs = BCBlockSize::GetSize(params.bcFormat);
        std::vector<std::vector<uint8_t>> centroids;
        std::vector<uint32_t> indices;
        // the three starred lines of the reference (:142-150) as one device call
        VQCodebook cb = enc.EncodeBCn(bcData.data(), bcData.size() / bs, channels, centroids, indices, params);
        timings_.vqMs = enc.LastVqMs();
        timings_.centroidEncodeMs = enc.LastCentroidEncodeMs();
        info.storedCodebookEntries = (uint32_t)cb.entries.size();
        payload = payload_of(cb, indices, bs);
    }
    const double tz = now_ms();
    if (!(info.compressionFlags & COMPRESSION_FLAGS_ZSTD_BYPASSED))
        payload = compressWithZstd(payload, params.zstdLevel, params.numThreads, enableLdm);
    timings_.zstdMs = now_ms() - tz;
    auto out = serialise(info, payload);
    timings_.totalMs = now_ms() - t0;
    return out;
}

std::vector<uint8_t> VQBCnCompressor::CompressHDR(const float* inData, uint32_t width, uint32_t height, uint8_t channels,
                                                  const CompressionParams& params) {
    TextureInfo info = make_info(width, height, channels, params, COMPRESSION_FLAGS_IS_HDR);
    const bool enableLdm = width >= 4000 || height >= 4000;
    timings_ = Timings();
    const double t0 = now_ms();
    std::vector<uint8_t> payload;
    if (info.compressionFlags & COMPRESSION_FLAGS_VQ_BYPASSED) {
        info.storedCodebookEntries = 0;
        payload = bcnCompressor.CompressHDR(inData, width, height, channels, params.bcFormat, params.numThreads, params.bcQuality);
        if (payload.empty()) throw std::runtime_error("HDR BCn compression failed");
    } else {
        VQEncoder enc(vq_config(params, true));
        enc.AttachContext(SharedContext());
        enc.SetFormat(params.bcFormat);
        enc.SetSeed(seed_);
        std::vector<std::vector<float>> centroids;
        std::vector<uint32_t> indices;
        VQCodebook cb = enc.EncodeHDR(inData, width, height, channels, centroids, indices, params);  // :230-256 on the device
        timings_.vqMs = enc.LastVqMs();
        timings_.centroidEncodeMs = enc.LastCentroidEncodeMs();
        info.storedCodebookEntries = (uint32_t)cb.entries.size();
        payload = payload_of(cb, indices, BCBlockSize::GetSize(params.bcFormat));
    }
    const double tz = now_ms();
    if (!(info.compressionFlags & COMPRESSION_FLAGS_ZSTD_BYPASSED))
        payload = compressWithZstd(payload, params.zstdLevel, params.numThreads, enableLdm);
    timings_.zstdMs = now_ms() - tz;
    auto out = serialise(info, payload);
    timings_.totalMs = now_ms() - t0;
    return out;
}

std::vector<uint8_t> VQBCnCompressor::DecompressToBCn(const std::vector<uint8_t>& compressedData, TextureInfo& outInfo, int) {  // :287-361
    if (compressedData.size() < sizeof(TextureInfo))
        throw std::runtime_error("Compressed data stream is too small to contain a header. Cannot decompress.");
    std::memcpy(&outInfo, compressedData.data(), sizeof(TextureInfo));
    const uint8_t* src = compressedData.data() + sizeof(TextureInfo);
    const size_t src_size = compressedData.size() - sizeof(TextureInfo);
    std::vector<uint8_t> payload;
    if (outInfo.compressionFlags & COMPRESSION_FLAGS_ZSTD_BYPASSED) {
        payload.assign(src, src + src_size);
    } else {
        Zstd& z = Zstd::get();
        zstdCtx->ensure();
        unsigned long long want = z.getFrameContentSize(src, src_size);
        if (want == Zstd::CONTENTSIZE_ERROR || want == Zstd::CONTENTSIZE_UNKNOWN) want = src_size * 10;
        payload.resize((size_t)want);
        const size_t got = zstdCtx->ddict ? z.decompress_usingDDict(zstdCtx->dctx, payload.data(), payload.size(), src, src_size, zstdCtx->ddict)
                                          : z.decompressDCtx(zstdCtx->dctx, payload.data(), payload.size(), src, src_size);
        if (z.isError(got)) throw std::runtime_error(std::string("Zstd decompression failed: ") + z.getErrorName(got));
        payload.resize(got);
    }
    if (outInfo.compressionFlags & COMPRESSION_FLAGS_VQ_BYPASSED) return payload;
    const size_t bs = BCBlockSize::GetSize(outInfo.format);
    const size_t cb_bytes = (size_t)outInfo.storedCodebookEntries * bs, n = outInfo.GetTotalBlocks();
    if (payload.size() != cb_bytes + n * sizeof(uint32_t))
        throw std::runtime_error("Decompressed data size mismatch. Expected codebook + indices.");
    // bc[i] = codebook[indices[i]] (:350-358) as a device gather
    std::vector<uint8_t> bc(n * bs);
    if (yupvq_gather_codebook(SharedContext(), payload.data(), outInfo.storedCodebookEntries, (uint32_t)bs,
                              reinterpret_cast<const uint32_t*>(payload.data() + cb_bytes), n, bc.data()))
        throw std::runtime_error(std::string("DecompressToBCn: ") + yupvq_last_error(gather_->Context()));
    return bc;
}

std::vector<uint8_t> VQBCnCompressor::Decompress(const std::vector<uint8_t>& compressedData, TextureInfo& outInfo) {  // :363-376
    auto bc = DecompressToBCn(compressedData, outInfo);
    if (outInfo.compressionFlags & COMPRESSION_FLAGS_IS_HDR)
        throw std::runtime_error("Cannot decompress HDR texture to 8-bit data. Use DecompressHDR instead.");
    return bcnCompressor.Decompress(bc.data(), outInfo.width, outInfo.height, outInfo.originalChannelCount, outInfo.format);
}

std::vector<float> VQBCnCompressor::DecompressHDR(const std::vector<uint8_t>& compressedData, TextureInfo& outInfo) {  // :378-390
    auto bc = DecompressToBCn(compressedData, outInfo);
    if (!(outInfo.compressionFlags & COMPRESSION_FLAGS_IS_HDR))
        throw std::runtime_error("Cannot decompress LDR texture to float data. Use Decompress instead.");
    return bcnCompressor.DecompressHDR(bc.data(), outInfo.width, outInfo.height, outInfo.originalChannelCount, outInfo.format);
}

// ---- plain-C entry points over the classes (ctypes / FFI callers: bench.py's whole-Compress breakdown) -----------------
// out_ms[5] = {BCn pre-pass encode, device VQ core, K centroid encodes (A16), zstd, whole call}; returns 0 on success,
// the stream size in *out_bytes and (if out_stream != null, capacity out_cap) the stream itself.
extern "C" {

YUPTEXTURE_API const char* yuptexture_codec_name(void) { return BCnCompressor::CodecName(); }

YUPTEXTURE_API int yuptexture_load_codec(const char* path) { return BCnCompressor::LoadCodecLibrary(path) ? 0 : 1; }
// -1 auto, 0 off, 1 on (BCnCompressor::SetGpuEncode); the encoder the calling thread's last Compress used
YUPTEXTURE_API void yuptexture_set_gpu_bcn_encode(int mode) {
    BCnCompressor::SetGpuEncode(mode < 0 ? BCnCompressor::kGpuAuto : (mode ? BCnCompressor::kGpuOn : BCnCompressor::kGpuOff));
}
YUPTEXTURE_API const char* yuptexture_last_encoder_name(void) { return BCnCompressor::LastEncoderName(); }

YUPTEXTURE_API int yuptexture_compress_timed(const void* image, int is_hdr, uint32_t width, uint32_t height, uint32_t channels,
                                             const CompressionParams* params, uint64_t seed, double out_ms[5], size_t* out_bytes,
                                             uint8_t* out_stream, size_t out_cap, char* err, size_t err_cap) {
    try {
        static thread_local std::unique_ptr<VQBCnCompressor> comp;   // keeps the device context (and its arena) across calls
        if (!comp) comp = std::make_unique<VQBCnCompressor>();
        comp->SetSeed(seed);
        const std::vector<uint8_t> s = is_hdr ? comp->CompressHDR(static_cast<const float*>(image), width, height, (uint8_t)channels, *params)
                                              : comp->Compress(static_cast<const uint8_t*>(image), width, height, (uint8_t)channels, *params);
        const VQBCnCompressor::Timings& t = comp->LastTimings();
        if (out_ms) {
            out_ms[0] = t.bcnEncodeMs; out_ms[1] = t.vqMs; out_ms[2] = t.centroidEncodeMs; out_ms[3] = t.zstdMs; out_ms[4] = t.totalMs;
        }
        if (out_bytes) *out_bytes = s.size();
        if (out_stream && out_cap >= s.size()) std::memcpy(out_stream, s.data(), s.size());
        return 0;
    } catch (const std::exception& e) {
        if (err && err_cap) std::snprintf(err, err_cap, "%s", e.what());
        return 1;
    }
}

// VQBCnCompressor::DecompressToBCn on a stream (ours or the reference's); out must hold GetTotalBlocks() * blockSize bytes
YUPTEXTURE_API int yuptexture_decompress_to_bcn(const uint8_t* stream, size_t stream_bytes, uint8_t* out, size_t out_cap, size_t* out_bytes,
                                                TextureInfo* out_info, char* err, size_t err_cap) {
    try {
        static thread_local std::unique_ptr<VQBCnCompressor> comp;
        if (!comp) comp = std::make_unique<VQBCnCompressor>();
        TextureInfo info;
        const std::vector<uint8_t> bc = comp->DecompressToBCn(std::vector<uint8_t>(stream, stream + stream_bytes), info);
        if (out_info) *out_info = info;
        if (out_bytes) *out_bytes = bc.size();
        if (out && out_cap >= bc.size()) std::memcpy(out, bc.data(), bc.size());
        return 0;
    } catch (const std::exception& e) {
        if (err && err_cap) std::snprintf(err, err_cap, "%s", e.what());
        return 1;
    }
}

}  // extern "C"
