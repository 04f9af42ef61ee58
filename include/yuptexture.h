// yuptexture.h -- C++ host API of the B200-native YupTexture VQ path.
//
// Source-compatible with the reference's public headers (YupCore/YupTexture @ 7c8d4366):
//   include/vq_bcn_types.h      -> BCFormat, BCBlockSize, CompressionFlags, TextureInfo, VQCodebook,
//                                  DistanceMetric, CompressionParams            (:27-126)
//   include/vq_encoder.h        -> CompressionConfig (:22-41), VQEncoder (:43-104)
//   include/bcn_compressor.h    -> BCnCompressor (:7-59)
//   include/vq_bcn_compressor.h -> VQBCnCompressor (:12-60)
// The four headers of those names in this directory simply include this file, so a caller such as the
// reference's example/example.cpp:297-300 recompiles unchanged.  Struct layouts are identical
// (sizeof(TextureInfo) == 24 with the offsets the reference streams to disk, src/vq_bcn_compressor.cpp:173-175).
//
// What is different underneath: every VQEncoder call runs on the GPU through the C ABI of
// include/yupvq.h (no CPU fallback; construction throws std::runtime_error without a B200), and
// VQBCnCompressor::Compress hands the BCn bytes to the device once (decode + seeding + Lloyd + final
// assignment fused behind yupvq_encode_bcn) instead of materialising N std::vector pixel blocks.
#pragma once

#include <cstddef>
#include <cstdint>
#include <functional>
#include <memory>
#include <string>
#include <vector>

#if defined(__GNUC__) && __GNUC__ >= 4
#define YUPTEXTURE_API __attribute__((visibility("default")))
#else
#define YUPTEXTURE_API
#endif

// ---- shared types (reference include/vq_bcn_types.h) ---------------------------------------------------

enum class BCFormat { Unknown = 0, BC1 = 1, BC2, BC3, BC4, BC5, BC6H, BC7 };

struct BCBlockSize {
    static constexpr size_t BC1 = 8, BC2 = 16, BC3 = 16, BC4 = 8, BC5 = 16, BC6H = 16, BC7 = 16;
    static size_t GetSize(BCFormat format) { return (format == BCFormat::BC1 || format == BCFormat::BC4) ? 8 : 16; }
};

enum CompressionFlags : uint32_t {
    COMPRESSION_FLAGS_DEFAULT = 0,
    COMPRESSION_FLAGS_VQ_BYPASSED = 1u << 0,    // payload is raw BCn data
    COMPRESSION_FLAGS_ZSTD_BYPASSED = 1u << 1,  // payload is not zstd-compressed
    COMPRESSION_FLAGS_IS_HDR = 1u << 2,
    COMPRESSION_FLAGS_USES_PQ = 1u << 3
};

// 24-byte stream header: width@0 height@4 format@8 originalChannelCount@12 storedCodebookEntries@16
// compressionFlags@20 (bytes 13-15 are padding the reference leaves uninitialised; zeroed here).
struct TextureInfo {
    uint32_t width = 0;
    uint32_t height = 0;
    BCFormat format = BCFormat::BC1;
    uint8_t originalChannelCount = 4;
    uint32_t storedCodebookEntries = 0;
    uint32_t compressionFlags = COMPRESSION_FLAGS_DEFAULT;

    size_t GetBlocksX() const { return (width + 3) / 4; }
    size_t GetBlocksY() const { return (height + 3) / 4; }
    size_t GetTotalBlocks() const { return GetBlocksX() * GetBlocksY(); }
};
static_assert(sizeof(TextureInfo) == 24, "TextureInfo is a serialised header");

struct VQCodebook {
    std::vector<std::vector<uint8_t>> entries;
    uint32_t blockSize = 0;
    uint32_t codebookSize = 0;
    VQCodebook() = default;
    VQCodebook(uint32_t bSize, uint32_t cbSize) : blockSize(bSize), codebookSize(cbSize) {}
};

enum class DistanceMetric { SAD_SIMD, PERCEPTUAL_OKLAB };

struct CompressionParams {
    BCFormat bcFormat = BCFormat::BC7;
    float bcQuality = 1.0f;
    int zstdLevel = 3;
    int numThreads = 16;
    uint8_t alphaThreshold = 128;
    bool useVQ = true;
    bool useZstd = true;
    float vq_FastModeSampleRatio = 1.0f;
    float quality = 0.5f;
    DistanceMetric vq_Metric = DistanceMetric::PERCEPTUAL_OKLAB;
    uint32_t vq_min_cb_power = 4;   // no effect, as in the reference
    uint32_t vq_max_cb_power = 10;  // no effect, as in the reference
    uint32_t vq_maxIterations = 32;
};

// ---- BCn codec wrapper (reference include/bcn_compressor.h) ---------------------------------------------
// The reference wraps AMD Compressonator here; that library stays a host-side dependency of the
// integrator.  This build decodes on the GPU (bit-exact with Compressonator 4.3 for
// BC1/BC3/BC4/BC5/BC7) and ships small bounding-box encoders for BC1/BC3/BC4/BC5 and a mode-6 BC7
// encoder so the API is usable stand-alone; an integrator installs the real encoders with SetCodec().
// The stand-ins are not silent: the first time one of them writes blocks the library prints one line on stderr naming the
// fallback (YUPTEXTURE_QUIET=1 suppresses it); LastEncoderName() reports which encoder a call used.
class YUPTEXTURE_API BCnCompressor {
public:
    struct Codec {  // any member left empty keeps the built-in behaviour
        std::function<std::vector<uint8_t>(const uint8_t*, uint32_t, uint32_t, uint32_t, BCFormat, int, float, uint8_t, bool)> compress;
        std::function<std::vector<uint8_t>(const float*, uint32_t, uint32_t, uint32_t, BCFormat, int, float)> compressHDR;
        std::function<std::vector<float>(const uint8_t*, uint32_t, uint32_t, uint32_t, BCFormat)> decompressHDR;
    };
    static void SetCodec(const Codec& codec);
    // Installs the encoders of a block-codec plugin (include/yupcodec.h), e.g. libyupcodec_cmp.so (Compressonator).
    // Without an explicit call the first use looks for $YUPTEXTURE_CODEC, then for libyupcodec_cmp.so next to
    // libyuptexture.so.  Returns false (and leaves the current codec) if the library cannot be loaded.
    static bool LoadCodecLibrary(const char* path);
    static const char* CodecName();   // "built-in" or the plugin's yupcodec_name()
    // The whole-texture pre-pass of Compress() on the GPU (yupvq_bcn_encode_image: BC1, BC3, BC4, BC5).
    // kGpuAuto (default): the GPU encodes when no codec is installed -- with a codec (Compressonator) the blocks stay the
    // reference's, byte for byte; kGpuOn: the GPU encodes whenever the format has a GPU encoder (a 4096^2 RGB pre-pass
    // drops from seconds to under a millisecond of kernel time, decoded PSNR >= Compressonator's); kGpuOff: never.
    // $YUPTEXTURE_GPU_BCN = 1 / 0 sets the default of a process.
    enum GpuEncode { kGpuAuto = -1, kGpuOff = 0, kGpuOn = 1 };
    static void SetGpuEncode(GpuEncode mode);
    static const char* LastEncoderName();   // which encoder the calling thread's last Compress() used

    std::vector<uint8_t> Compress(const uint8_t* inData, uint32_t width, uint32_t height, uint32_t channelCount,
                                  BCFormat format, int numThreads, float quality = 1.0f, uint8_t alphaThreshold = 128,
                                  bool flipRGB = false);
    std::vector<uint8_t> CompressHDR(const float* inData, uint32_t width, uint32_t height, uint32_t channelCount,
                                     BCFormat format, int numThreads, float quality = 1.0f);
    std::vector<uint8_t> Decompress(const uint8_t* bcData, uint32_t width, uint32_t height, uint32_t channelCount,
                                    BCFormat format);
    std::vector<float> DecompressHDR(const uint8_t* bcData, uint32_t width, uint32_t height, uint32_t channelCount,
                                     BCFormat format);
};

// ---- VQ encoder (reference include/vq_encoder.h) -----------------------------------------------------------

struct CompressionConfig {
    uint32_t maxIterations = 32;
    DistanceMetric metric = DistanceMetric::PERCEPTUAL_OKLAB;
    float fastModeSampleRatio = 1.0f;
    uint32_t min_cb_power = 4;
    uint32_t max_cb_power = 10;
    CompressionConfig(float quality_level = 0.5f);
    void SetQuality(float quality_level);  // codebookSize = 1 << (min + round(q * (max - min)))
    float quality = 0.5f;
    uint32_t codebookSize;
};

struct yupvq_ctx;  // include/yupvq.h

class YUPTEXTURE_API VQEncoder {
public:
    VQEncoder(const CompressionConfig& cfg = CompressionConfig());
    ~VQEncoder();
    VQEncoder(const VQEncoder&) = delete;
    VQEncoder& operator=(const VQEncoder&) = delete;
    void SetFormat(BCFormat format);

    VQCodebook BuildCodebook(const std::vector<std::vector<uint8_t>>& allPixelBlocks, uint8_t channels,
                             std::vector<std::vector<uint8_t>>& outPixelCentroids, const CompressionParams& params);
    std::vector<uint32_t> QuantizeBlocks(const std::vector<std::vector<uint8_t>>& pixelBlocks, uint8_t channels,
                                         const std::vector<std::vector<uint8_t>>& pixelCentroids,
                                         const CompressionParams& params);
    VQCodebook BuildCodebookHDR(const std::vector<std::vector<float>>& allPixelFloatBlocks, uint8_t channels,
                                std::vector<std::vector<float>>& outPixelFloatCentroids, const CompressionParams& params);
    std::vector<uint32_t> QuantizeBlocksHDR(const std::vector<std::vector<float>>& pixelFloatBlocks, uint8_t channels,
                                            const std::vector<std::vector<float>>& pixelFloatCentroids,
                                            const CompressionParams& params);

    // ---- extensions (not in the reference) ----
    // The reference seeds std::mt19937 from std::random_device and has no seed parameter
    // (src/vq_encoder.cpp:24-26).  0 keeps that behaviour; the YUPVQ_SEED environment variable sets
    // the default for processes that cannot call this.
    void SetSeed(uint64_t seed);
    // CUDA device of this encoder (default: YUPVQ_DEVICE or 0).  Must be called before the first use.
    void SetDevice(int device);
    // The fused path VQBCnCompressor uses: BCn bytes in, pixel centroids + indices out, one H2D/D2H.
    VQCodebook EncodeBCn(const uint8_t* bcData, size_t numBlocks, uint8_t channels,
                         std::vector<std::vector<uint8_t>>& outPixelCentroids, std::vector<uint32_t>& outIndices,
                         const CompressionParams& params);
    VQCodebook EncodeHDR(const float* image, uint32_t width, uint32_t height, uint8_t channels,
                         std::vector<std::vector<float>>& outPixelFloatCentroids, std::vector<uint32_t>& outIndices,
                         const CompressionParams& params);
    yupvq_ctx* Context();  // lazily created
    // Use a context owned by someone else (VQBCnCompressor keeps one across its Compress calls, so the per-call VQEncoder
    // of the reference's structure, src/vq_bcn_compressor.cpp:136, does not pay for a new device arena each time).
    void AttachContext(yupvq_ctx* shared);
    // wall-clock milliseconds of the last Encode* / BuildCodebook* call: the device VQ core (incl. its H2D / D2H) and
    // the K host-side centroid encodes (A16), timed separately
    double LastVqMs() const { return lastVqMs; }
    double LastCentroidEncodeMs() const { return lastCentroidMs; }

private:
    CompressionConfig config;
    BCFormat bcFormat = BCFormat::BC1;
    BCnCompressor bcnCompressor;
    uint64_t seed = 0;
    int device = 0;
    yupvq_ctx* ctx = nullptr;
    bool ownsCtx = true;
    double lastVqMs = 0.0, lastCentroidMs = 0.0;
    VQCodebook EncodeCentroids(const std::vector<std::vector<uint8_t>>& pixelCentroids, uint8_t channels,
                               const CompressionParams& params);
    VQCodebook EncodeCentroidsHDR(const std::vector<std::vector<float>>& centroids, uint8_t channels,
                                  const CompressionParams& params);
};

// ---- container + orchestration (reference include/vq_bcn_compressor.h) --------------------------------------

class YUPTEXTURE_API VQBCnCompressor {
public:
    VQBCnCompressor();
    ~VQBCnCompressor();
    void LoadDictionary(const uint8_t* dictData, size_t dictSize);
    std::vector<uint8_t> Compress(const uint8_t* inData, uint32_t width, uint32_t height, uint8_t channels,
                                  const CompressionParams& params);
    std::vector<uint8_t> CompressHDR(const float* inData, uint32_t width, uint32_t height, uint8_t channels,
                                     const CompressionParams& params);
    std::vector<uint8_t> DecompressToBCn(const std::vector<uint8_t>& compressedData, TextureInfo& outInfo, int numThreads = 16);
    std::vector<uint8_t> Decompress(const std::vector<uint8_t>& compressedData, TextureInfo& outInfo);
    std::vector<float> DecompressHDR(const std::vector<uint8_t>& compressedData, TextureInfo& outInfo);

    // extension: seed forwarded to the VQEncoder of each Compress call (0 = std::random_device)
    void SetSeed(uint64_t seed) { seed_ = seed; }
    // extension: stage times of the last Compress / CompressHDR call in wall-clock milliseconds, the breakdown the
    // reference's example prints around its call (example/example.cpp:293-304)
    struct Timings {
        double bcnEncodeMs = 0, vqMs = 0, centroidEncodeMs = 0, zstdMs = 0, totalMs = 0;
    };
    const Timings& LastTimings() const { return timings_; }

private:
    Timings timings_;
    BCnCompressor bcnCompressor;
    struct ZstdContext;
    std::unique_ptr<ZstdContext> zstdCtx;
    std::unique_ptr<VQEncoder> gather_;  // owns the device context shared by the Compress calls and the DecompressToBCn gather
    yupvq_ctx* SharedContext();
    uint64_t seed_ = 0;
    std::vector<uint8_t> compressWithZstd(const std::vector<uint8_t>& payload, int level, int numThreads, bool enableLdm);
};
