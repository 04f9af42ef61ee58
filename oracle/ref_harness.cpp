// TEST INFRASTRUCTURE — not product code.
//
// C-ABI shim around the UNMODIFIED reference sources (compiled where they lie under
// /root/reference by oracle/build_ref.sh; nothing from the reference is copied into this
// repository).  The resulting oracle/_ref/libyupref.so is the "real reference" oracle:
//   * it pins the CPU restatement in oracle/vq_oracle.cpp (tests/test_oracle_vs_ref.py),
//   * it generates the committed golden fixtures (tests/golden/make_golden.py),
//   * it is the timed CPU baseline of bench.py (cpu_baseline.kind == "reference").
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg
// may load it.
//
// The reference has no seed parameter (src/vq_encoder.cpp:24-26 seeds std::mt19937 from
// std::random_device) and its conversion/distance helpers are private (include/vq_encoder.h:44-70),
// so the harness uses the access trick from SURVEY.md Appendix C: include every std header the
// reference header pulls in, then re-include it with `private` defined away.
#include <random>
#include <cstdio>
#include <vector>
#include <atomic>
#include <algorithm>
#include <execution>
#include <numeric>
#include <limits>
#include <cmath>
#include <stdexcept>
#include <string>
#include <fstream>
#include <array>
#include <memory>
#include <cstring>
#include <cassert>
#include <sstream>
#include <iomanip>
#include <chrono>
#include <cstdint>
#include <iostream>
#include <map>
#include <functional>
#include <thread>

#define private public
#include "vq_encoder.h"
#undef private
#ifdef YUPREF_HAVE_CONTAINER
#include "vq_bcn_compressor.h"   // the container (src/vq_bcn_compressor.cpp), built when libzstd.so.1 is present
#endif

namespace {

double now_s() {
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

CompressionParams make_params(int bc_format, float bc_quality, int threads) {
    CompressionParams p;
    p.bcFormat = static_cast<BCFormat>(bc_format);
    p.bcQuality = bc_quality;
    p.numThreads = threads;
    return p;
}

// Mirrors the config copy order of src/vq_bcn_compressor.cpp:128-134 (SetQuality before the
// dead min/max power copies).
CompressionConfig make_config(float quality, uint32_t max_iter, int metric, float ratio) {
    CompressionConfig cfg;
    cfg.metric = static_cast<DistanceMetric>(metric);
    cfg.fastModeSampleRatio = ratio;
    cfg.maxIterations = max_iter;
    cfg.SetQuality(quality);
    return cfg;
}

template <typename T>
std::vector<std::vector<T>> to_blocks(const T* flat, size_t n, size_t elems) {
    std::vector<std::vector<T>> v(n);
    for (size_t i = 0; i < n; ++i) v[i].assign(flat + i * elems, flat + (i + 1) * elems);
    return v;
}

}  // namespace

extern "C" {

// SetQuality -> codebookSize (src/vq_encoder.cpp:12-18).
uint32_t ref_codebook_size(float quality) {
    CompressionConfig cfg;
    cfg.SetQuality(quality);
    return cfg.codebookSize;
}

// Whole-texture BCn encode through the reference wrapper (src/bcn_compressor.cpp:45).
// Returns bytes written or 0 on failure.
size_t ref_bcn_encode(const uint8_t* img, uint32_t w, uint32_t h, uint32_t ch, int bc_format,
                      float quality, int threads, int alpha_threshold, uint8_t* out, size_t out_cap) {
    BCnCompressor c;
    auto v = c.Compress(img, w, h, ch, static_cast<BCFormat>(bc_format), threads, quality,
                        (uint8_t)alpha_threshold, false);
    if (v.empty() || v.size() > out_cap) return 0;
    std::memcpy(out, v.data(), v.size());
    return v.size();
}

size_t ref_bcn_encode_hdr(const float* img, uint32_t w, uint32_t h, uint32_t ch, int bc_format,
                          float quality, int threads, uint8_t* out, size_t out_cap) {
    BCnCompressor c;
    auto v = c.CompressHDR(img, w, h, ch, static_cast<BCFormat>(bc_format), threads, quality);
    if (v.empty() || v.size() > out_cap) return 0;
    std::memcpy(out, v.data(), v.size());
    return v.size();
}

// Whole-texture decode (src/bcn_compressor.cpp:191) tiled into the per-block layout the VQ path
// consumes (src/vq_bcn_compressor.cpp:142-146).  Byte-identical to per-block calls for BC1/BC3/BC7
//  and the only feasible way for BC7.
int ref_bcn_decode_blocks(const uint8_t* bc, uint32_t w, uint32_t h, uint32_t ch, int bc_format,
                          uint8_t* out_blocks) {
    BCnCompressor c;
    auto img = c.Decompress(bc, w, h, ch, static_cast<BCFormat>(bc_format));
    if (img.empty()) return 1;
    const size_t bx = (w + 3) / 4, by = (h + 3) / 4;
    for (size_t b = 0; b < bx * by; ++b) {
        const size_t x0 = (b % bx) * 4, y0 = (b / bx) * 4;
        for (size_t y = 0; y < 4; ++y)
            for (size_t x = 0; x < 4; ++x) {
                const size_t px = std::min<size_t>(x0 + x, w - 1), py = std::min<size_t>(y0 + y, h - 1);
                std::memcpy(out_blocks + (b * 16 + y * 4 + x) * ch, img.data() + (py * w + px) * ch, ch);
            }
    }
    return 0;
}

// The reference's literal per-block call (src/vq_bcn_compressor.cpp:145) for n independent blocks.
int ref_bcn_decode_each(const uint8_t* bc, size_t n, uint32_t ch, int bc_format, uint8_t* out_blocks) {
    BCnCompressor c;
    const size_t bs = BCBlockSize::GetSize(static_cast<BCFormat>(bc_format));
    for (size_t i = 0; i < n; ++i) {
        auto px = c.Decompress(bc + i * bs, 4, 4, ch, static_cast<BCFormat>(bc_format));
        if (px.size() != 16 * ch) return 1;
        std::memcpy(out_blocks + i * 16 * ch, px.data(), 16 * ch);
    }
    return 0;
}

// PixelBlockToOklabBlock (src/vq_encoder.cpp:96-102).
void ref_pixels_to_oklab(const uint8_t* blocks, size_t n, uint32_t ch, float* out) {
    VQEncoder enc;
#pragma omp parallel for
    for (int64_t i = 0; i < (int64_t)n; ++i) {
        std::vector<uint8_t> b(blocks + i * 16 * ch, blocks + (i + 1) * 16 * ch);
        auto o = enc.PixelBlockToOklabBlock(b, (uint8_t)ch);
        std::memcpy(out + i * 64, o.data(), 64 * sizeof(float));
    }
}

// OklabBlockToPixelBlock (src/vq_encoder.cpp:105-111).
void ref_oklab_to_pixels(const float* oklab, size_t n, uint32_t ch, uint8_t* out) {
    VQEncoder enc;
    for (size_t i = 0; i < n; ++i) {
        OklabBlock b(oklab + i * 64, oklab + (i + 1) * 64);
        auto p = enc.OklabBlockToPixelBlock(b, (uint8_t)ch);
        std::memcpy(out + i * 16 * ch, p.data(), 16 * ch);
    }
}

// RgbaFloatBlockToOklabBlock (src/vq_encoder.cpp:179-185).
void ref_floats_to_oklab(const float* blocks, size_t n, uint32_t ch, float* out) {
    VQEncoder enc;
#pragma omp parallel for
    for (int64_t i = 0; i < (int64_t)n; ++i) {
        std::vector<float> b(blocks + i * 16 * ch, blocks + (i + 1) * 16 * ch);
        auto o = enc.RgbaFloatBlockToOklabBlock(b, (uint8_t)ch);
        std::memcpy(out + i * 64, o.data(), 64 * sizeof(float));
    }
}

// OklabBlockToRgbaFloatBlock (src/vq_encoder.cpp:188-194).
void ref_oklab_to_floats(const float* oklab, size_t n, uint32_t ch, float* out) {
    VQEncoder enc;
    for (size_t i = 0; i < n; ++i) {
        OklabFloatBlock b(oklab + i * 64, oklab + (i + 1) * 64);
        auto p = enc.OklabBlockToRgbaFloatBlock(b, (uint8_t)ch);
        std::memcpy(out + i * 16 * ch, p.data(), 16 * ch * sizeof(float));
    }
}

// OklabBlockDistanceSq_SIMD (src/vq_encoder.cpp:212-237): out[i*nb+j] = d(a_i, b_j).
void ref_oklab_distance(const float* a, size_t na, const float* b, size_t nb, float* out) {
    VQEncoder enc;
    std::vector<OklabBlock> A(na), B(nb);
    for (size_t i = 0; i < na; ++i) A[i].assign(a + i * 64, a + (i + 1) * 64);
    for (size_t j = 0; j < nb; ++j) B[j].assign(b + j * 64, b + (j + 1) * 64);
    for (size_t i = 0; i < na; ++i)
        for (size_t j = 0; j < nb; ++j) out[i * nb + j] = enc.OklabBlockDistanceSq_SIMD(A[i], B[j]);
}

// BlockDistanceSAD (src/vq_encoder.cpp:201-208).
void ref_block_sad(const uint8_t* a, size_t na, const uint8_t* b, size_t nb, uint32_t ch, float* out) {
    VQEncoder enc;
    for (size_t i = 0; i < na; ++i)
        for (size_t j = 0; j < nb; ++j)
            out[i * nb + j] = enc.BlockDistanceSAD(a + i * 16 * ch, b + j * 16 * ch, (uint8_t)ch);
}

// BuildCodebook with a pinned seed (src/vq_encoder.cpp:260-441).
// out_centroids: K x 16*ch bytes.  out_entries (may be null): K x blockSize BCn-encoded entries.
// times[0] = BuildCodebook wall seconds.
int ref_build_codebook(const uint8_t* blocks, size_t n, uint32_t ch, int bc_format, float quality,
                       uint32_t max_iter, int metric, float ratio, uint32_t seed, int threads,
                       float bc_quality, uint8_t* out_centroids, uint8_t* out_entries, double* times) {
    auto px = to_blocks<uint8_t>(blocks, n, 16 * ch);
    VQEncoder enc(make_config(quality, max_iter, metric, ratio));
    enc.SetFormat(static_cast<BCFormat>(bc_format));
    enc.rng.seed(seed);
    auto params = make_params(bc_format, bc_quality, threads);
    std::vector<std::vector<uint8_t>> cent;
    const double t0 = now_s();
    VQCodebook cb = enc.BuildCodebook(px, (uint8_t)ch, cent, params);
    if (times) times[0] = now_s() - t0;
    for (size_t k = 0; k < cent.size(); ++k) std::memcpy(out_centroids + k * 16 * ch, cent[k].data(), 16 * ch);
    if (out_entries)
        for (size_t k = 0; k < cb.entries.size(); ++k)
            if (cb.entries[k].size() == cb.blockSize)
                std::memcpy(out_entries + k * cb.blockSize, cb.entries[k].data(), cb.blockSize);
    return (int)cent.size();
}

// QuantizeBlocks (src/vq_encoder.cpp:447-486).
int ref_quantize_blocks(const uint8_t* blocks, size_t n, uint32_t ch, const uint8_t* centroids, uint32_t k,
                        int metric, int threads, uint32_t* out_idx, double* times) {
    auto px = to_blocks<uint8_t>(blocks, n, 16 * ch);
    auto ce = to_blocks<uint8_t>(centroids, k, 16 * ch);
    CompressionConfig cfg;
    cfg.metric = static_cast<DistanceMetric>(metric);
    VQEncoder enc(cfg);
    auto params = make_params(1, 1.0f, threads);
    const double t0 = now_s();
    auto idx = enc.QuantizeBlocks(px, (uint8_t)ch, ce, params);
    if (times) times[0] = now_s() - t0;
    std::memcpy(out_idx, idx.data(), idx.size() * sizeof(uint32_t));
    return (int)idx.size();
}

// BuildCodebookHDR with a pinned seed (src/vq_encoder.cpp:493-622).
int ref_build_codebook_hdr(const float* blocks, size_t n, uint32_t ch, int bc_format, float quality,
                           uint32_t max_iter, float ratio, uint32_t seed, int threads, float bc_quality,
                           float* out_centroids, uint8_t* out_entries, double* times) {
    auto px = to_blocks<float>(blocks, n, 16 * ch);
    VQEncoder enc(make_config(quality, max_iter, (int)DistanceMetric::PERCEPTUAL_OKLAB, ratio));
    enc.SetFormat(static_cast<BCFormat>(bc_format));
    enc.rng.seed(seed);
    auto params = make_params(bc_format, bc_quality, threads);
    std::vector<std::vector<float>> cent;
    const double t0 = now_s();
    VQCodebook cb = enc.BuildCodebookHDR(px, (uint8_t)ch, cent, params);
    if (times) times[0] = now_s() - t0;
    for (size_t k = 0; k < cent.size(); ++k)
        std::memcpy(out_centroids + k * 16 * ch, cent[k].data(), 16 * ch * sizeof(float));
    if (out_entries)
        for (size_t k = 0; k < cb.entries.size(); ++k)
            if (cb.entries[k].size() == cb.blockSize)
                std::memcpy(out_entries + k * cb.blockSize, cb.entries[k].data(), cb.blockSize);
    return (int)cent.size();
}

// QuantizeBlocksHDR (src/vq_encoder.cpp:629-658).
int ref_quantize_blocks_hdr(const float* blocks, size_t n, uint32_t ch, const float* centroids, uint32_t k,
                            int threads, uint32_t* out_idx, double* times) {
    auto px = to_blocks<float>(blocks, n, 16 * ch);
    auto ce = to_blocks<float>(centroids, k, 16 * ch);
    VQEncoder enc;
    auto params = make_params(6, 1.0f, threads);
    const double t0 = now_s();
    auto idx = enc.QuantizeBlocksHDR(px, (uint8_t)ch, ce, params);
    if (times) times[0] = now_s() - t0;
    std::memcpy(out_idx, idx.data(), idx.size() * sizeof(uint32_t));
    return (int)idx.size();
}

// One BCn 4x4 encode as the reference's finalise step does it (src/vq_encoder.cpp:248-250).
size_t ref_encode_single_block(const uint8_t* px, uint32_t ch, int bc_format, float bc_quality, int threads,
                               uint8_t* out) {
    BCnCompressor c;
    auto v = c.Compress(px, 4, 4, ch, static_cast<BCFormat>(bc_format), threads, bc_quality);
    std::memcpy(out, v.data(), v.size());
    return v.size();
}

// One BC6H 4x4 encode as BuildCodebookHDR's finalise step does it (src/vq_encoder.cpp:252-254, :617-620).
size_t ref_encode_single_block_hdr(const float* px, uint32_t ch, int bc_format, float bc_quality, int threads, uint8_t* out) {
    BCnCompressor c;
    auto v = c.CompressHDR(px, 4, 4, ch, static_cast<BCFormat>(bc_format), threads, bc_quality);
    std::memcpy(out, v.data(), v.size());
    return v.size();
}

int ref_num_threads() { return (int)std::thread::hardware_concurrency(); }

// ---- the container: VQBCnCompressor::Compress / CompressHDR / DecompressToBCn (src/vq_bcn_compressor.cpp) ----------
int ref_have_container() {
#ifdef YUPREF_HAVE_CONTAINER
    return 1;
#else
    return 0;
#endif
}

#ifdef YUPREF_HAVE_CONTAINER
// params_blob: a CompressionParams (44 bytes, include/vq_bcn_types.h:110-126) filled by the caller.
// The reference seeds its VQEncoder from std::random_device inside Compress (:136, src/vq_encoder.cpp:24-26), so the
// stream differs from run to run; times[0] = wall seconds of the call.
long long ref_compress(const void* image, int is_hdr, uint32_t w, uint32_t h, uint32_t ch, const void* params_blob,
                       uint8_t* out, size_t out_cap, double* times) {
    try {
        CompressionParams p;
        std::memcpy(static_cast<void*>(&p), params_blob, sizeof(p));
        VQBCnCompressor c;
        const double t0 = now_s();
        // Compress may swizzle its input in place when useVQ is false (src/bcn_compressor.cpp:61-76): work on a copy
        std::vector<uint8_t> copy;
        std::vector<uint8_t> s;
        if (is_hdr) {
            s = c.CompressHDR(static_cast<const float*>(image), w, h, (uint8_t)ch, p);
        } else {
            copy.assign(static_cast<const uint8_t*>(image), static_cast<const uint8_t*>(image) + (size_t)w * h * ch);
            s = c.Compress(copy.data(), w, h, (uint8_t)ch, p);
        }
        if (times) times[0] = now_s() - t0;
        if (s.size() > out_cap) return -2;
        std::memcpy(out, s.data(), s.size());
        return (long long)s.size();
    } catch (const std::exception& e) {
        std::fprintf(stderr, "ref_compress: %s\n", e.what());
        return -1;
    }
}

// The reference's DecompressToBCn (:287-361) on any .yupt2 stream; info24 receives the 24-byte TextureInfo.
long long ref_decompress_to_bcn(const uint8_t* stream, size_t n, uint8_t* out, size_t out_cap, uint8_t* info24, int threads) {
    try {
        VQBCnCompressor c;
        TextureInfo info;
        const std::vector<uint8_t> v(stream, stream + n);
        const std::vector<uint8_t> bc = c.DecompressToBCn(v, info, threads);
        if (info24) std::memcpy(info24, &info, sizeof(TextureInfo));
        if (bc.size() > out_cap) return -2;
        std::memcpy(out, bc.data(), bc.size());
        return (long long)bc.size();
    } catch (const std::exception& e) {
        std::fprintf(stderr, "ref_decompress_to_bcn: %s\n", e.what());
        return -1;
    }
}
#endif

}  // extern "C"
