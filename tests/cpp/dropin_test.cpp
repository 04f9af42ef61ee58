// Drop-in check of the container and the host-side codec stages against the UNMODIFIED reference
// (oracle/_ref/libyupref.so, loaded with dlopen: test side only -- the product never links it).
//   * VQBCnCompressor::Compress on BASELINE config 1 (1024x1024 RGB, BC1, quality 0.8, zstd on) through the
//     Compressonator codec plugin; the stream is decoded by the REFERENCE's DecompressToBCn (src/vq_bcn_compressor.cpp:287-361)
//     and by ours: same bytes.  A stream written by the reference's Compress is decoded by ours: same bytes as the reference's.
//   * VQCodebook.entries equal the reference's CompressSingleBlock[HDR] on the same pixel centroids, byte for byte
//     (src/vq_encoder.cpp:248-254, :432-440, :614-620), for BC1, BC3, BC7 and BC6H.
//   * dumps pixel blocks / centroids / indices for tests/test_gpu_dropin.py (PSNR within 0.05 dB of the reference).
// usage: dropin_test <out_dir> <libyupref.so> <rgb_1024.u8>
#include <dlfcn.h>

#include <algorithm>
#include <cmath>
#include <chrono>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <stdexcept>
#include <string>
#include <thread>
#include <vector>

#include "vq_bcn_compressor.h"
#include "vq_encoder.h"

static int failures = 0;
#define CHECK(cond)                                                     \
    do {                                                                \
        if (!(cond)) {                                                  \
            std::printf("FAIL %s:%d: %s\n", __FILE__, __LINE__, #cond); \
            ++failures;                                                 \
        }                                                               \
    } while (0)

template <typename T>
static void dump(const std::string& path, const std::vector<T>& v) {
    std::ofstream f(path, std::ios::binary);
    f.write(reinterpret_cast<const char*>(v.data()), (std::streamsize)(v.size() * sizeof(T)));
}
template <typename T>
static std::vector<T> flat(const std::vector<std::vector<T>>& vv) {
    std::vector<T> out;
    for (const auto& v : vv) out.insert(out.end(), v.begin(), v.end());
    return out;
}
static std::vector<uint8_t> read_file(const std::string& path) {
    std::ifstream f(path, std::ios::binary);
    return std::vector<uint8_t>((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>());
}

// reference entry points (oracle/ref_harness.cpp)
typedef long long (*ref_compress_fn)(const void*, int, uint32_t, uint32_t, uint32_t, const void*, uint8_t*, size_t, double*);
typedef long long (*ref_decompress_fn)(const uint8_t*, size_t, uint8_t*, size_t, uint8_t*, int);
typedef size_t (*ref_single_fn)(const uint8_t*, uint32_t, int, float, int, uint8_t*);
typedef size_t (*ref_single_hdr_fn)(const float*, uint32_t, int, float, int, uint8_t*);
typedef int (*ref_decode_blocks_fn)(const uint8_t*, uint32_t, uint32_t, uint32_t, int, uint8_t*);

struct Ref {
    ref_compress_fn compress = nullptr;
    ref_decompress_fn decompress_to_bcn = nullptr;
    ref_single_fn single = nullptr;
    ref_single_hdr_fn single_hdr = nullptr;
    ref_decode_blocks_fn decode_blocks = nullptr;
};

static Ref load_ref(const char* path) {
    void* h = dlopen(path, RTLD_NOW | RTLD_LOCAL | RTLD_DEEPBIND);   // its classes share names with ours: bind inside first
    if (!h) throw std::runtime_error(std::string("cannot load ") + path + ": " + dlerror());
    Ref r;
    r.compress = (ref_compress_fn)dlsym(h, "ref_compress");
    r.decompress_to_bcn = (ref_decompress_fn)dlsym(h, "ref_decompress_to_bcn");
    r.single = (ref_single_fn)dlsym(h, "ref_encode_single_block");
    r.single_hdr = (ref_single_hdr_fn)dlsym(h, "ref_encode_single_block_hdr");
    r.decode_blocks = (ref_decode_blocks_fn)dlsym(h, "ref_bcn_decode_blocks");
    if (!r.compress || !r.decompress_to_bcn || !r.single || !r.single_hdr || !r.decode_blocks)
        throw std::runtime_error("libyupref.so lacks the container entry points (rebuild with oracle/build_ref.sh)");
    return r;
}

static bool header_equal(const TextureInfo& a, const uint8_t* ref24) {   // padding bytes 13..15 are uninitialised in the reference (Q9)
    uint8_t mine[24];
    std::memcpy(mine, &a, 24);
    return std::memcmp(mine, ref24, 13) == 0 && std::memcmp(mine + 16, ref24 + 16, 8) == 0;
}

// progress line with the wall time since start (the test log shows where the minutes go)
static void stamp(const char* what) {
    static const auto t0 = std::chrono::steady_clock::now();
    std::printf("[%7.1f s] %s\n", std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count(), what);
    std::fflush(stdout);
}

// Runs one LDR case: Compress (seed 42), both decoders on our stream, entries vs the reference's single-block encoder.
static void ldr_case(const Ref& ref, const std::string& out, const std::string& tag, const std::vector<uint8_t>& img, uint32_t W,
                     uint32_t H, uint8_t CH, CompressionParams p, bool reference_stream = true) {
    stamp((tag + ": start").c_str());
    VQBCnCompressor comp;
    comp.SetSeed(42);
    const auto stream = comp.Compress(img.data(), W, H, CH, p);
    stamp((tag + ": our Compress done").c_str());
    const VQBCnCompressor::Timings t = comp.LastTimings();
    std::printf("TIMINGS %s {\"bcn_encode_ms\": %.3f, \"vq_ms\": %.3f, \"centroid_encode_ms\": %.3f, \"zstd_ms\": %.3f, \"total_ms\": %.3f, "
                "\"stream_bytes\": %zu}\n", tag.c_str(), t.bcnEncodeMs, t.vqMs, t.centroidEncodeMs, t.zstdMs, t.totalMs, stream.size());
    const size_t bs = BCBlockSize::GetSize(p.bcFormat);
    const size_t n = (size_t)((W + 3) / 4) * ((H + 3) / 4);
    // (1) our stream through both decoders
    TextureInfo info;
    const auto bc_ours = comp.DecompressToBCn(stream, info);
    std::vector<uint8_t> bc_ref(n * bs);
    uint8_t info_ref[24];
    const long long got = ref.decompress_to_bcn(stream.data(), stream.size(), bc_ref.data(), bc_ref.size(), info_ref, p.numThreads);
    CHECK(got == (long long)(n * bs));
    CHECK(bc_ours.size() == n * bs && bc_ours == bc_ref);
    CHECK(header_equal(info, info_ref));
    CHECK(info.storedCodebookEntries == (1u << (4 + (uint32_t)std::lround(p.quality * 6))));
    CHECK(info.compressionFlags == (p.useZstd ? COMPRESSION_FLAGS_DEFAULT : COMPRESSION_FLAGS_ZSTD_BYPASSED));
    stamp((tag + ": both decoders done").c_str());
    // (2) the same path by hand, to get at the pixel centroids: entries == the reference's CompressSingleBlock
    BCnCompressor codec;
    const auto bc_in = codec.Compress(img.data(), W, H, CH, p.bcFormat, p.numThreads, p.bcQuality, p.alphaThreshold, false);
    CHECK(bc_in.size() == n * bs);
    CompressionConfig cfg;
    cfg.metric = p.vq_Metric;
    cfg.maxIterations = p.vq_maxIterations;
    cfg.SetQuality(p.quality);
    VQEncoder enc(cfg);
    enc.SetFormat(p.bcFormat);
    enc.SetSeed(42);
    std::vector<std::vector<uint8_t>> cent;
    std::vector<uint32_t> idx;
    const VQCodebook cb = enc.EncodeBCn(bc_in.data(), n, CH, cent, idx, p);
    stamp((tag + ": EncodeBCn by hand done").c_str());
    size_t same_entries = 0;
    for (size_t k = 0; k < cent.size(); ++k) {
        uint8_t e[16] = {0};
        const size_t sz = ref.single(cent[k].data(), CH, (int)p.bcFormat, p.bcQuality, 1, e);   // one thread: a 4x4 encode does not depend on it and Compressonator spins up its pool per call
        same_entries += sz == bs && cb.entries[k].size() == bs && std::memcmp(e, cb.entries[k].data(), bs) == 0;
    }
    std::printf("%s: %zu / %zu codebook entries byte-identical to the reference's CompressSingleBlock\n", tag.c_str(), same_entries, cent.size());
    CHECK(same_entries == cent.size());
    stamp((tag + ": reference single-block encodes done").c_str());
    // the stream holds exactly these entries and indices (same seed): DecompressToBCn == codebook[indices]
    bool same = idx.size() == n;
    for (size_t i = 0; same && i < n; ++i) same = std::memcmp(&bc_ours[i * bs], cb.entries[idx[i]].data(), bs) == 0;
    CHECK(same);
    // (3) dumps for the PSNR gate: the decoded pixel blocks the VQ stage saw (the REFERENCE's decoder), centroids, indices
    std::vector<uint8_t> px(n * 16 * CH);
    CHECK(ref.decode_blocks(bc_in.data(), W, H, CH, (int)p.bcFormat, px.data()) == 0);
    dump(out + "/" + tag + "_blocks.u8", px);
    dump(out + "/" + tag + "_centroids.u8", flat(cent));
    dump(out + "/" + tag + "_indices.u32", idx);
    // (4) a stream written by the reference's own Compress (its seed is random) decodes to the same bytes with both decoders
    std::vector<uint8_t> rs(stream.size() * 2 + (1u << 20));
    // (the container does not depend on how long the reference's CPU clustering ran: one Lloyd pass on a 5 % sample keeps
    // this leg to seconds instead of minutes of host time per case)
    CompressionParams pr = p;
    pr.vq_maxIterations = 1;
    pr.vq_FastModeSampleRatio = 0.05f;
    pr.quality = 0.0f;   // K = 16: the reference hands numThreads to each of its K single-block Compressonator calls, which
                         // spin up a thread pool apiece (numThreads itself must stay at the core count: the reference sizes
                         // per-thread buffers by it while OpenMP may run more threads)
    stamp((tag + ": dumps done").c_str());
    // BC7: the reference decodes block by block through Compressonator (src/vq_bcn_compressor.cpp:142-146), 55 ms per 4x4 call
    // with its per-call thread pool -- 225 s for these 4096 blocks, measured; the container layout is format-independent, so
    // the BC1 and BC3 cases carry this leg
    if (!reference_stream) return;
    const long long rn = ref.compress(img.data(), 0, W, H, CH, &pr, rs.data(), rs.size(), nullptr);
    stamp((tag + ": reference Compress done").c_str());
    CHECK(rn > 24);
    if (rn > 24) {
        rs.resize((size_t)rn);
        TextureInfo i2;
        const auto a = comp.DecompressToBCn(rs, i2);
        std::vector<uint8_t> b(n * bs);
        uint8_t i2_ref[24];
        CHECK(ref.decompress_to_bcn(rs.data(), rs.size(), b.data(), b.size(), i2_ref, p.numThreads) == (long long)(n * bs));
        CHECK(a == b && header_equal(i2, i2_ref));
    }
}

int main(int argc, char** argv) {
    if (argc < 4) {
        std::printf("usage: dropin_test <out_dir> <libyupref.so> <rgb_1024.u8>\n");
        return 2;
    }
    const std::string out = argv[1];
    try {
        const Ref ref = load_ref(argv[2]);
        std::printf("codec: %s\n", BCnCompressor::CodecName());
        CHECK(std::string(BCnCompressor::CodecName()).find("Compressonator") != std::string::npos);
        const int threads = std::max(1u, std::thread::hardware_concurrency());

        // ---- BASELINE config 1: 1024x1024 RGB, BC1, quality 0.8 (K = 512), zstd on, numThreads = host cores ----
        const auto img = read_file(argv[3]);
        CHECK(img.size() == 1024u * 1024u * 3u);
        CompressionParams p;
        p.bcFormat = BCFormat::BC1;
        p.quality = 0.8f;
        p.useZstd = true;
        p.numThreads = threads;
        ldr_case(ref, out, "c1_bc1", img, 1024, 1024, 3, p);

        // ---- RGBA with alpha through BC3 and Compressonator-encoded BC7 (config 3's formats, on a crop) ----
        std::vector<uint8_t> rgba((size_t)256 * 256 * 4);
        for (uint32_t y = 0; y < 256; ++y)
            for (uint32_t x = 0; x < 256; ++x) {
                const uint8_t* s = &img[((size_t)y * 1024 + x) * 3];
                uint8_t* d = &rgba[((size_t)y * 256 + x) * 4];
                d[0] = s[0]; d[1] = s[1]; d[2] = s[2];
                const double r = std::hypot(x - 128.0, y - 128.0) / 181.0;
                d[3] = ((x * 7 + y * 13) % 19 == 0) ? 0 : (uint8_t)std::lround(255.0 * (1.0 - r));
            }
        CompressionParams p3 = p;
        p3.bcFormat = BCFormat::BC3;
        p3.quality = 0.5f;   // K = 128 on 4096 blocks
        ldr_case(ref, out, "c3_bc3", rgba, 256, 256, 4, p3);
        CompressionParams p7 = p3;
        p7.bcFormat = BCFormat::BC7;
        p7.bcQuality = 0.25f;
        ldr_case(ref, out, "c3_bc7", rgba, 256, 256, 4, p7, false);
        // the example's BC7 preset uses the SAD metric (example/example.cpp:271-277)
        CompressionParams p7s = p7;
        p7s.vq_Metric = DistanceMetric::SAD_SIMD;
        ldr_case(ref, out, "c3_bc7_sad", rgba, 256, 256, 4, p7s, false);

        // ---- HDR: BC6H through CompressHDR (the example's HDR preset at a small size) ----
        const uint32_t HW = 256, HH = 128;
        std::vector<float> hdr((size_t)HW * HH * 3);
        uint32_t s = 4242;
        for (uint32_t y = 0; y < HH; ++y)
            for (uint32_t x = 0; x < HW; ++x)
                for (uint32_t c = 0; c < 3; ++c) {
                    s = s * 1664525u + 1013904223u;
                    float v = 0.1f + 0.6f * (float)(((x / 16) * 37 + (y / 16) * 101 + c * 17) % 64) / 64.0f + 0.02f * (float)((s >> 8) / 16777216.0);
                    if (((x * 31 + y * 17) % 509) == 0) v = 20.0f + 80.0f * (float)((s >> 8) / 16777216.0);
                    hdr[((size_t)y * HW + x) * 3 + c] = v;
                }
        CompressionParams ph;
        ph.bcFormat = BCFormat::BC6H;
        ph.quality = 0.5f;
        ph.bcQuality = 0.25f;
        ph.vq_maxIterations = 16;
        ph.numThreads = threads;
        VQBCnCompressor comp;
        comp.SetSeed(42);
        stamp("hdr: start");
        const auto hs = comp.CompressHDR(hdr.data(), HW, HH, 3, ph);
        stamp("hdr: our CompressHDR done");
        TextureInfo hi;
        const auto hbc = comp.DecompressToBCn(hs, hi);
        const size_t hn = (size_t)(HW / 4) * (HH / 4);
        CHECK((hi.compressionFlags & COMPRESSION_FLAGS_IS_HDR) && hi.storedCodebookEntries == 128 && hbc.size() == hn * 16);
        std::vector<uint8_t> hbc_ref(hn * 16);
        uint8_t hinfo_ref[24];
        CHECK(ref.decompress_to_bcn(hs.data(), hs.size(), hbc_ref.data(), hbc_ref.size(), hinfo_ref, threads) == (long long)(hn * 16));
        CHECK(hbc == hbc_ref && header_equal(hi, hinfo_ref));
        const auto hpix = comp.DecompressHDR(hs, hi);   // BC6H decode through the plugin
        CHECK(hpix.size() == (size_t)HW * HH * 3);
        CompressionConfig hc;
        hc.maxIterations = ph.vq_maxIterations;
        hc.SetQuality(ph.quality);
        VQEncoder henc(hc);
        henc.SetFormat(BCFormat::BC6H);
        henc.SetSeed(42);
        std::vector<std::vector<float>> hcent;
        std::vector<uint32_t> hidx;
        const VQCodebook hcb = henc.EncodeHDR(hdr.data(), HW, HH, 3, hcent, hidx, ph);
        stamp("hdr: EncodeHDR by hand done");
        size_t same_h = 0;
        for (size_t k = 0; k < hcent.size(); ++k) {
            uint8_t e[16] = {0};
            const size_t sz = ref.single_hdr(hcent[k].data(), 3, 6, ph.bcQuality, 1, e);
            same_h += sz == 16 && hcb.entries[k].size() == 16 && std::memcmp(e, hcb.entries[k].data(), 16) == 0;
        }
        std::printf("hdr_bc6h: %zu / %zu codebook entries byte-identical to the reference's CompressSingleBlockHDR\n", same_h, hcent.size());
        CHECK(same_h == hcent.size());
        bool hsame = hidx.size() == hn;
        for (size_t i = 0; hsame && i < hn; ++i) hsame = std::memcmp(&hbc[i * 16], hcb.entries[hidx[i]].data(), 16) == 0;
        CHECK(hsame);
        dump(out + "/hdr_image.f32", hdr);
        dump(out + "/hdr_centroids.f32", flat(hcent));
        dump(out + "/hdr_indices.u32", hidx);
    } catch (const std::exception& e) {
        std::printf("FAIL exception: %s\n", e.what());
        return 2;
    }
    std::printf(failures ? "DROPIN FAIL (%d)\n" : "DROPIN PASS\n", failures);
    return failures ? 1 : 0;
}
