// Exercises the drop-in C++ API (include/vq_bcn_compressor.h, include/vq_encoder.h) the way the
// reference's example/example.cpp does, and dumps raw arrays for tests/test_gpu_host_api.py to check
// against the oracle.  usage: host_api_test <out_dir>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <iostream>
#include <stdexcept>
#include <string>

#include "vq_bcn_compressor.h"
#include "vq_encoder.h"

static int failures = 0;
#define CHECK(cond)                                                       \
    do {                                                                  \
        if (!(cond)) {                                                    \
            std::printf("FAIL %s:%d: %s\n", __FILE__, __LINE__, #cond);   \
            ++failures;                                                   \
        }                                                                 \
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

static std::vector<uint8_t> make_rgb(uint32_t w, uint32_t h, uint32_t ch) {
    std::vector<uint8_t> img((size_t)w * h * ch);
    uint32_t s = 12345;
    for (uint32_t y = 0; y < h; ++y)
        for (uint32_t x = 0; x < w; ++x)
            for (uint32_t c = 0; c < ch; ++c) {
                s = s * 1664525u + 1013904223u;
                const double v = 0.5 + 0.4 * std::sin(0.031 * x * (c + 1) + 0.017 * y * (3 - c % 3)) + 0.07 * ((s >> 8) / 16777216.0 - 0.5);
                img[((size_t)y * w + x) * ch + c] = (uint8_t)std::lround(255.0 * std::min(1.0, std::max(0.0, v)));
            }
    return img;
}

static std::vector<std::vector<uint8_t>> tile(const std::vector<uint8_t>& img, uint32_t w, uint32_t h, uint32_t ch) {
    const uint32_t bx = (w + 3) / 4, by = (h + 3) / 4;
    std::vector<std::vector<uint8_t>> blocks((size_t)bx * by, std::vector<uint8_t>(16 * ch));
    for (uint32_t b = 0; b < bx * by; ++b)
        for (uint32_t p = 0; p < 16; ++p) {
            const uint32_t x = std::min((b % bx) * 4 + (p & 3), w - 1), y = std::min((b / bx) * 4 + (p >> 2), h - 1);
            std::memcpy(&blocks[b][p * ch], &img[((size_t)y * w + x) * ch], ch);
        }
    return blocks;
}

int main(int argc, char** argv) {
    const std::string out = argc > 1 ? argv[1] : ".";
    try {
        // ---- LDR: Compress -> DecompressToBCn round trip (reference README.md:45-55 usage) ----
        const uint32_t W = 256, H = 192, CH = 3;
        const auto img = make_rgb(W, H, CH);
        CompressionParams p;
        p.bcFormat = BCFormat::BC1;
        p.quality = 0.5f;  // K = 128
        p.vq_maxIterations = 8;
        p.numThreads = 4;
        VQBCnCompressor comp;
        comp.SetSeed(42);
        const auto stream = comp.Compress(img.data(), W, H, CH, p);
        TextureInfo info;
        const auto bc = comp.DecompressToBCn(stream, info);
        CHECK(info.width == W && info.height == H && info.format == BCFormat::BC1 && info.originalChannelCount == CH);
        CHECK(info.storedCodebookEntries == 128 && info.compressionFlags == COMPRESSION_FLAGS_DEFAULT);
        CHECK(bc.size() == info.GetTotalBlocks() * 8);
        const auto rgb = comp.Decompress(stream, info);
        CHECK(rgb.size() == (size_t)W * H * CH);

        // the same path by hand: BCn encode, fused device VQ, codebook gather
        BCnCompressor codec;
        const auto bc_in = codec.Compress(img.data(), W, H, CH, BCFormat::BC1, 4);
        CompressionConfig cfg;
        cfg.maxIterations = 8;
        cfg.SetQuality(0.5f);
        CHECK(cfg.codebookSize == 128);
        CompressionConfig q8(0.8f);
        CHECK(q8.codebookSize == 512);
        VQEncoder enc(cfg);
        enc.SetFormat(BCFormat::BC1);
        enc.SetSeed(42);
        std::vector<std::vector<uint8_t>> cent;
        std::vector<uint32_t> idx;
        const VQCodebook cb = enc.EncodeBCn(bc_in.data(), bc_in.size() / 8, CH, cent, idx, p);
        CHECK(cb.entries.size() == 128 && cb.blockSize == 8 && cent.size() == 128 && cent[0].size() == 48);
        bool same = idx.size() * 8 == bc.size();
        for (size_t i = 0; same && i < idx.size(); ++i) same = std::memcmp(&bc[i * 8], cb.entries[idx[i]].data(), 8) == 0;
        CHECK(same);  // DecompressToBCn == codebook[indices]

        // the reference's two-call form on decoded pixel blocks gives the same centroids and indices
        const auto decoded = codec.Decompress(bc_in.data(), W, H, CH, BCFormat::BC1);
        const auto blocks = tile(decoded, W, H, CH);
        std::vector<std::vector<uint8_t>> cent2;
        VQEncoder enc2(cfg);
        enc2.SetFormat(BCFormat::BC1);
        enc2.SetSeed(42);
        const VQCodebook cb2 = enc2.BuildCodebook(blocks, CH, cent2, p);
        const auto idx2 = enc2.QuantizeBlocks(blocks, CH, cent2, p);
        CHECK(cent2 == cent && idx2 == idx && cb2.entries == cb.entries);
        dump(out + "/ldr_blocks.u8", flat(blocks));
        dump(out + "/ldr_centroids.u8", flat(cent));
        dump(out + "/ldr_indices.u32", idx);

        // zstd off, VQ off, tiny textures: bypass flags (src/vq_bcn_compressor.cpp:97-125)
        CompressionParams raw = p;
        raw.useZstd = false;
        const auto s2 = comp.Compress(img.data(), W, H, CH, raw);
        CHECK(s2.size() == sizeof(TextureInfo) + 128 * 8 + idx.size() * 4);
        TextureInfo i2;
        CHECK(comp.DecompressToBCn(s2, i2) == bc && (i2.compressionFlags & COMPRESSION_FLAGS_ZSTD_BYPASSED));
        const auto small = make_rgb(16, 16, 3);
        const auto s3 = comp.Compress(small.data(), 16, 16, 3, p);
        TextureInfo i3;
        const auto bc3 = comp.DecompressToBCn(s3, i3);
        CHECK(i3.compressionFlags == (COMPRESSION_FLAGS_VQ_BYPASSED | COMPRESSION_FLAGS_ZSTD_BYPASSED) && bc3.size() == 16 * 8);

        // vq_Metric == SAD (the example's BC4 / BC5 / BC7 presets, example/example.cpp:260-277): all-integer path
        {
            CompressionParams ps = p;
            ps.vq_Metric = DistanceMetric::SAD_SIMD;
            CompressionConfig cs;
            cs.maxIterations = 8;
            cs.SetQuality(0.5f);
            cs.metric = DistanceMetric::SAD_SIMD;
            VQEncoder es(cs);
            es.SetFormat(BCFormat::BC1);
            es.SetSeed(42);
            std::vector<std::vector<uint8_t>> cs_cent;
            const VQCodebook cbs = es.BuildCodebook(blocks, CH, cs_cent, ps);
            const auto is = es.QuantizeBlocks(blocks, CH, cs_cent, ps);
            CHECK(cbs.entries.size() == 128 && cs_cent.size() == 128 && is.size() == blocks.size());
            dump(out + "/sad_centroids.u8", flat(cs_cent));
            dump(out + "/sad_indices.u32", is);
            const auto ss = comp.Compress(img.data(), W, H, CH, ps);
            TextureInfo isad;
            CHECK(comp.DecompressToBCn(ss, isad).size() == bc.size() && isad.storedCodebookEntries == 128);
        }

        // RGBA through BC3 and BC7
        const auto rgba = make_rgb(128, 128, 4);
        for (BCFormat f : {BCFormat::BC3, BCFormat::BC7}) {
            CompressionParams pa = p;
            pa.bcFormat = f;
            const auto sa = comp.Compress(rgba.data(), 128, 128, 4, pa);
            TextureInfo ia;
            const auto px = comp.Decompress(sa, ia);
            CHECK(ia.format == f && px.size() == 128 * 128 * 4);
        }

        // ---- HDR ----
        const uint32_t HW = 128, HH = 64;
        std::vector<float> hdr((size_t)HW * HH * 3);
        uint32_t s = 777;
        for (auto& v : hdr) {
            s = s * 1664525u + 1013904223u;
            v = 0.1f + 0.6f * (float)((s >> 8) / 16777216.0);
            if ((s & 1023) == 0) v *= 60.0f;
        }
        CompressionConfig hc(1.0f);
        CHECK(hc.codebookSize == 1024);
        hc.maxIterations = 6;
        VQEncoder henc(hc);
        henc.SetFormat(BCFormat::BC6H);
        henc.SetSeed(7);
        std::vector<std::vector<float>> hblocks((size_t)(HW / 4) * (HH / 4), std::vector<float>(48));
        for (uint32_t b = 0; b < hblocks.size(); ++b)
            for (uint32_t px = 0; px < 16; ++px)
                std::memcpy(&hblocks[b][px * 3], &hdr[((size_t)((b / (HW / 4)) * 4 + (px >> 2)) * HW + (b % (HW / 4)) * 4 + (px & 3)) * 3], 12);
        std::vector<std::vector<float>> hcent;
        bool threw = false;
        try {
            henc.BuildCodebookHDR(hblocks, 3, hcent, p);   // centroids are produced, the BC6H entry encode has no codec
        } catch (const std::runtime_error& e) {
            threw = std::string(e.what()).find("HDR block encoder") != std::string::npos;
        }
        CHECK(threw && hcent.size() == 1024);
        const auto hidx = henc.QuantizeBlocksHDR(hblocks, 3, hcent, p);
        CHECK(hidx.size() == hblocks.size());
        dump(out + "/hdr_blocks.f32", flat(hblocks));
        dump(out + "/hdr_centroids.f32", flat(hcent));
        dump(out + "/hdr_indices.u32", hidx);
        // with a codec installed the HDR container path works end to end
        BCnCompressor::Codec codec_hooks;
        codec_hooks.compressHDR = [](const float*, uint32_t w, uint32_t h, uint32_t, BCFormat, int, float) {
            return std::vector<uint8_t>((size_t)((w + 3) / 4) * ((h + 3) / 4) * 16, 0xAB);
        };
        BCnCompressor::SetCodec(codec_hooks);
        CompressionParams hp;
        hp.bcFormat = BCFormat::BC6H;
        hp.quality = 0.5f;
        hp.vq_maxIterations = 4;
        const auto hs = comp.CompressHDR(hdr.data(), HW, HH, 3, hp);
        TextureInfo hi;
        const auto hbc = comp.DecompressToBCn(hs, hi);
        CHECK((hi.compressionFlags & COMPRESSION_FLAGS_IS_HDR) && hi.storedCodebookEntries == 128 && hbc.size() == hblocks.size() * 16);
        BCnCompressor::SetCodec(BCnCompressor::Codec());

        // errors keep the reference's convention: std::runtime_error
        threw = false;
        try {
            TextureInfo t;
            comp.DecompressToBCn(std::vector<uint8_t>(5), t);
        } catch (const std::runtime_error&) {
            threw = true;
        }
        CHECK(threw);
    } catch (const std::exception& e) {
        std::printf("FAIL exception: %s\n", e.what());
        return 2;
    }
    std::printf(failures ? "HOST_API FAIL (%d)\n" : "HOST_API PASS\n", failures);
    return failures ? 1 : 0;
}
