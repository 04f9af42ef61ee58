// libyupcodec_cmp.so -- the BCn block-codec plugin (include/yupcodec.h) backed by AMD Compressonator 4.3,
// the SDK the reference uses for its host-side BCn stages (src/bcn_compressor.cpp).  Host code only: it is
// not on the GPU hot path; it supplies the whole-texture pre-pass, the K centroid encodes
// and the BC6H decode so that VQBCnCompressor::Compress / CompressHDR produce the reference's codebook entries.
// The option block below is the one the reference passes (src/bcn_compressor.cpp:92-107 for LDR, :179-182 for
// HDR): the encoder is deterministic, so identical options give byte-identical blocks.
// Built by yuptexture_b200/host/Makefile when the SDK's static libraries are found ($CMP_SDK).
#include <compressonator.h>

#include <cstring>
#include <vector>

#include "../../include/yupcodec.h"

namespace {

CMP_FORMAT block_format(int bc) {
    static const CMP_FORMAT table[8] = {CMP_FORMAT_Unknown, CMP_FORMAT_BC1, CMP_FORMAT_BC2, CMP_FORMAT_BC3,
                                        CMP_FORMAT_BC4,     CMP_FORMAT_BC5, CMP_FORMAT_BC6H, CMP_FORMAT_BC7};
    return (bc >= 1 && bc <= 7) ? table[bc] : CMP_FORMAT_Unknown;
}

// uncompressed side: 8-bit or 32-bit float, 1..4 channels (3 float channels travel as "RGBE_32F", the only
// three-float layout the SDK accepts: src/bcn_compressor.cpp:30)
CMP_FORMAT pixel_format(uint32_t channels, bool is_float) {
    static const CMP_FORMAT u8[5] = {CMP_FORMAT_Unknown, CMP_FORMAT_R_8, CMP_FORMAT_RG_8, CMP_FORMAT_RGB_888, CMP_FORMAT_RGBA_8888};
    static const CMP_FORMAT f32[5] = {CMP_FORMAT_Unknown, CMP_FORMAT_R_32F, CMP_FORMAT_RG_32F, CMP_FORMAT_RGBE_32F, CMP_FORMAT_RGBA_32F};
    if (channels < 1 || channels > 4) return CMP_FORMAT_Unknown;
    return is_float ? f32[channels] : u8[channels];
}

size_t block_bytes(int bc) { return (bc == 1 || bc == 4) ? 8 : 16; }

CMP_Texture texture(uint32_t w, uint32_t h, CMP_FORMAT fmt, uint32_t pitch, const void* data, size_t bytes) {
    CMP_Texture t = {};
    t.dwSize = sizeof(CMP_Texture);
    t.dwWidth = w;
    t.dwHeight = h;
    t.dwPitch = pitch;
    t.format = fmt;
    t.dwDataSize = (CMP_DWORD)bytes;
    t.pData = static_cast<CMP_BYTE*>(const_cast<void*>(data));
    return t;
}

size_t bc_bytes(uint32_t w, uint32_t h, int bc) { return (size_t)((w + 3) / 4) * ((h + 3) / 4) * block_bytes(bc); }

}  // namespace

extern "C" {

__attribute__((visibility("default"))) const char* yupcodec_name(void) { return "AMD Compressonator 4.3 (static SDK libraries)"; }

__attribute__((visibility("default"))) int yupcodec_compress(const uint8_t* image, uint32_t width, uint32_t height,
                                                             uint32_t channels, int bc_format, int num_threads, float quality,
                                                             int alpha_threshold, int flip_rgb, uint8_t* out, size_t out_bytes) {
    const CMP_FORMAT src_fmt = pixel_format(channels, false), dst_fmt = block_format(bc_format);
    if (!image || !out || src_fmt == CMP_FORMAT_Unknown || dst_fmt == CMP_FORMAT_Unknown) return 1;
    if (out_bytes < bc_bytes(width, height, bc_format)) return 1;
    const size_t n = (size_t)width * height * channels;
    std::vector<uint8_t> swapped;
    if (flip_rgb && channels == 3) {   // red <-> blue on a private copy (the reference swaps the caller's buffer)
        swapped.assign(image, image + n);
        for (size_t i = 0; i + 2 < n; i += 3) std::swap(swapped[i], swapped[i + 2]);
        image = swapped.data();
    }
    CMP_Texture src = texture(width, height, src_fmt, width * channels, image, n);
    CMP_Texture dst = texture(width, height, dst_fmt, 0, out, bc_bytes(width, height, bc_format));
    CMP_CompressOptions opt = {};
    opt.dwSize = sizeof(opt);
    opt.fquality = quality;
    opt.dwnumThreads = (CMP_DWORD)num_threads;
    opt.bUseAdaptiveWeighting = true;
    opt.bUseRefinementSteps = true;
    opt.doDeltaEncodeBRLG = true;
    opt.doPreconditionBRLG = true;
    opt.doSwizzleBRLG = true;
    if (channels == 4 && bc_format == 1) {   // BC1 with alpha: punch-through
        opt.bDXT1UseAlpha = true;
        opt.nAlphaThreshold = (CMP_BYTE)alpha_threshold;
    }
    return CMP_ConvertTexture(&src, &dst, &opt, nullptr) == CMP_OK ? 0 : 2;
}

__attribute__((visibility("default"))) int yupcodec_compress_hdr(const float* image, uint32_t width, uint32_t height,
                                                                 uint32_t channels, int bc_format, int num_threads, float quality,
                                                                 uint8_t* out, size_t out_bytes) {
    if (!image || !out || out_bytes < bc_bytes(width, height, bc_format)) return 1;
    // BC6H / BC7 from three floats: the SDK wants RGBA, alpha = 1 (src/bcn_compressor.cpp:128-147)
    std::vector<float> rgba;
    uint32_t ch = channels;
    if (channels == 3 && (bc_format == 6 || bc_format == 7)) {
        rgba.resize((size_t)width * height * 4);
        for (size_t i = 0; i < (size_t)width * height; ++i) {
            rgba[i * 4 + 0] = image[i * 3 + 0];
            rgba[i * 4 + 1] = image[i * 3 + 1];
            rgba[i * 4 + 2] = image[i * 3 + 2];
            rgba[i * 4 + 3] = 1.0f;
        }
        image = rgba.data();
        ch = 4;
    }
    const CMP_FORMAT src_fmt = pixel_format(ch, true), dst_fmt = block_format(bc_format);
    if (src_fmt == CMP_FORMAT_Unknown || dst_fmt == CMP_FORMAT_Unknown) return 1;
    CMP_Texture src = texture(width, height, src_fmt, width * ch * (uint32_t)sizeof(float), image,
                              (size_t)width * height * ch * sizeof(float));
    CMP_Texture dst = texture(width, height, dst_fmt, 0, out, bc_bytes(width, height, bc_format));
    CMP_CompressOptions opt = {};
    opt.dwSize = sizeof(opt);
    opt.fquality = quality;
    opt.dwnumThreads = (CMP_DWORD)num_threads;
    return CMP_ConvertTexture(&src, &dst, &opt, nullptr) == CMP_OK ? 0 : 2;
}

__attribute__((visibility("default"))) int yupcodec_decompress_hdr(const uint8_t* bc, uint32_t width, uint32_t height,
                                                                   uint32_t channels, int bc_format, float* out) {
    const CMP_FORMAT src_fmt = block_format(bc_format);
    if (!bc || !out || src_fmt == CMP_FORMAT_Unknown || channels < 1 || channels > 4) return 1;
    // the SDK decodes BC6H to RGBA float; extra channels are dropped afterwards (src/bcn_compressor.cpp:246-299)
    std::vector<float> rgba((size_t)width * height * 4);
    CMP_Texture src = texture(width, height, src_fmt, 0, bc, bc_bytes(width, height, bc_format));
    CMP_Texture dst = texture(width, height, CMP_FORMAT_RGBA_32F, width * 4 * (uint32_t)sizeof(float), rgba.data(),
                              rgba.size() * sizeof(float));
    if (CMP_ConvertTexture(&src, &dst, nullptr, nullptr) != CMP_OK) return 2;
    if (channels == 4) {
        std::memcpy(out, rgba.data(), rgba.size() * sizeof(float));
    } else {
        for (size_t i = 0; i < (size_t)width * height; ++i)
            for (uint32_t c = 0; c < channels; ++c) out[i * channels + c] = rgba[i * 4 + c];
    }
    return 0;
}

}  // extern "C"
