/*
 * yupcodec.h -- C ABI of a BCn block-codec plugin for the host side of VQBCnCompressor.
 *
 * The VQ hot path runs on the GPU (include/yupvq.h).  Three host stages of the reference stay on the
 * host behind the same C++ API (last row): the whole-texture BCn encode
 * (src/vq_bcn_compressor.cpp:94 -> src/bcn_compressor.cpp:45), the K centroid encodes
 * (src/vq_encoder.cpp:248-254, :432-440, :614-620) and the BC6H decode of DecompressHDR
 * (src/bcn_compressor.cpp:229).  The reference does them with AMD Compressonator 4.3, a closed-source
 * SDK.  A shared library exporting the symbols below supplies those encoders to libyuptexture.so:
 *
 *     BCnCompressor::LoadCodecLibrary("/path/libyupcodec_cmp.so");      // explicit
 *     YUPTEXTURE_CODEC=/path/libyupcodec_cmp.so                          // environment
 *     libyupcodec_cmp.so next to libyuptexture.so                        // picked up on first use
 *
 * yuptexture_b200/host/codec_compressonator.cpp is the Compressonator-backed implementation (built when
 * the SDK's static libraries are available).  Without a plugin libyuptexture.so falls back to its small
 * built-in bounding-box encoders for BC1/3/4/5/7 and has no BC6H encoder.
 *
 * All functions return 0 on success.  Images are row-major, channels interleaved; `bc_format` takes the
 * BCFormat values of include/vq_bcn_types.h:27-36; outputs are caller-allocated
 * (ceil(w/4) * ceil(h/4) * blockSize bytes, or w * h * channels floats).
 */
#ifndef YUPCODEC_H
#define YUPCODEC_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* BCnCompressor::Compress, src/bcn_compressor.cpp:45-116.  The input is never modified: with flip_rgb
 * (3-channel input only) the plugin encodes a red/blue-swapped copy (the reference swaps in place). */
int yupcodec_compress(const uint8_t* image, uint32_t width, uint32_t height, uint32_t channels, int bc_format,
                      int num_threads, float quality, int alpha_threshold, int flip_rgb, uint8_t* out, size_t out_bytes);
/* BCnCompressor::CompressHDR, src/bcn_compressor.cpp:119-189. */
int yupcodec_compress_hdr(const float* image, uint32_t width, uint32_t height, uint32_t channels, int bc_format,
                          int num_threads, float quality, uint8_t* out, size_t out_bytes);
/* BCnCompressor::DecompressHDR, src/bcn_compressor.cpp:229-300 (out: width * height * channels floats). */
int yupcodec_decompress_hdr(const uint8_t* bc, uint32_t width, uint32_t height, uint32_t channels, int bc_format,
                            float* out);
const char* yupcodec_name(void);

#ifdef __cplusplus
}
#endif
#endif /* YUPCODEC_H */
