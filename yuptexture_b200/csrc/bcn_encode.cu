// N4: whole-texture BCn encoders (BC1, BC3, BC4, BC5) on the GPU -- the pre-pass of VQBCnCompressor::Compress
// (reference src/vq_bcn_compressor.cpp:94 -> BCnCompressor::Compress, src/bcn_compressor.cpp:45-116, which hands the
// image to the closed-source Compressonator encoders: 5.3 s for a 4096^2 RGB texture).
//
// Compressonator's bit patterns cannot be reproduced (binary-only); what can be matched is the QUALITY of the blocks,
// measured the way the VQ stage sees them: PSNR of the decoded texels against the source.  Gated against Compressonator's
// own output: BC1 (RGB, and RGBA punch-through), BC3 and BC4 on the host-compiled build of this block code
// (tests/test_bcn_encode_cpu.py), BC1 again on the GPU at 4096^2 (tests/test_gpu_bcn_encode.py), and the kernel's bytes
// equal that host build's for all four formats.  BC5 has no gate of its own: it is two BC4 halves from the gated scalar
// encoder (the reference library aborts on 2-channel sources here, so there is nothing to compare with).
// One thread encodes one 4x4 block:
//   colour (BC1 / the colour half of BC3): principal axis of the 16 texels by power iteration on the covariance matrix,
//     endpoints = the extreme projections, then least-squares refits of the endpoints for the current selectors; each
//     candidate is scored with the DECODER's integer palette ((2a+b)/3 truncating, bcn_decode.cuh) and the best kept.
//     BC1 with an alpha channel: texels below the threshold select the transparent entry of the 3-colour mode
//     (CMP_CompressOptions::bDXT1UseAlpha / nAlphaThreshold, src/bcn_compressor.cpp:103-107).
//   scalar (BC4, both halves of BC5, the alpha half of BC3): least-squares refit of the two endpoints plus an exhaustive
//     +-2 search around them, in the 8-value mode and -- when the block touches 0 or 255 -- in the 6-value mode whose
//     palette holds 0 and 255 explicitly.
// HBM traffic: 16 * ch bytes read + 8 / 16 bytes written per block: the kernel is ALU-bound (about 3k integer operations
// per scalar block), a fraction of a millisecond at 4096^2.
#include "bcn_encode.cuh"
#include "common.cuh"
#include "kernels.h"

namespace yupvq {

namespace {

// One thread per block.  The image is row-major, channels interleaved, edge texels repeated for partial blocks.
// 1- and 2-channel sources expand to (v, v, v, 255) / (v, v, v, a) the way the VQ path's decode reads them back.
__global__ void __launch_bounds__(128) bcn_encode_kernel(const uint8_t* __restrict__ img, uint32_t w, uint32_t h, uint32_t ch, uint32_t bx,
                                                         uint32_t n_blocks, int fmt, int swap_rb, uint32_t alpha_threshold,
                                                         uint8_t* __restrict__ out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_blocks) return;
    const uint32_t x0 = (i % bx) * 4, y0 = (i / bx) * 4;
    uint32_t px[16];
#pragma unroll
    for (int y = 0; y < 4; ++y) {
        const size_t row = (size_t)min(y0 + (uint32_t)y, h - 1) * w;
#pragma unroll
        for (int x = 0; x < 4; ++x) {
            const uint8_t* s = img + (row + min(x0 + (uint32_t)x, w - 1)) * ch;
            uint32_t r, g, b, a = 255;
            if (ch >= 3) {
                r = s[swap_rb ? 2 : 0]; g = s[1]; b = s[swap_rb ? 0 : 2];
                if (ch == 4) a = s[3];
            } else {
                r = g = b = s[0];
                if (ch == 2) a = s[1];
            }
            px[y * 4 + x] = r | (g << 8) | (b << 16) | (a << 24);
        }
    }
    if (fmt == YUPVQ_FMT_BC1) {
        const bcnenc::U2 c = bcnenc::encode_colour_block(px, true, ch == 4, alpha_threshold);
        reinterpret_cast<uint2*>(out)[i] = make_uint2(c.x, c.y);
        return;
    }
    int v[16];
    if (fmt == YUPVQ_FMT_BC3) {
#pragma unroll
        for (int p = 0; p < 16; ++p) v[p] = (int)(px[p] >> 24);
        const bcnenc::U2 a = bcnenc::encode_scalar_block(v);
        const bcnenc::U2 c = bcnenc::encode_colour_block(px, false, false, 0);
        reinterpret_cast<uint4*>(out)[i] = make_uint4(a.x, a.y, c.x, c.y);
        return;
    }
    // BC4: the first channel; BC5: first channel, then the second (2-channel sources keep theirs in alpha)
#pragma unroll
    for (int p = 0; p < 16; ++p) v[p] = (int)(px[p] & 255u);
    const bcnenc::U2 r = bcnenc::encode_scalar_block(v);
    if (fmt == YUPVQ_FMT_BC4) {
        reinterpret_cast<uint2*>(out)[i] = make_uint2(r.x, r.y);
        return;
    }
#pragma unroll
    for (int p = 0; p < 16; ++p) v[p] = ch == 2 ? (int)(px[p] >> 24) : (int)((px[p] >> 8) & 255u);
    const bcnenc::U2 g = bcnenc::encode_scalar_block(v);
    reinterpret_cast<uint4*>(out)[i] = make_uint4(r.x, r.y, g.x, g.y);
}

}  // namespace

bool bcn_encode_supported(int fmt) {
    return fmt == YUPVQ_FMT_BC1 || fmt == YUPVQ_FMT_BC3 || fmt == YUPVQ_FMT_BC4 || fmt == YUPVQ_FMT_BC5;
}

cudaError_t launch_bcn_encode(const uint8_t* image, uint32_t width, uint32_t height, uint32_t channels, int fmt, int swap_rb,
                              uint32_t alpha_threshold, uint8_t* out, cudaStream_t s) {
    if (!bcn_encode_supported(fmt) || channels < 1 || channels > 4 || width == 0 || height == 0) return cudaErrorInvalidValue;
    const uint32_t bx = (width + 3) / 4, by = (height + 3) / 4;
    const uint64_t n = (uint64_t)bx * by;
    if (n > 0xffffffffull) return cudaErrorInvalidValue;
    bcn_encode_kernel<<<(unsigned)((n + 127) / 128), 128, 0, s>>>(image, width, height, channels, bx, (uint32_t)n, fmt, swap_rb, alpha_threshold, out);
    return cudaGetLastError();
}

}  // namespace yupvq
