// Test infrastructure (NOT product code): the block encoders of yuptexture_b200/csrc/bcn_encode.cuh compiled as host
// C++ and run over an image with the texel fetch of bcn_encode_kernel, so that the CPU suite (no GPU in the build
// container) can gate the encoders' decoded PSNR against Compressonator's.  The GPU tests run the real kernel and
// check that it produces exactly these bytes.
//   g++ -O2 -shared -fPIC -std=c++17 tests/cpp/bcn_encode_host.cpp -o tests/cpp/_build/libbcnenc_host.so
#include <cstdint>
#include <cstdlib>
#include <cstring>

#include "../../yuptexture_b200/csrc/bcn_encode.cuh"

using namespace yupvq::bcnenc;

extern "C" int bcnenc_host_encode(const uint8_t* img, uint32_t w, uint32_t h, uint32_t ch, int fmt, int swap_rb,
                                  uint32_t alpha_threshold, uint8_t* out) {
    if (fmt != 1 && fmt != 3 && fmt != 4 && fmt != 5) return 1;
    const uint32_t bx = (w + 3) / 4, by = (h + 3) / 4;
    const size_t bs = (fmt == 1 || fmt == 4) ? 8 : 16;
    for (size_t i = 0; i < (size_t)bx * by; ++i) {
        const uint32_t x0 = (uint32_t)(i % bx) * 4, y0 = (uint32_t)(i / bx) * 4;
        uint32_t px[16];
        for (uint32_t y = 0; y < 4; ++y)
            for (uint32_t x = 0; x < 4; ++x) {
                const uint8_t* s = img + ((size_t)(y0 + y < h ? y0 + y : h - 1) * w + (x0 + x < w ? x0 + x : w - 1)) * ch;
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
        uint32_t words[4];
        int v[16];
        if (fmt == 1) {
            const U2 c = encode_colour_block(px, true, ch == 4, alpha_threshold);
            words[0] = c.x; words[1] = c.y;
        } else if (fmt == 3) {
            for (int p = 0; p < 16; ++p) v[p] = (int)(px[p] >> 24);
            const U2 a = encode_scalar_block(v);
            const U2 c = encode_colour_block(px, false, false, 0);
            words[0] = a.x; words[1] = a.y; words[2] = c.x; words[3] = c.y;
        } else {
            for (int p = 0; p < 16; ++p) v[p] = (int)(px[p] & 255u);
            const U2 r = encode_scalar_block(v);
            words[0] = r.x; words[1] = r.y;
            if (fmt == 5) {
                for (int p = 0; p < 16; ++p) v[p] = ch == 2 ? (int)(px[p] >> 24) : (int)((px[p] >> 8) & 255u);
                const U2 g = encode_scalar_block(v);
                words[2] = g.x; words[3] = g.y;
            }
        }
        std::memcpy(out + i * bs, words, bs);
    }
    return 0;
}
