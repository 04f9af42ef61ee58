// Microbenchmark: throughput of byte-SAD formulations on sm_100a (per-SM lanes/clk).
#include <cstdio>
#include <cstdint>
#include <cuda_fp16.h>
template <int MODE>
__global__ void __launch_bounds__(1024) k(unsigned* out, unsigned seed, int iters, long long* cyc) {
    unsigned a0 = threadIdx.x * 2654435761u + seed, a1 = a0 * 3u + 1, a2 = a0 * 5u + 7, a3 = a0 * 7u + 3;
    unsigned s0 = seed * 11u, s1 = seed * 13u, s2 = seed * 17u, s3 = seed * 19u;
    unsigned acc0 = 0, acc1 = 0, acc2 = 0, acc3 = 0;
    __syncthreads();
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
        if (MODE == 0) {
            acc0 = __vsadu4(a0, s0) + acc0; acc1 = __vsadu4(a1, s1) + acc1; acc2 = __vsadu4(a2, s2) + acc2; acc3 = __vsadu4(a3, s3) + acc3;
        } else if (MODE == 1) {  // max/min + dp4a
            acc0 = __dp4a(__vmaxu4(a0, s0), 0x01010101u, acc0) - __dp4a(__vminu4(a0, s0), 0x01010101u, 0u);
            acc1 = __dp4a(__vmaxu4(a1, s1), 0x01010101u, acc1) - __dp4a(__vminu4(a1, s1), 0x01010101u, 0u);
            acc2 = __dp4a(__vmaxu4(a2, s2), 0x01010101u, acc2) - __dp4a(__vminu4(a2, s2), 0x01010101u, 0u);
            acc3 = __dp4a(__vmaxu4(a3, s3), 0x01010101u, acc3) - __dp4a(__vminu4(a3, s3), 0x01010101u, 0u);
        } else if (MODE == 2) {  // half2: bytes -> (1024 + b) via PRMT, |a - s| accumulated in half2
            const unsigned magic = 0x64006400u;
            __half2 h0 = __halves2half2(__ushort_as_half(0), __ushort_as_half(0)), h1 = h0;
            unsigned as_[4] = {a0, a1, a2, a3}, ss_[4] = {s0, s1, s2, s3};
#pragma unroll
            for (int w = 0; w < 4; ++w) {
                unsigned lo = __byte_perm(as_[w], magic, 0x7170), hi = __byte_perm(as_[w], magic, 0x7372);
                unsigned slo = __byte_perm(ss_[w], magic, 0x7170), shi = __byte_perm(ss_[w], magic, 0x7372);
                h0 = __hadd2(h0, __habs2(__hsub2(*(__half2*)&lo, *(__half2*)&slo)));
                h1 = __hadd2(h1, __habs2(__hsub2(*(__half2*)&hi, *(__half2*)&shi)));
            }
            __half2 h = __hadd2(h0, h1);
            acc0 += (unsigned)__half2float(__low2half(h)) + (unsigned)__half2float(__high2half(h));
        } else if (MODE == 3) {  // plain IMAD reference (full-rate integer)
            acc0 = a0 * s0 + acc0; acc1 = a1 * s1 + acc1; acc2 = a2 * s2 + acc2; acc3 = a3 * s3 + acc3;
        }
        a0 += acc3; a1 ^= acc0; a2 += 0x01010101u; a3 ^= acc1;  // keep the compiler honest (2 IADD/LOP per 4 ops)
    }
    long long t1 = clock64();
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc0 + acc1 + acc2 + acc3;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
int main() {
    unsigned* out; long long* cyc; cudaMalloc(&out, 148 * 1024 * 4); cudaMallocManaged(&cyc, 8);
    const int iters = 4096;
    const char* names[] = {"VABSDIFF4.ACC", "vmax/vmin+dp4a", "half2 PRMT", "IMAD"};
    for (int m = 0; m < 4; ++m) {
        for (int rep = 0; rep < 2; ++rep) {
            if (m == 0) k<0><<<148, 1024>>>(out, 12345u + rep, iters, cyc);
            if (m == 1) k<1><<<148, 1024>>>(out, 12345u + rep, iters, cyc);
            if (m == 2) k<2><<<148, 1024>>>(out, 12345u + rep, iters, cyc);
            if (m == 3) k<3><<<148, 1024>>>(out, 12345u + rep, iters, cyc);
            cudaDeviceSynchronize();
        }
        double words = 4.0 * iters * 1024;  // 32-bit word SADs per SM
        printf("%-16s %lld cycles: %.2f word-SADs/clk/SM (%.1f bytes/clk/SM)\n", names[m], *cyc, words / *cyc, 4 * words / *cyc);
    }
    return 0;
}
