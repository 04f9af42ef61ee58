// Issue rate of the integer byte-SIMD instructions the seeding / SAD kernels lean on, per SM per clock.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/vabsdiff_rate tools/micro/vabsdiff_rate.cu
#include <cstdio>
#include <cuda_runtime.h>

template <int OP>
__global__ void __launch_bounds__(1024, 1) rate_kernel(unsigned* out, unsigned a0, unsigned b0, int iters, long long* cyc) {
    unsigned a[8], acc[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) { a[i] = a0 + threadIdx.x * 0x01030507u + i; acc[i] = i; }
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (OP == 0) acc[i] = __vsadu4(a[i], b0 + it) + acc[i];          // VABSDIFF4.U8.ACC
            if (OP == 1) acc[i] = __byte_perm(acc[i], a[i], b0 + it);        // PRMT
            if (OP == 2) acc[i] = __dp4a(a[i], b0 + it, acc[i]);             // IDP.4A
            if (OP == 3) acc[i] = __vminu4(a[i] + it, acc[i]);               // VIMNMX? / emulated
            if (OP == 4) acc[i] = (acc[i] ^ a[i]) + (b0 + it);               // LOP3 + IADD3 (reference: plain ALU)
        }
    }
    const long long t1 = clock64();
    unsigned r = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) r ^= acc[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

template <int OP>
void run(const char* name, int per_iter) {
    unsigned* out; long long* cyc; long long h;
    cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&cyc, 8);
    const int iters = 4096;
    rate_kernel<OP><<<148, 1024>>>(out, 1, 2, iters, cyc);
    rate_kernel<OP><<<148, 1024>>>(out, 1, 2, iters, cyc);
    cudaDeviceSynchronize();
    cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    const double thread_ops = 1024.0 * iters * 8 * per_iter;
    printf("%-28s %8.1f thread-ops/clk/SM  (%lld cycles)\n", name, thread_ops / (double)h, h);
    cudaFree(out); cudaFree(cyc);
}

int main() {
    run<0>("VABSDIFF4.U8.ACC", 1);
    run<1>("PRMT", 1);
    run<2>("IDP.4A", 1);
    run<3>("__vminu4", 1);
    run<4>("LOP3+IADD3 (2 ops)", 2);
    return 0;
}
