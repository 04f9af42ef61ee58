// Issue rates of the front-end epilogue's instructions (assign_tc.cu): FMNMX3 (3-input max), FMNMX, FADD, SHF funnel shift,
// and the two mixes the epilogue can be written with, per SM per clock.  8 independent chains per thread, 1024 threads per SM.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/fp_alu_rate tools/micro/fp_alu_rate.cu && build/fp_alu_rate
#include <cstdio>
#include <cuda_runtime.h>

template <int OP, int WARPS>
__global__ void __launch_bounds__(WARPS * 32, 1) rate_kernel(float* out, float a0, float b0, int iters, long long* cyc) {
    float a[8], acc[8];
    unsigned k[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) { a[i] = a0 + threadIdx.x * 0.37f + i; acc[i] = (float)i; k[i] = i; }
    const long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < iters; ++it) {   // not unrolled: ptxas folds max(max(x, a), b) across iterations into FMNMX3
        const float b = b0 + (float)it;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (OP == 0) asm volatile("max.f32 %0, %0, %1, %2;" : "+f"(acc[i]) : "f"(a[i]), "f"(b));                 // FMNMX3
            if (OP == 1) asm volatile("max.f32 %0, %0, %1;" : "+f"(acc[i]) : "f"(b));                                // FMNMX
            if (OP == 2) asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(acc[i]) : "f"(b));                             // FADD
            if (OP == 3) asm volatile("shf.l.wrap.b32 %0, %1, %0, 1;" : "+r"(k[i]) : "r"(__float_as_uint(a[i])));    // SHF
            if (OP == 4) {   // the current epilogue's mix per two scores: one max3, two FADD, two SHF
                asm volatile("max.f32 %0, %0, %1, %2;" : "+f"(acc[i]) : "f"(a[i]), "f"(b));
                float d0, d1;
                asm volatile("add.rn.f32 %0, %1, %2;" : "=f"(d0) : "f"(a[i]), "f"(b));
                asm volatile("add.rn.f32 %0, %1, %2;" : "=f"(d1) : "f"(acc[i]), "f"(b));
                asm volatile("shf.l.wrap.b32 %0, %1, %0, 1;" : "+r"(k[i]) : "r"(__float_as_uint(d0)));
                asm volatile("shf.l.wrap.b32 %0, %1, %0, 1;" : "+r"(k[i]) : "r"(__float_as_uint(d1)));
            }
            if (OP == 5) {   // the same two scores with two-input maxima
                asm volatile("max.f32 %0, %0, %1;" : "+f"(acc[i]) : "f"(b));
                asm volatile("max.f32 %0, %0, %1;" : "+f"(a[i]) : "f"(b));     // a second chain: nothing nested to fold
                float d0, d1;
                asm volatile("add.rn.f32 %0, %1, %2;" : "=f"(d0) : "f"(a[i]), "f"(b));
                asm volatile("add.rn.f32 %0, %1, %2;" : "=f"(d1) : "f"(acc[i]), "f"(b));
                asm volatile("shf.l.wrap.b32 %0, %1, %0, 1;" : "+r"(k[i]) : "r"(__float_as_uint(d0)));
                asm volatile("shf.l.wrap.b32 %0, %1, %0, 1;" : "+r"(k[i]) : "r"(__float_as_uint(d1)));
            }
            if (OP == 6) {   // sign bits collected on the fma pipe: mad.hi.u32(d, 2, 2k) = (d >> 31) + 2k
                asm volatile("max.f32 %0, %0, %1;" : "+f"(acc[i]) : "f"(a[i]));
                float d0;
                asm volatile("add.rn.f32 %0, %1, %2;" : "=f"(d0) : "f"(a[i]), "f"(b));
                unsigned k2;
                asm volatile("mul.lo.u32 %0, %1, 2;" : "=r"(k2) : "r"(k[i]));
                asm volatile("mad.hi.u32 %0, %1, 2, %2;" : "=r"(k[i]) : "r"(__float_as_uint(d0)), "r"(k2));
            }
        }
    }
    const long long t1 = clock64();
    float r = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) r += acc[i] + (float)k[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

template <int OP, int WARPS>
void run(const char* name, int per_iter) {
    float* out; long long* cyc; long long h;
    cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&cyc, 8);
    const int iters = 4096;
    rate_kernel<OP, WARPS><<<148, WARPS * 32>>>(out, 1.f, 2.f, iters, cyc);
    rate_kernel<OP, WARPS><<<148, WARPS * 32>>>(out, 1.f, 2.f, iters, cyc);
    cudaDeviceSynchronize();
    cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    const double warp_instr = (double)WARPS * iters * 8 * per_iter;
    printf("%-44s warps %2d: %6.2f warp-instr/clk/SM (%5.2f per SMSP), %lld cycles\n", name, WARPS, warp_instr / (double)h, warp_instr / (double)h / 4, h);
    cudaFree(out); cudaFree(cyc);
}

int main() {
    run<0, 32>("FMNMX3 (max of three)", 1);
    run<0, 16>("FMNMX3 (max of three)", 1);
    run<1, 32>("FMNMX (max of two)", 1);
    run<2, 32>("FADD", 1);
    run<3, 32>("SHF.L.W", 1);
    run<4, 16>("epilogue mix now: max3 + 2 FADD + 2 SHF", 5);
    run<5, 16>("epilogue mix: 2 max + 2 FADD + 2 SHF", 6);
    run<6, 16>("per score: max + FADD + IMAD + IMAD.HI", 4);
    return 0;
}
