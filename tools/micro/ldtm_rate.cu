// TMEM -> register read rate (tcgen05.ld.32x32b.x32) per SM as a function of the number of reading warps, with and without
// a consumer of the loaded values: the front-end's epilogue reads every score from TMEM, so this is one of its floors.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/ldtm_rate tools/micro/ldtm_rate.cu && build/ldtm_rate
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ void tc_ld32(uint32_t taddr, uint32_t (&v)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
          "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
          "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
          "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void tc_wait_ld(uint32_t (&v)[32]) {
    asm volatile("tcgen05.wait::ld.sync.aligned;"
                 : "+r"(v[0]), "+r"(v[1]), "+r"(v[2]), "+r"(v[3]), "+r"(v[4]), "+r"(v[5]), "+r"(v[6]), "+r"(v[7]), "+r"(v[8]), "+r"(v[9]), "+r"(v[10]), "+r"(v[11]), "+r"(v[12]), "+r"(v[13]), "+r"(v[14]), "+r"(v[15]), "+r"(v[16]), "+r"(v[17]), "+r"(v[18]), "+r"(v[19]), "+r"(v[20]), "+r"(v[21]), "+r"(v[22]), "+r"(v[23]), "+r"(v[24]), "+r"(v[25]), "+r"(v[26]), "+r"(v[27]), "+r"(v[28]), "+r"(v[29]), "+r"(v[30]), "+r"(v[31])
                 :
                 : "memory");
}

// MODE 0: loads only; 1: loads + a max3 reduction of every value (the epilogue's first pass); 2: two loads in flight per wait
template <int MODE>
__global__ void __launch_bounds__(1024, 1) ldtm_kernel(float* out, int iters, long long* cyc) {
    __shared__ uint32_t tmem_slot;
    const int warp = threadIdx.x >> 5;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(&tmem_slot)), "r"(512u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t base = tmem_slot + ((uint32_t)((warp & 3) * 32) << 16);
    float m = -1e30f;
    uint32_t va[32], vb[32];
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
        const uint32_t col = (uint32_t)(((warp >> 2) * 64 + (it & 1) * 256) & 511);
        if (MODE == 2) {
            tc_ld32(base + col, va);
            tc_ld32(base + col + 32, vb);
            tc_wait_ld(va);
            tc_wait_ld(vb);
#pragma unroll
            for (int i = 0; i < 32; i += 2) {
                asm("max.f32 %0, %0, %1, %2;" : "+f"(m) : "f"(__uint_as_float(va[i])), "f"(__uint_as_float(va[i + 1])));
                asm("max.f32 %0, %0, %1, %2;" : "+f"(m) : "f"(__uint_as_float(vb[i])), "f"(__uint_as_float(vb[i + 1])));
            }
        } else {
            tc_ld32(base + col, va);
            tc_wait_ld(va);
            tc_ld32(base + col + 32, vb);
            tc_wait_ld(vb);
            if (MODE == 1) {
#pragma unroll
                for (int i = 0; i < 32; i += 2) {
                    asm("max.f32 %0, %0, %1, %2;" : "+f"(m) : "f"(__uint_as_float(va[i])), "f"(__uint_as_float(va[i + 1])));
                    asm("max.f32 %0, %0, %1, %2;" : "+f"(m) : "f"(__uint_as_float(vb[i])), "f"(__uint_as_float(vb[i + 1])));
                }
            } else {
                m = fmaxf(m, __uint_as_float(va[0] ^ vb[31]));
            }
        }
    }
    const long long t1 = clock64();
    out[blockIdx.x * blockDim.x + threadIdx.x] = m;
    __syncthreads();
    if (threadIdx.x == 0) *cyc = t1 - t0;
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_slot), "r"(512u) : "memory");
}

int main() {
    float* out;
    long long* cyc;
    cudaMalloc(&out, 148 * 1024 * 4);
    cudaMallocManaged(&cyc, 8);
    const int iters = 4000;
    for (int mode = 0; mode < 3; ++mode)
        for (int warps : {4, 8, 16, 32}) {
            for (int rep = 0; rep < 2; ++rep) {
                if (mode == 0) ldtm_kernel<0><<<1, warps * 32>>>(out, iters, cyc);
                if (mode == 1) ldtm_kernel<1><<<1, warps * 32>>>(out, iters, cyc);
                if (mode == 2) ldtm_kernel<2><<<1, warps * 32>>>(out, iters, cyc);
                cudaError_t e = cudaDeviceSynchronize();
                if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
            }
            const double bytes = (double)iters * warps * 2 * 32 * 32 * 4;
            printf("mode %d warps %2d: %lld cycles, %.1f B/clk/SM, %.1f cycles per 64-column load pair per warp\n", mode, warps, *cyc,
                   bytes / (double)*cyc, (double)*cyc / iters);
        }
    return 0;
}
