// The fused tail of a Lloyd pass: everything the reference does between its assignment loop and the next one
// (src/vq_encoder.cpp:341-375 / :572-605), plus the operand preparation for the next pass of the tensor front-end, in
// ONE cooperative launch instead of six kernels and two memsets:
//   1. the convergence test `if (!hasChanged && iter > N) break;` on the (all-reduced) accumulator   (lloyd_control)
//   2. rare: errors[i] = d(x_i, centroid of i) with the OLD centroids when a cluster came out empty   (exact_errors)
//   3. centroid = float(sum) * (1.0f / count)                                                          (finalize)
//   4. rare: the serial-semantics empty-cluster reseed (ascending c, first maximum of errors, zero it) (reseed)
//   5. B' rows / folded norms / cmax of the new centroids for the next front-end launch                (prep_centroids)
// Steps 3 and 5 are per centroid, so a warp that finalises a centroid prepares its operand row at once.  Steps 2 and 4
// need every row (exact_errors) and a grid-wide arg-max per empty cluster: they use grid.sync(), which is why the kernel
// is launched cooperatively (one CTA per SM).  A lock-step batch runs texture-major over the same grid (texture t's words
// at the given strides); the rare path exists for a single texture only -- a block-sharded batch cannot reseed on the
// device (the winner may live on another GPU) and raises `sticky` instead, as lloyd_control_batch_kernel does.
// Every arithmetic step is the one of the kernel it replaces (update.cu, assign.cu, tc_prep.cuh): results are identical.
#include <cfloat>
#include <cooperative_groups.h>

#include "common.cuh"
#include "host_util.h"
#include "kernels.h"
#include "tc_prep.cuh"

namespace cg = cooperative_groups;

namespace yupvq {

namespace {

constexpr int kTailThreads = 256;

__device__ __forceinline__ void red_release_sys_add_u32(uint32_t* p, uint32_t v) {
    asm volatile("red.release.sys.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys_u32(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned long long global_timer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

__global__ void __launch_bounds__(kTailThreads) lloyd_tail_kernel(TailArgs a) {
    cg::grid_group grid = cg::this_grid();
    __shared__ unsigned int s_empty;
    __shared__ float s_val[kTailThreads / 32];
    __shared__ unsigned long long s_idx[kTailThreads / 32];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const uint32_t K = a.K;
    // ---- 0. block-sharded batch: the exchange step (XchgArgs, kernels.h) -- after it the sums of all ranks can be read.
    // No grid-wide barrier: every CTA pushes its part, bumps a counter on each peer (release, system scope) and waits on
    // this rank's own counters, which are monotonic: exchange number `epoch` is complete once a counter has reached
    // epoch * gridDim.x.  (The data buffers alternate with the epoch's parity; the counters need not.)
    const XchgArgs& x = a.x;
    const bool sum_on_read = x.n_ranks > 1 && x.one_shot != 0;
    if (x.n_ranks > 1 && !(x.err && __ldcg(x.err) != 0u)) {   // a timed-out exchange earlier in this call: do not wait again
        const int R = x.n_ranks;
        const size_t gtid = (size_t)blockIdx.x * kTailThreads + tid, gsz = (size_t)gridDim.x * kTailThreads;
        const uint32_t target = x.epoch * gridDim.x;
        auto signal_and_wait = [&](int phase) {
            __syncthreads();
            if (tid < R) {   // release: cumulative over the CTA's stores ordered before it by the barrier
                red_release_sys_add_u32(x.peer_flags[tid] + phase * kXchgMaxRanks + x.rank, 1u);
                const uint32_t* mine = x.peer_flags[x.rank] + phase * kXchgMaxRanks + tid;
                // the clock is looked at once per 1024 polls: the common case (counter already there, or a few microseconds
                // away) never reads it
                uint32_t polls = 0;
                unsigned long long t0 = 0;
                while ((int32_t)(ld_acquire_sys_u32(mine) - target) < 0) {
                    if (x.err == nullptr || (++polls & 1023u) != 0u) continue;
                    const unsigned long long now = global_timer_ns();
                    if (t0 == 0) {
                        t0 = now;
                    } else if (now - t0 > x.timeout_ns || __ldcg(x.err) != 0u) {
                        atomicMax(x.err, 1u + (uint32_t)phase * kXchgMaxRanks + (uint32_t)tid);   // which counter: phase, source rank
                        break;
                    }
                }
            }
            __syncthreads();
        };
        if (x.one_shot) {
            // one shot: the whole vector goes to every peer's slot for this rank; readers add the R vectors as they go
            for (int j = 0; j < R; ++j) {
                if (j == x.rank) continue;
                long long* dst = x.peer_stage[j] + (size_t)x.rank * x.slot_stride;
                for (size_t w = gtid; w < x.words; w += gsz) dst[w] = x.src[w];
            }
            signal_and_wait(0);
        } else {
            // two shots: slice j of this rank's partial sums goes to rank j's slot for this rank ...
            for (size_t w = gtid; w < x.words; w += gsz) {
                const size_t j = w / x.slice;
                x.peer_stage[j][(size_t)x.rank * x.slot_stride + (w - j * x.slice)] = x.src[w];
            }
            signal_and_wait(0);
            // ... rank j adds the R contributions of its slice and pushes the totals to every rank
            const size_t s0 = (size_t)x.rank * x.slice;
            const size_t s1 = s0 + x.slice < x.words ? s0 + x.slice : x.words;
            const long long* stage = x.peer_stage[x.rank];
            for (size_t w = s0 + gtid; w < s1; w += gsz) {
                long long t = 0;
                for (int r = 0; r < R; ++r) t += __ldcg(stage + (size_t)r * x.slot_stride + (w - s0));
                for (int j = 0; j < R; ++j) x.peer_result[j][w] = t;
            }
            signal_and_wait(1);
        }
    }
    // word w of the (all-reduced) accumulator vector
    auto ld_acc = [&](size_t w) -> long long {
        if (!sum_on_read) return __ldcg(a.acc + w);
        long long t = __ldcg(x.src + w);
        const long long* stage = x.peer_stage[x.rank];
        for (int r = 0; r < x.n_ranks; ++r)
            if (r != x.rank) t += __ldcg(stage + (size_t)r * x.slot_stride + w);
        return t;
    };
    bool any_rare = false;   // uniform across the grid: every CTA evaluates the same words
    for (uint32_t tex = 0; tex < a.n_textures; ++tex) {
        const size_t acc0 = (size_t)tex * a.acc_stride;   // this texture's words in the accumulator vector
        uint32_t* ctl = a.ctl + (size_t)tex * 3;
        uint32_t* flags = a.flags + (size_t)tex * a.flags_stride;
        // ---- 1. control (every CTA computes the same decision; CTA 0 publishes it)
        const bool stopped = __ldcg(ctl) != 0u;
        if (tid == 0) s_empty = 0;
        __syncthreads();
        unsigned int local = 0;
        if (!stopped)
            for (uint32_t c = tid; c < K; c += kTailThreads) local += ld_acc(acc0 + (size_t)K * 64 + c) == 0 ? 1u : 0u;
        local = warp_sum(local);
        if (lane == 0 && local) atomicAdd(&s_empty, local);
        __syncthreads();
        const unsigned int n_empty = s_empty;
        const bool changed = ld_acc(acc0 + (size_t)K * 65) != 0;
        const bool stop_now = !stopped && !changed && a.iter > a.break_after;
        __syncthreads();
        if (stopped || stop_now) {
            if (!stopped && blockIdx.x == 0 && tid == 0) {
                ctl[1] += 1u;
                ctl[2] = n_empty;
                ctl[0] = 1u;     // stop: this pass' update and every later pass are skipped (the centroids stay as they are)
                flags[0] = 0u;
                flags[1] = 0u;
            }
            continue;
        }
        if (blockIdx.x == 0 && tid == 0) {
            ctl[1] += 1u;        // assignment passes run
            ctl[2] = n_empty;    // empty clusters of this pass
            flags[0] = 0u;
            flags[1] = 0u;
            if (n_empty && a.sticky) *a.sticky = 1u;
        }
        const bool rare = n_empty != 0 && a.errors != nullptr && a.n_textures == 1;
        any_rare |= rare;
        float* cent = a.cent + (size_t)tex * K * 64;
        // ---- 2. rare: errors[] against the OLD centroids (src/vq_encoder.cpp:336), all rows, whole grid
        if (rare) {
            for (size_t blk = (size_t)blockIdx.x * kTailThreads + tid; blk < a.n_rows; blk += (size_t)gridDim.x * kTailThreads) {
                const float4* x = reinterpret_cast<const float4*>(a.feat) + blk * 16;
                const float4* c = reinterpret_cast<const float4*>(cent) + (size_t)a.assign[blk] * 16;
                DistAcc d;
                d.reset();
#pragma unroll
                for (int i = 0; i < 8; ++i) d.step(x[2 * i], x[2 * i + 1], c[2 * i], c[2 * i + 1]);
                a.errors[blk] = d.finish();
            }
            grid.sync();
        }
        // ---- 3 + 5. finalize and operand row, one warp per centroid (padding rows K..k_pad-1 included)
        const double inv_scale = a.inv_scales ? a.inv_scales[tex] : a.inv_scale;
        const float* mu = reinterpret_cast<const float*>(a.small + (size_t)tex * a.small_stride + a.mu_off);
        unsigned int* cmax_next = reinterpret_cast<unsigned int*>(a.small + (size_t)tex * a.small_stride + a.cmax_off) + ((a.iter + 1u) & 1u);
        unsigned int* cmax_cur = reinterpret_cast<unsigned int*>(a.small + (size_t)tex * a.small_stride + a.cmax_off) + (a.iter & 1u);
        uint16_t* b_main = reinterpret_cast<uint16_t*>(a.b_main) + (size_t)tex * a.k_pad * 128;
        uint16_t* b_ext = reinterpret_cast<uint16_t*>(a.b_ext) + (size_t)tex * a.k_pad * 16;
        const uint32_t warps_total = gridDim.x * (kTailThreads / 32);
        auto finalize_and_prep = [&](bool empties_only) {
            for (uint32_t c = blockIdx.x * (kTailThreads / 32) + wid; c < a.k_pad; c += warps_total) {
                const bool real = c < K;
                float x0 = 0.0f, x1 = 0.0f;
                if (real) {
                    const uint32_t cnt = (uint32_t)ld_acc(acc0 + (size_t)K * 64 + c);
                    if (empties_only && cnt != 0) continue;
                    if (cnt > 0) {
                        const float inv = __fdiv_rn(1.0f, (float)cnt);                                 // 1.0f / counts[c]
                        x0 = __fmul_rn((float)((double)ld_acc(acc0 + (size_t)c * 64 + lane) * inv_scale), inv);   // the sum, rounded to fp32 once
                        x1 = __fmul_rn((float)((double)ld_acc(acc0 + (size_t)c * 64 + 32 + lane) * inv_scale), inv);
                        cent[(size_t)c * 64 + lane] = x0;
                        cent[(size_t)c * 64 + 32 + lane] = x1;
                    } else {   // empty: the old centroid stays (or the reseeded row, second call)
                        x0 = __ldcg(cent + (size_t)c * 64 + lane);
                        x1 = __ldcg(cent + (size_t)c * 64 + 32 + lane);
                    }
                } else if (empties_only) {
                    continue;
                }
                tc_prep_centroid_warp(x0, x1, real, mu, b_main + (size_t)c * 128, b_ext + (size_t)c * 16, cmax_next, a.rgb, lane);
            }
        };
        finalize_and_prep(false);
        if (blockIdx.x == 0 && tid == 0) *cmax_cur = 0u;   // the slot the tail of the NEXT pass max-reduces into (nobody reads it before)
        // ---- 4. rare: reseed, ascending c, first maximum of errors (std::max_element), then its operand row again
        if (rare) {
            grid.sync();
            const size_t stripe = (a.n_rows + gridDim.x - 1) / gridDim.x;
            const size_t r0 = (size_t)blockIdx.x * stripe, r1 = r0 + stripe < a.n_rows ? r0 + stripe : a.n_rows;
            for (uint32_t c = 0; c < K; ++c) {
                if (ld_acc(acc0 + (size_t)K * 64 + c) != 0) continue;
                // this CTA's stripe: (max, first index)
                float bv = -FLT_MAX;
                unsigned long long bi = ~0ull;
                for (size_t j = r0 + tid; j < r1; j += kTailThreads) {
                    const float e = __ldcg(a.errors + j);
                    if (bi == ~0ull || e > bv) {   // ascending strided scan: strict '>' keeps the first maximum
                        bv = e;
                        bi = j;
                    }
                }
                auto merge = [&](float ov, unsigned long long oi) {
                    if (oi != ~0ull && (bi == ~0ull || ov > bv || (ov == bv && oi < bi))) {
                        bv = ov;
                        bi = oi;
                    }
                };
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) {
                    const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
                    const unsigned long long oi = __shfl_xor_sync(0xffffffffu, bi, o);
                    merge(ov, oi);
                }
                if (lane == 0) {
                    s_val[wid] = bv;
                    s_idx[wid] = bi;
                }
                __syncthreads();
                if (tid == 0) {
                    for (int w = 1; w < kTailThreads / 32; ++w) merge(s_val[w], s_idx[w]);
                    a.scratch_val[blockIdx.x] = bv;
                    a.scratch_idx[blockIdx.x] = bi;
                }
                grid.sync();
                if (blockIdx.x == 0) {
                    // stripes ascend with the CTA index, so the first CTA holding the maximum holds its first occurrence
                    if (wid == 0) {
                        float gv = -FLT_MAX;
                        unsigned long long gi = ~0ull;
                        for (uint32_t b = lane; b < gridDim.x; b += 32) {
                            const float ov = __ldcg(a.scratch_val + b);
                            const unsigned long long oi = __ldcg(a.scratch_idx + b);
                            if (oi != ~0ull && (gi == ~0ull || ov > gv || (ov == gv && oi < gi))) {
                                gv = ov;
                                gi = oi;
                            }
                        }
#pragma unroll
                        for (int o = 16; o > 0; o >>= 1) {
                            const float ov = __shfl_xor_sync(0xffffffffu, gv, o);
                            const unsigned long long oi = __shfl_xor_sync(0xffffffffu, gi, o);
                            if (oi != ~0ull && (gi == ~0ull || ov > gv || (ov == gv && oi < gi))) {
                                gv = ov;
                                gi = oi;
                            }
                        }
                        if (lane == 0) s_idx[0] = gi;
                    }
                    __syncthreads();
                    const unsigned long long worst = s_idx[0];
                    if (tid < 64) cent[(size_t)c * 64 + tid] = a.feat[worst * 64 + tid];
                    if (tid == 0) {
                        a.errors[worst] = 0.0f;
                        atomicAdd(flags + 2, 1u);
                    }
                }
                grid.sync();
            }
            finalize_and_prep(true);   // operand rows of the reseeded centroids
        }
    }
    (void)any_rare;
}

}  // namespace

cudaError_t launch_lloyd_tail(const TailArgs& args, cudaStream_t s) {
    int dev = 0, sms = 0, coop = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev);
    if (!coop || sms <= 0) return cudaErrorNotSupported;
    TailArgs a = args;
    void* params[] = {&a};
    return cudaLaunchCooperativeKernel((const void*)lloyd_tail_kernel, dim3((unsigned)sms), dim3(kTailThreads), params, 0, s);
}

}  // namespace yupvq
