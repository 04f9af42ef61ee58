// K3 (tensor engine): nearest-centroid assignment as a dense N x K x 64 contraction on the tcgen05
// tensor cores, followed by an exact fp32 re-rank so the argmin is the reference's argmin.
// Reference: the loops of src/vq_encoder.cpp:330-340, :462-471, :561-571, :643-656 over
// OklabBlockDistanceSq_SIMD (:212-237).
//
//   d(x, c) = sum_i (w_i (x_i - c_i))^2 = |x'|^2 - 2 x'.c' + |c'|^2,   x' = w (x - mu), c' = w (c - mu)
//
// (w = [2,1,1,1] per pixel, mu = the texture's mean feature vector: distances are translation
// invariant and centring removes ~1.5 decimal digits of cancellation).  For a
// fixed block only  s(c) = x'.c' - |c'|^2 / 2  matters (argmin d == argmax s).  The front-end
// computes s for all K centroids with kind::f16 (bf16) MMAs on two-term bf16 splits of both operands,
//   x' = xh + xl (+ 2^-18 |x'|),  c' = ch + cl (+ 2^-18 |c'|),     s ~ xh.ch + xh.cl + xl.ch - |c'|^2/2
// which costs 13 K=16 MMAs per 128 x 256 tile instead of the 9 K=8 MMAs of a single-pass tf32 product but
// is 2^6 times more accurate (error 2^-16 |x'||c'| instead of 2^-10): the candidate lists shrink from
// 2.2 to ~1.0 entries per block and the exact re-rank has almost nothing to do.  Operand rows keep the
// tf32 design's 256 bytes:
//   A  = [xh | xl]  128 blocks    x (64 + 64) bf16   TMA, 128B swizzle, 2-stage ring
//   B  = [ch | cl]  256 centroids x (64 + 64) bf16   TMA, resident in smem when K <= 512
//   the three partial products re-use the same shared-memory tiles through their descriptors; a 13th
//   K-step multiplies a column of ones in A by a three-term bf16 split of -|c'|^2/2 in B, so the norm
//   is folded into the accumulator (24 bits) and the epilogue has no FFMA.
//   D  = 128 lanes x 256 columns fp32 in TMEM, two accumulator stages (512 columns).
// The epilogue (16 warps, one TMEM lane == one block per thread, four column slices) makes two passes over each
// accumulator stage: max3 reduction, then a sign-bit sweep that collects every centroid whose score
// is within the error margin (header comment below) of the running maximum.  Blocks end up with a short candidate list
// (uint16 x 7 + count in one uint4); rerank_kernel evaluates those candidates with the reference's
// exact operation order (common.cuh DistAcc) and resolves ties to the lowest index.
//
// Error margin: each operand is represented to 2^-18 relative, the dropped xl.cl term is
// 2^-18 |x_i c_i|, so
//   |x'.c' - mma| <= 3 * 2^-18 |x'||c'| + fp32 accumulation  <=  kDotErr * |x'| * Cmax,  kDotErr = 1.5 * 2^-16
// with Cmax = max_c |c'|.  A centroid can only be the reference argmin if
//   s(c) >= s_max - (2 * E + reference-rounding slack).
// The list is therefore a superset of the possible winners; an empty or overflowing list falls back
// to a full exact scan of that block inside rerank_kernel (flat textures, NaNs).
//
// Roofline: tensor pipe (2 * N * K * 208 FLOP per pass at the bf16 rate) against an epilogue that is
// ALU-issue bound (about 4 issue slots per block-centroid pair); HBM traffic is 256 B per block.
#include <cfloat>
#include <cstdlib>
#include <cuda.h>
#include <cuda_bf16.h>

#include "common.cuh"
#include "host_util.h"
#include "kernels.h"
#include "tc_prep.cuh"

namespace yupvq {

namespace {

constexpr int BM = 128;            // blocks per tile (MMA M, TMEM lanes)
constexpr int BN = 256;            // centroids per chunk (MMA N, TMEM columns per accumulator stage)
constexpr int kEpiWarps = 16;      // epilogue warps: 4 TMEM lane quadrants x 4 column slices
constexpr int kListCap = 4;        // candidates one (row, column slice) thread can record per tile
constexpr int kThreadsTc = (2 + kEpiWarps) * 32;
// v2 role layout (MW = merge warps): warps 0-15 = the 16 epilogue warps; warps 16-19 = one merge warp per row quadrant;
// warps 20-23 = TMA producer, MMA issuer, two idle.  Roles sit on warpgroup boundaries so that setmaxnreg can move registers
// from the control / merge warpgroups to the epilogue, and the tensor-feeding warps have the highest ids (scheduler priority).
constexpr int kThreadsTc2 = (8 + kEpiWarps) * 32;
constexpr int kMaxCand = 7;

constexpr uint32_t A_HALF_BYTES = BM * 128;        // 128 rows x 64 bf16 (xh, then xl)
constexpr uint32_t A_STAGE_BYTES = 2 * A_HALF_BYTES;
constexpr uint32_t B_HALF_BYTES = BN * 128;
constexpr uint32_t BX_BYTES = BN * 32;             // 256 rows x 16 bf16 (32B swizzle)
constexpr uint32_t B_STAGE_BYTES = 2 * B_HALF_BYTES + BX_BYTES;
constexpr uint32_t ONES_BYTES = BM * 32;

// dynamic shared memory map (offsets from a 1024-byte aligned base)
constexpr uint32_t OFF_A = 0;
constexpr uint32_t OFF_B = OFF_A + 2 * A_STAGE_BYTES;
constexpr uint32_t OFF_ONES = OFF_B + 2 * B_STAGE_BYTES;
constexpr uint32_t OFF_LIST = OFF_ONES + ONES_BYTES;                 // u16 [2 parity][4 slices][128][kListCap]
constexpr uint32_t OFF_MAX = OFF_LIST + 2 * 4 * 128 * kListCap * 2;  // float [2 parity][4 slices][128]
constexpr uint32_t OFF_CNT = OFF_MAX + 2 * 4 * 128 * 4;              // u8 [2 parity][4 slices][128]
constexpr uint32_t OFF_BAR = OFF_CNT + 2 * 4 * 128;                  // 28 mbarriers + tmem ptr
constexpr uint32_t SMEM_TC = OFF_BAR + 32 * 8 + 1024 /* alignment slack */;

// MFULL / MEMPTY [tile parity][row quadrant]: epilogue slices -> merge warp hand-over of the per-row candidate lists (v2)
enum { BAR_FULL_A = 0, BAR_EMPTY_A = 2, BAR_FULL_B = 4, BAR_EMPTY_B = 6, BAR_TFULL = 8, BAR_TEMPTY = 10, BAR_MFULL = 12,
       BAR_MEMPTY = 20, BAR_COUNT = 28 };

// error model constants (see header comment)
constexpr float kDotErr = 0.0000152587890625f * 1.5f;                             // 1.5 * 2^-16
constexpr float kRefRel = 0.000030517578125f;                                     // 2^-15

// ---- PTX wrappers ----------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
// Waiting warps must not spend issue slots: in the r02d capture 18 % of all executed warp instructions were the
// YIELD / TRYWAIT / BRA of these loops (9 wake-ups per accumulator wait: the epilogue warps of a sub-partition drift up to
// two chunks apart and the leaders wait for the MMAs the laggards have not released yet), on an issue port that is the
// kernel's bound.  A failed probe therefore sleeps (plain nanosleep: the .SYNCS form the suspend-time hint compiles to
// wakes on every mbarrier event of the CTA) before it probes again.
template <int kSleepNs>
__device__ __forceinline__ void mbar_wait_sleep(uint32_t bar, uint32_t parity) {
    for (;;) {
        uint32_t done;
        asm volatile(
            "{\n\t"
            ".reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t"
            "}\n"
            : "=r"(done)
            : "r"(bar), "r"(parity)
            : "memory");
        if (done) break;
        __nanosleep(kSleepNs);
    }
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) { mbar_wait_sleep<32>(bar, parity); }
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, int c0, int c1, uint32_t bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(dst),
        "l"(map), "r"(bar), "r"(c0), "r"(c1)
        : "memory");
}
// one lane of a converged warp (elect.sync): the compiler knows the guarded region runs on a single thread with
// warp-uniform operands, so tcgen05 / TMA operands stay in uniform registers (a `lane == 0` branch makes it wrap every
// such instruction in a broadcast loop)
__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "elect.sync _|p, 0xffffffff;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t"
        "}\n"
        : "=r"(pred));
    return pred != 0;
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// The two shared-memory descriptors travel as (low word, constant high word): stepping an operand along K only adds
// to the low word's address field (bytes >> 4), so a K-step costs one 32-bit add per operand.
__device__ __forceinline__ void tc_mma_bf16(uint32_t tmem_d, uint32_t a_lo, uint32_t a_hi, uint32_t b_lo, uint32_t b_hi, uint32_t idesc,
                                            uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        ".reg .b64 da, db;\n\t"
        "mov.b64 da, {%1, %2};\n\t"
        "mov.b64 db, {%3, %4};\n\t"
        "setp.ne.b32 p, %6, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %5, p;\n\t"
        "}\n" ::"r"(tmem_d),
        "r"(a_lo), "r"(a_hi), "r"(b_lo), "r"(b_hi), "r"(idesc), "r"(accumulate)
        : "memory");
}
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
// tcgen05.wait::ld with the destination registers as read-write operands, so the compiler cannot move a
// use of the tile above the wait
__device__ __forceinline__ void tc_wait_ld(uint32_t (&v)[32]) {
    asm volatile("tcgen05.wait::ld.sync.aligned;"
                 : "+r"(v[0]), "+r"(v[1]), "+r"(v[2]), "+r"(v[3]), "+r"(v[4]), "+r"(v[5]), "+r"(v[6]), "+r"(v[7]), "+r"(v[8]), "+r"(v[9]), "+r"(v[10]), "+r"(v[11]), "+r"(v[12]), "+r"(v[13]), "+r"(v[14]), "+r"(v[15]), "+r"(v[16]), "+r"(v[17]), "+r"(v[18]), "+r"(v[19]), "+r"(v[20]), "+r"(v[21]), "+r"(v[22]), "+r"(v[23]), "+r"(v[24]), "+r"(v[25]), "+r"(v[26]), "+r"(v[27]), "+r"(v[28]), "+r"(v[29]), "+r"(v[30]), "+r"(v[31])
                 :
                 : "memory");
}
// Read-only global load that neither nvcc nor ptxas can schedule above the computation of `after`: the address carries
// the sign bit of `after`, which is zero at run time (after = kDotErr * |x'| * Cmax >= 0) but not provably so.
__device__ __forceinline__ float ldg_f32_pinned(const float* p, float after) {
    return __ldg(p + (__float_as_uint(after) >> 31));
}
__device__ __forceinline__ float fmax3(float a, float b, float c) {
    float d;
    asm("max.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
    return d;
}
// two-input maxima for the -DYUPVQ_EPI_MAX2 A/B build (nvcc and ptxas both fold max(max(a, b), c) into FMNMX3; a plain
// max under a NaN-propagating one cannot be folded)
__device__ __forceinline__ float fmax2(float a, float b) {
    float d;
    asm volatile("max.f32 %0, %1, %2;" : "=f"(d) : "f"(a), "f"(b));
    return d;
}
__device__ __forceinline__ float fmax2n(float a, float b) {   // NaN-propagating form: cannot be folded with a plain max
    float d;
    asm volatile("max.NaN.f32 %0, %1, %2;" : "=f"(d) : "f"(a), "f"(b));
    return d;
}
__device__ __forceinline__ void epi_bar() { asm volatile("bar.sync 1, %0;" ::"n"(kEpiWarps * 32) : "memory"); }

// K-major shared-memory matrix descriptor (cute::UMMA::SmemDescriptor, version 1 == Blackwell).
// layout: 2 = 128B swizzle (rows of 128 B, 8-row atoms 1024 B apart), 6 = 32B swizzle (rows of 32 B,
// 8-row atoms 256 B apart).  LBO is unused for swizzled K-major operands.
__device__ __forceinline__ constexpr uint64_t make_desc(uint32_t saddr, uint32_t sbo_bytes, uint32_t layout) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3ffffu) >> 4);
    d |= (uint64_t)1 << 16;                          // leading byte offset (ignored)
    d |= (uint64_t)((sbo_bytes >> 4) & 0x3fffu) << 32;
    d |= (uint64_t)1 << 46;                          // descriptor version
    d |= (uint64_t)layout << 61;
    return d;
}

// kind::f16, D = f32, A/B = bf16, both K-major, M = 128, N = 256 (cute::UMMA::InstrDescriptor)
constexpr uint32_t kIdesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);

template <int N>
__device__ __forceinline__ void setmaxnreg_inc() { asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(N)); }
template <int N>
__device__ __forceinline__ void setmaxnreg_dec() { asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(N)); }

// threshold below which a score cannot belong to the reference argmin, given the best score m
__device__ __forceinline__ float cand_threshold(float m, float xn, float e_dot, bool& bad) {
    const float d_up = fmaxf(xn * xn - 2.0f * m + 2.0f * e_dot, 0.0f);
    const float margin = 2.0f * e_dot + kRefRel * d_up;
    // NaN / Inf anywhere (or only padding columns): no proof possible, take the exact scan
    if (!(m > -1.0e29f) || !(margin < 1.0e30f)) bad = true;
    float thr = m - margin;
    thr = __uint_as_float(__float_as_uint(thr) + (thr > 0.0f ? -1 : (thr < 0.0f ? 1 : 0)));  // one ulp down
    if (thr == 0.0f) thr = -FLT_MIN;
    return thr;
}

// (tile, texture) walk of one CTA through the batch without a division per tile
struct TileWalk {
    uint32_t t, tex, off, step, tpt;
    __device__ __forceinline__ TileWalk(uint32_t t0, uint32_t step_, uint32_t tpt_, bool batch)
        : t(t0), tex(batch ? t0 / tpt_ : 0u), off(0u), step(step_), tpt(batch ? tpt_ : 0xffffffffu) {
        off = batch ? t0 - tex * tpt_ : 0u;
    }
    __device__ __forceinline__ void advance() {
        t += step;
        off += step;
        while (off >= tpt) {
            off -= tpt;
            ++tex;
        }
    }
};

// ---- the front-end kernel -----------------------------------------------------------------------------
// One row's merge of its four column slices (see the epilogue comment below): executed by the slice-0 epilogue thread
// of the row (MW = false) or by the row quadrant's merge warp (MW = true).  Returns true when the row's assignment moved.
__device__ __forceinline__ bool merge_row(const uint16_t* __restrict__ lists, const float* __restrict__ s_max,
                                          const uint8_t* __restrict__ s_cnt, uint32_t tpar, int r, size_t row, float xn, float e_dot,
                                          uint32_t old_c, uint32_t tex, uint4* __restrict__ cand, uint32_t* __restrict__ assign,
                                          uint32_t* __restrict__ flags, uint32_t* __restrict__ amb_list, unsigned int* __restrict__ amb_count,
                                          const TcBatch& bt) {
    constexpr int kSplit = kEpiWarps / 4;
    float mq[kSplit];
#pragma unroll
    for (int q = 0; q < kSplit; ++q) mq[q] = s_max[(tpar * kSplit + q) * 128 + r];
    bool any_bad = false;
    const float m_all = fmaxf(fmaxf(mq[0], mq[1]), fmaxf(mq[2], mq[3]));
    const float thr_all = cand_threshold(m_all, xn, e_dot, any_bad);
    // up to 8 candidate ids, 16 bits each, in two 64-bit words (shifts instead of a dynamically indexed array: registers)
    unsigned long long lo = 0, hi = 0;
    uint32_t total = 0;
    bool ovf = any_bad;
#pragma unroll
    for (int q = 0; q < kSplit; ++q) {
        const uint32_t cq = s_cnt[(tpar * kSplit + q) * 128 + r];
        if (!(mq[q] > thr_all)) continue;      // this slice's best score cannot be the winner's
        ovf |= cq > (uint32_t)kListCap;
        // the slice's list: kListCap u16 = one 8-byte word
        const unsigned long long lw = *reinterpret_cast<const unsigned long long*>(lists + ((tpar * kSplit + q) * 128 + r) * kListCap);
        for (uint32_t i = 0; i < cq && !ovf && total < 8; ++i, ++total) {
            const unsigned long long c = (lw >> (16 * i)) & 0xffffull;
            if (total < 4) lo |= c << (16 * total);
            else hi |= c << (16 * (total - 4));
        }
    }
    bool moved = false;
    if (!ovf && total == 1 && assign) {
        // a single candidate is the exact argmin: final
        const uint32_t c = (uint32_t)(lo & 0xffffull);
        moved = old_c != c;
        if (flags && moved) flags[(size_t)tex * bt.flags_stride] = 1u;
        assign[row] = c;
    } else {
        uint32_t w0 = (uint32_t)lo, w1 = (uint32_t)(lo >> 32), w2 = (uint32_t)hi, w3 = (uint32_t)(hi >> 32);
        if (!ovf && total >= 1 && total <= (uint32_t)kMaxCand) {
            w3 = (w3 & 0xffffu) | (total << 16);
        } else {
            w0 = w1 = w2 = 0;
            w3 = 0xffffu << 16;  // empty (NaN) or overflow: full exact scan in the re-rank
        }
        cand[row] = make_uint4(w0, w1, w2, w3);
        if (amb_list) amb_list[atomicAdd(amb_count, 1u)] = (uint32_t)row;
    }
    return moved;
}

// delta update: (row, previous centroid) of every row that moved -- one atomic per warp and tile, and its result is only
// consumed one tile later, so the round trip to L2 hides behind the next tile
struct MovedList {
    uint32_t pend_mask = 0, pend_base = 0, pend_row = 0, pend_old = 0;
    __device__ __forceinline__ void flush(const TcBatch& bt, int lane) {
        if (pend_mask) {
            const uint32_t slot = __shfl_sync(0xffffffffu, pend_base, __ffs(pend_mask) - 1) + (uint32_t)__popc(pend_mask & ((1u << lane) - 1u));
            if (pend_mask & (1u << lane)) {
                bt.chg_rows[slot] = pend_row;
                bt.chg_old[slot] = pend_old;
            }
        }
    }
    __device__ __forceinline__ void push(const TcBatch& bt, int lane, bool moved, uint32_t row, uint32_t old_c) {
        flush(bt, lane);
        pend_mask = __ballot_sync(0xffffffffu, moved);
        pend_row = row;
        pend_old = old_c;
        if (pend_mask && lane == __ffs(pend_mask) - 1) pend_base = atomicAdd(bt.chg_count, (unsigned)__popc(pend_mask));
    }
};

static_assert(512 * 96 + 256 * 48 == kThreadsTc2 * 80, "v2 register budgets must add up to the launch allocation (80 per thread)");

template <bool BATCH, bool MW>
__global__ void __launch_bounds__(MW ? kThreadsTc2 : kThreadsTc, 1)
assign_tc_kernel(const __grid_constant__ CUtensorMap tm_a, const __grid_constant__ CUtensorMap tm_b,
                 const __grid_constant__ CUtensorMap tm_bx, const float* __restrict__ xnorm,
                 const float* __restrict__ cmax_ptr, uint32_t n, uint32_t num_tiles, uint32_t num_chunks,
                 uint32_t resident_b, uint4* __restrict__ cand, uint32_t* __restrict__ assign, uint32_t* __restrict__ flags,
                 uint32_t* __restrict__ amb_list, unsigned int* __restrict__ amb_count, const uint32_t* __restrict__ stop,
                 TcBatch bt) {
    constexpr int kThreadsCta = MW ? kThreadsTc2 : kThreadsTc;
    // Role -> warp id.  The warp scheduler of a sub-partition prefers the eligible warp with the HIGHEST id (measured,
    // B300_MICROARCH "arbiter priority"), so in v2 the two warps that feed the tensor pipe sit above the 16 epilogue warps:
    // with the issuer as warp 1 every always-eligible epilogue warp of its sub-partition went first and the MMAs of the
    // next chunk were issued late (r02e capture: 9 wake-ups per accumulator wait while the tensor pipe was 49 % active).
    constexpr int kEpiBase = MW ? 0 : 2;     // first epilogue warp
    constexpr int kMergeBase = 16;           // v2: warps 16-19 merge, 20 = TMA producer, 21 = MMA issuer, 22-23 idle
    constexpr int kProdWarp = MW ? 20 : 0, kMmaWarp = MW ? 21 : 1;
    // Lock-step batch (bt.tiles_per_tex < num_tiles): the rows of all textures are one contiguous matrix, tile t belongs
    // to texture t / tiles_per_tex; that texture's centroid operand is rows [tex * k_pad, (tex + 1) * k_pad) of the B maps
    // and its cmax / flags / stop words sit at the given strides.  A CTA reloads its resident B when it crosses
    // into the next texture (about once per 1/T of its tiles).  Single texture: tiles_per_tex == num_tiles.
    // The single-texture instantiation (BATCH = false) has none of the per-tile texture bookkeeping.
    if (!BATCH && stop && __ldg(stop)) return;  // converged: passes enqueued ahead of time do nothing
    // batch: a texture's stop word is read once per texture a role walks into, not per tile (the words share cache lines
    // with flags the epilogue writes)
    uint32_t chk_tex = 0xffffffffu;
    bool chk_live = true;
    auto live = [&](uint32_t x) {   // x = texture index
        if (!BATCH || !stop) return true;
        if (x != chk_tex) {
            chk_tex = x;
            chk_live = __ldg(stop + (size_t)x * bt.stop_stride) == 0u;
        }
        return chk_live;
    };
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t* base_ptr = smem_raw + (base - smem_u32(smem_raw));
    const uint32_t bar0 = base + OFF_BAR;
    auto bar = [&](int i) { return bar0 + 8u * (uint32_t)i; };
    volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(base_ptr + OFF_BAR + BAR_COUNT * 8);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (warp == kProdWarp && lane == 0) {
        for (int i = 0; i < 2; ++i) {
            mbar_init(bar(BAR_FULL_A + i), 1);
            mbar_init(bar(BAR_EMPTY_A + i), 1);
            mbar_init(bar(BAR_FULL_B + i), 1);
            mbar_init(bar(BAR_EMPTY_B + i), 1);
            mbar_init(bar(BAR_TFULL + i), 1);
            mbar_init(bar(BAR_TEMPTY + i), kEpiWarps);
        }
        for (int i = 0; i < 8; ++i) {
            mbar_init(bar(BAR_MFULL + i), kEpiWarps / 4);   // the four slice warps of a row quadrant
            mbar_init(bar(BAR_MEMPTY + i), 1);              // its merge warp
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == kMmaWarp) {  // TMEM: all 512 columns (two 256-column accumulator stages)
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32((const void*)tmem_slot)),
                     "r"(512u)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    // the column of ones that multiplies the folded norm terms (layout-invariant: every element is 1)
    for (uint32_t i = threadIdx.x; i < ONES_BYTES / 4; i += kThreadsCta) reinterpret_cast<uint32_t*>(base_ptr + OFF_ONES)[i] = 0x3f803f80u;  // bf16 1.0 x 2
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    // v2: registers follow the work (setmaxnreg at the head of each role's branch, so the role's code is compiled against
    // its own budget): 16 epilogue warps hold two 32-column score tiles each (96), the control and merge warpgroups need
    // little (48 each).  The pool setmaxnreg moves registers inside is what the CTA got at launch -- 768 threads x 80 --
    // so the budgets must add up to it: 512 x 96 + 256 x 48 = 61440 (an .inc beyond the pool would wait forever).
    uint16_t* const lists = reinterpret_cast<uint16_t*>(base_ptr + OFF_LIST);   // [2 parity][4 slices][128][kListCap]
    float* const s_max = reinterpret_cast<float*>(base_ptr + OFF_MAX);          // [2 parity][4 slices][128]
    uint8_t* const s_cnt = reinterpret_cast<uint8_t*>(base_ptr + OFF_CNT);      // [2 parity][4 slices][128]

    if (MW && warp >= kMergeBase + 4) {
        setmaxnreg_dec<48>();
    }
    if (warp == kProdWarp) {
        // ===== TMA producer ==============================================================================
        if (lane == 0) {
            uint32_t it = 0, u = 0, cur_tex = 0xffffffffu, bgen = 0;
            if (!BATCH && resident_b) {
                for (uint32_t j = 0; j < num_chunks; ++j) {
                    const uint32_t dstb = base + OFF_B + j * B_STAGE_BYTES;
                    mbar_expect_tx(bar(BAR_FULL_B + j), B_STAGE_BYTES);
                    tma_load_2d(dstb, &tm_b, 0, (int)(j * BN), bar(BAR_FULL_B + j));
                    tma_load_2d(dstb + B_HALF_BYTES, &tm_b, 64, (int)(j * BN), bar(BAR_FULL_B + j));
                    tma_load_2d(dstb + 2 * B_HALF_BYTES, &tm_bx, 0, (int)(j * BN), bar(BAR_FULL_B + j));
                }
                cur_tex = 0;
                bgen = 1;
            }
            for (TileWalk w(blockIdx.x, gridDim.x, bt.tiles_per_tex, BATCH); w.t < num_tiles; w.advance()) {
                const uint32_t t = w.t, tex = w.tex;
                if (!live(tex)) continue;
                const uint32_t sa = it & 1, pa = (it >> 1) & 1;
                ++it;
                mbar_wait(bar(BAR_EMPTY_A + sa), pa ^ 1);
                const uint32_t dst = base + OFF_A + sa * A_STAGE_BYTES;
                mbar_expect_tx(bar(BAR_FULL_A + sa), A_STAGE_BYTES);
                tma_load_2d(dst, &tm_a, 0, (int)(t * BM), bar(BAR_FULL_A + sa));
                tma_load_2d(dst + A_HALF_BYTES, &tm_a, 64, (int)(t * BM), bar(BAR_FULL_A + sa));
                if (BATCH && resident_b && tex != cur_tex) {
                    // (re)load the resident centroid operand (after this tile's A load is in flight); a reload waits until the
                    // MMAs of the previous texture are done
                    for (uint32_t j = 0; j < num_chunks; ++j) {
                        if (bgen) mbar_wait(bar(BAR_EMPTY_B + j), (bgen - 1) & 1);
                        const uint32_t dstb = base + OFF_B + j * B_STAGE_BYTES;
                        const int brow = (int)(tex * bt.k_pad + j * BN);
                        mbar_expect_tx(bar(BAR_FULL_B + j), B_STAGE_BYTES);
                        tma_load_2d(dstb, &tm_b, 0, brow, bar(BAR_FULL_B + j));
                        tma_load_2d(dstb + B_HALF_BYTES, &tm_b, 64, brow, bar(BAR_FULL_B + j));
                        tma_load_2d(dstb + 2 * B_HALF_BYTES, &tm_bx, 0, brow, bar(BAR_FULL_B + j));
                    }
                    ++bgen;
                    cur_tex = tex;
                }
                if (!resident_b) {
                    for (uint32_t j = 0; j < num_chunks; ++j, ++u) {
                        const uint32_t sb = u & 1, pb = (u >> 1) & 1;
                        mbar_wait(bar(BAR_EMPTY_B + sb), pb ^ 1);
                        const uint32_t dstb = base + OFF_B + sb * B_STAGE_BYTES;
                        const int brow = (int)(tex * bt.k_pad + j * BN);
                        mbar_expect_tx(bar(BAR_FULL_B + sb), B_STAGE_BYTES);
                        tma_load_2d(dstb, &tm_b, 0, brow, bar(BAR_FULL_B + sb));
                        tma_load_2d(dstb + B_HALF_BYTES, &tm_b, 64, brow, bar(BAR_FULL_B + sb));
                        tma_load_2d(dstb + 2 * B_HALF_BYTES, &tm_bx, 0, brow, bar(BAR_FULL_B + sb));
                    }
                }
            }
        }
    } else if (warp == kMmaWarp) {
        // ===== MMA issuer: the whole warp walks the tiles (uniform control flow), one elected lane issues ========
        {
            uint32_t it = 0, u = 0, cur_tex = 0xffffffffu, bgen = 0;
            const uint64_t ones_desc = make_desc(base + OFF_ONES, 256, 6);
            for (TileWalk w(blockIdx.x, gridDim.x, bt.tiles_per_tex, BATCH); w.t < num_tiles; w.advance()) {
                const uint32_t tex = w.tex;
                if (!live(tex)) continue;
                const bool first_of_tex = tex != cur_tex;   // wait for this texture's resident B
                if (first_of_tex) {
                    cur_tex = tex;
                    ++bgen;
                }
                // last tile this CTA runs on this texture: hand the resident B back to the producer afterwards
                // (a texture is live or stopped as a whole, so the next tile of the same texture is live too)
                bool last_of_tex = false;
                if (BATCH) {
                    TileWalk nx = w;
                    nx.advance();
                    last_of_tex = !(nx.t < num_tiles && nx.tex == tex);
                }
                const uint32_t sa = it & 1, pa = (it >> 1) & 1;
                ++it;
                mbar_wait(bar(BAR_FULL_A + sa), pa);
                const uint32_t a_addr = base + OFF_A + sa * A_STAGE_BYTES;
                for (uint32_t j = 0; j < num_chunks; ++j, ++u) {
                    const uint32_t acc = u & 1, pacc = (u >> 1) & 1;
                    mbar_wait(bar(BAR_TEMPTY + acc), pacc ^ 1);
                    uint32_t sb;
                    if (resident_b) {
                        sb = j;
                        if (first_of_tex) mbar_wait(bar(BAR_FULL_B + sb), (bgen - 1) & 1);
                    } else {
                        sb = u & 1;
                        mbar_wait(bar(BAR_FULL_B + sb), (u >> 1) & 1);
                    }
                    tc_fence_after();
                    const uint32_t b_addr = base + OFF_B + sb * B_STAGE_BYTES;
                    const uint32_t d_tmem = tmem_base + acc * BN;
                    // descriptor words: low = address field (+ the ignored LBO), high = stride / version / swizzle
                    constexpr uint32_t kHi128 = (uint32_t)(make_desc(0, 1024, 2) >> 32), kHi32 = (uint32_t)(make_desc(0, 256, 6) >> 32);
                    const uint32_t a_lo = (uint32_t)make_desc(a_addr, 1024, 2), b_lo = (uint32_t)make_desc(b_addr, 1024, 2);
                    // xl.ch + xh.cl + xh.ch: twelve K = 16 steps over the two 128-byte halves of each operand row
                    // (small terms first), then the folded norm
                    if (elect_one()) {   // one elected lane issues the chunk's MMAs and its commits
                        if (bt.rgb) {
                            // 48 colour columns: xl.ch and xh.cl over three K-steps each, xh.ch over four -- the fourth
                            // holds the ones / folded-norm columns
#pragma unroll
                            for (int k = 0; k < 10; ++k) {
                                const int seg = k < 3 ? 0 : (k < 6 ? 1 : 2);
                                const uint32_t kk = (uint32_t)(k < 3 ? k : (k < 6 ? k - 3 : k - 6)) * 32u;
                                const uint32_t koff = (seg == 0 ? A_HALF_BYTES : 0u) + kk;
                                const uint32_t boff = (seg == 1 ? B_HALF_BYTES : 0u) + kk;
                                tc_mma_bf16(d_tmem, a_lo + (koff >> 4), kHi128, b_lo + (boff >> 4), kHi128, kIdesc, k > 0 ? 1u : 0u);
                            }
                        } else {
#pragma unroll
                            for (int k = 0; k < 12; ++k) {
                                const int seg = k >> 2;
                                const uint32_t koff = (seg == 0 ? A_HALF_BYTES : 0u) + (uint32_t)(k & 3) * 32u;
                                const uint32_t boff = (seg == 1 ? B_HALF_BYTES : 0u) + (uint32_t)(k & 3) * 32u;
                                tc_mma_bf16(d_tmem, a_lo + (koff >> 4), kHi128, b_lo + (boff >> 4), kHi128, kIdesc, k > 0 ? 1u : 0u);
                            }
                            tc_mma_bf16(d_tmem, (uint32_t)ones_desc, kHi32, (uint32_t)make_desc(b_addr + 2 * B_HALF_BYTES, 256, 6), kHi32, kIdesc, 1u);
                        }
                        tc_commit(bar(BAR_TFULL + acc));
                        if (!resident_b || last_of_tex) tc_commit(bar(BAR_EMPTY_B + sb));
                    }
                    __syncwarp();
                }
                if (elect_one()) tc_commit(bar(BAR_EMPTY_A + sa));
                __syncwarp();
            }
        }
    } else if (MW && warp >= kMergeBase + 4) {
        // idle warps of the control warpgroup (they exist so that the roles fall on warpgroup boundaries)
    } else if (MW && warp >= kMergeBase) {
        // ===== merge warps (v2): warp 4 + m merges the four column slices of rows [32 m, 32 m + 32) of every tile ======
        // The epilogue slices hand their per-row maxima / counts / lists over through shared memory (double buffered by
        // tile parity) and an mbarrier pair, so no epilogue warp ever waits for a merge: the merge of tile t runs while
        // the 16 epilogue warps are already sweeping tile t + 1.
        setmaxnreg_dec<48>();
        const int m = warp - kMergeBase;
        const int r = m * 32 + lane;
        TileWalk w(blockIdx.x, gridDim.x, bt.tiles_per_tex, BATCH);
        auto next_live = [&](bool step) {
            if (step) w.advance();
            while (w.t < num_tiles && !live(w.tex)) w.advance();
            return w.t;
        };
        float cmax = BATCH ? 0.0f : __ldg(cmax_ptr);
        uint32_t cmax_tex = BATCH ? 0xffffffffu : 0u;
        const uint32_t t_first = next_live(false);
        uint32_t tex = w.tex;
        float xn_next = 0.0f;
        uint32_t old_next = 0;
        if (t_first < num_tiles && (size_t)t_first * BM + r < n) {
            xn_next = __ldg(xnorm + (size_t)t_first * BM + r);
            if (assign) old_next = assign[(size_t)t_first * BM + r];
        }
        MovedList ml;
        uint32_t ti = 0;
        for (uint32_t t = t_first, t_next, tex_next; t < num_tiles; t = t_next, tex = tex_next, ++ti) {
            t_next = next_live(true);
            tex_next = w.tex;
            if (BATCH && tex != cmax_tex) {
                cmax_tex = tex;
                cmax = __ldg(cmax_ptr + (size_t)tex * bt.cmax_stride);
            }
            const size_t row = (size_t)t * BM + r;
            const float xn = xn_next;
            const uint32_t old_c = old_next;
            const float e_dot = kDotErr * xn * cmax;
            {   // prefetch the next tile's |x'| and previous assignment -- AFTER the last use of this tile's values: loads issued
                // before it share its scoreboard, and the use then waits for the new loads' DRAM latency (r02e capture)
                const size_t nrow = (size_t)t_next * BM + r;
                const bool has = t_next < num_tiles && nrow < n;
                xn_next = has ? ldg_f32_pinned(xnorm + nrow, e_dot) : 0.0f;
                if (assign) old_next = has ? assign[nrow] : 0u;
            }
            const uint32_t tpar = ti & 1u;
            mbar_wait(bar(BAR_MFULL + tpar * 4 + m), (ti >> 1) & 1u);
            bool moved = false;
            if (row < n)
                moved = merge_row(lists, s_max, s_cnt, tpar, r, row, xn, e_dot, old_c, tex, cand, assign, flags, amb_list,
                                  amb_count, bt);
            __syncwarp();
            if (lane == 0) mbar_arrive(bar(BAR_MEMPTY + tpar * 4 + m));
            if (bt.chg_count) ml.push(bt, lane, moved, (uint32_t)row, old_c);
        }
        if (bt.chg_count) ml.flush(bt, lane);
    } else {
        // ===== epilogue: 16 warps, thread == (TMEM lane, column quarter) =================================
        // Every (row, column slice) thread works on its own: running maximum, threshold and candidate list
        // are slice-local, so there is no barrier between the two passes over an accumulator stage and the
        // 16 warps drift apart (one warp's TMEM latency hides behind another's ALU work).  The four slices
        // of a row meet once per tile: a slice whose maximum is below the row's final threshold cannot hold
        // the winner and is dropped whole; the others are concatenated.  (A slice-local threshold is never
        // above the final one, so the true winner is always recorded in its slice -- see the margin proof in
        // the header.)  Rows with exactly one candidate are final: their assignment is written here; the
        // rest go to the exact re-rank through a compact list.
        constexpr int kSplit = kEpiWarps / 4;        // column slices per accumulator stage
        constexpr int kCols = BN / kSplit;           // columns per thread per stage
        constexpr int kTiles = kCols / 32;           // 32-column register tiles per pass
        static_assert(kTiles == 2, "the unrolled load schedule below is written for two tiles per pass");
        if (MW) setmaxnreg_inc<96>();
        const int ew = warp - kEpiBase;
        const int quad = warp & 3;          // TMEM lane quadrant this warp may access (kEpiBase is a multiple of 4 or 2: see r)
        const int slice = ew >> 2;          // which kCols of the stage's 256 columns
        const int r = quad * 32 + lane;     // row inside the tile
        uint32_t lane_taddr = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(slice * kCols);
        uint32_t tfull0 = bar(BAR_TFULL);
        // opaque to the optimiser: without this it recomputes both addresses from %tid / the shared window base in every
        // chunk (about 25 instructions of a 260-instruction loop)
        asm volatile("" : "+r"(lane_taddr), "+r"(tfull0));
        uint32_t u = 0, tpar = 0;
        float xn_next = 0.0f;
        uint32_t old_next = 0;
        // tiles of stopped textures are skipped (by all three roles alike)
        // (tile, texture) walk without a division per tile: the epilogue is ALU-issue bound
        TileWalk w(blockIdx.x, gridDim.x, bt.tiles_per_tex, BATCH);
        auto next_live = [&](bool step) {
            if (step) w.advance();
            while (w.t < num_tiles && !live(w.tex)) w.advance();
            return w.t;
        };
        float cmax = BATCH ? 0.0f : __ldg(cmax_ptr);
        uint32_t cmax_tex = BATCH ? 0xffffffffu : 0u;
        const uint32_t t_first = next_live(false);
        uint32_t tex = w.tex;
        if (t_first < num_tiles && (size_t)t_first * BM + r < n) {
            xn_next = __ldg(xnorm + (size_t)t_first * BM + r);
            if (!MW && slice == 0 && assign) old_next = assign[(size_t)t_first * BM + r];
        }
        MovedList ml;
        uint32_t ti = 0;   // tiles this CTA has run (hand-over barrier phases)
        for (uint32_t t = t_first, t_next, tex_next; t < num_tiles; t = t_next, tex = tex_next, tpar ^= 1, ++ti) {
            t_next = next_live(true);
            tex_next = w.tex;
            if (BATCH && tex != cmax_tex) {
                cmax_tex = tex;
                cmax = __ldg(cmax_ptr + (size_t)tex * bt.cmax_stride);
            }
            // v2: this parity's hand-over buffers are free once the merge warp has consumed the tile before last
            if (MW) mbar_wait(bar(BAR_MEMPTY + tpar * 4 + quad), ((ti >> 1) & 1u) ^ 1u);
            const size_t row = (size_t)t * BM + r;
            const float xn = xn_next;
            const uint32_t old_c = old_next;
            const float e_dot = kDotErr * xn * cmax;
            {   // prefetch the next tile's |x'| (and previous assignment) so its DRAM latency hides behind this tile; issued after
                // e_dot, the only use of this tile's |x'| (see the merge warps)
                const size_t nrow = (size_t)t_next * BM + r;
                const bool has = t_next < num_tiles && nrow < n;
                xn_next = has ? ldg_f32_pinned(xnorm + nrow, e_dot) : 0.0f;
                if (!MW && slice == 0 && assign) old_next = has ? assign[nrow] : 0u;
            }
            float m_run = -INFINITY;
            uint32_t cnt = 0;
            bool local_bad = false;  // ignored: a slice of padding columns is not an error; the merge decides
            uint16_t* my_list = lists + ((tpar * kSplit + slice) * 128 + r) * kListCap;
            for (uint32_t j = 0; j < num_chunks; ++j, ++u) {
                const uint32_t acc = u & 1, pacc = (u >> 1) & 1;
                mbar_wait(tfull0 + 8u * acc, pacc);
                tc_fence_after();
                const uint32_t taddr = lane_taddr + acc * BN;
                uint32_t va[32], vb[32];
                float m0 = -INFINITY, m1 = -INFINITY, m2 = -INFINITY, m3 = -INFINITY;
                // FMNMX3 issues at FMNMX's rate (0.42 warp instructions / clk / SMSP, tools/micro/fp_alu_rate.cu), so the
                // three-input form halves the ALU-pipe time of this pass (-DYUPVQ_EPI_MAX2: two-input form, measured 4 % slower).
                auto reduce_max = [&](const uint32_t (&v)[32]) {
#pragma unroll
                    for (int i = 0; i < 32; i += 8) {
#ifndef YUPVQ_EPI_MAX2
                        m0 = fmax3(m0, __uint_as_float(v[i]), __uint_as_float(v[i + 1]));
                        m1 = fmax3(m1, __uint_as_float(v[i + 2]), __uint_as_float(v[i + 3]));
                        m2 = fmax3(m2, __uint_as_float(v[i + 4]), __uint_as_float(v[i + 5]));
                        m3 = fmax3(m3, __uint_as_float(v[i + 6]), __uint_as_float(v[i + 7]));
#else
                        m0 = fmax2n(fmax2(m0, __uint_as_float(v[i])), __uint_as_float(v[i + 1]));
                        m1 = fmax2n(fmax2(m1, __uint_as_float(v[i + 2])), __uint_as_float(v[i + 3]));
                        m2 = fmax2n(fmax2(m2, __uint_as_float(v[i + 4])), __uint_as_float(v[i + 5]));
                        m3 = fmax2n(fmax2(m3, __uint_as_float(v[i + 6])), __uint_as_float(v[i + 7]));
#endif
                    }
                };
                // ---- one read of the accumulator: both 32-column register tiles of this thread's slice are loaded,
                // the stage goes back to the MMA warp at once (the next chunk's MMAs run under the whole reduction /
                // sweep below), and both passes work from registers.
                tc_ld32(taddr, va);
                tc_ld32(taddr + 32, vb);
                tc_wait_ld(va);
                tc_wait_ld(vb);
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(tfull0 + 8u * (BAR_TEMPTY - BAR_TFULL) + 8u * acc);
                // ---- pass 1 (registers): maximum score over this thread's columns
                reduce_max(va);
                reduce_max(vb);
                const float m_prev = m_run;
                m_run = fmaxf(fmax3(m_run, m0, m1), fmaxf(m2, m3));
                const float thr = cand_threshold(m_run, xn, e_dot, local_bad);
                // every candidate recorded in earlier chunks has score <= m_prev: once the threshold has moved
                // past that, none of them can win any more
                if (m_prev < thr) cnt = 0;
                // ---- pass 2: sign-bit sweep, candidates are the columns with (thr - s) < 0.  Two independent
                // 16-bit funnel-shift chains per register tile (four per chunk) keep the ALU pipe busy; one multiply-add
                // joins them.
                auto sweep = [&](const uint32_t (&v)[32], uint32_t col0) {
                    uint32_t k0 = 0, k1 = 0;
#pragma unroll
                    for (int i = 0; i < 16; ++i) {
                        k0 = __funnelshift_l(__float_as_uint(thr - __uint_as_float(v[i])), k0, 1);
                        k1 = __funnelshift_l(__float_as_uint(thr - __uint_as_float(v[16 + i])), k1, 1);
                    }
                    uint32_t mask = k0 * 65536u + k1;  // bit 31 - i <-> column col0 + i
                    while (mask) {
                        const int b = __clz(mask);
                        mask &= ~(0x80000000u >> b);
                        if (cnt < (uint32_t)kListCap) my_list[cnt] = (uint16_t)(col0 + b);
                        ++cnt;
                    }
                };
                const uint32_t cbase = j * BN + slice * kCols;
                sweep(va, cbase);
                sweep(vb, cbase + 32);
            }
            // ---- the four slices of a row meet.  Lists, maxima and counts are double buffered by tile parity.
            // count 100 = list overflow.
            s_cnt[(tpar * kSplit + slice) * 128 + r] = (uint8_t)(cnt > (uint32_t)kListCap ? 100u : cnt);
            s_max[(tpar * kSplit + slice) * 128 + r] = m_run;
            if (MW) {
                // v2: hand the tile to the quadrant's merge warp and go on with the next one
                __syncwarp();
                if (lane == 0) mbar_arrive(bar(BAR_MFULL + tpar * 4 + quad));
            } else {
                // v1: one barrier per tile, the slice-0 thread of each row merges
                epi_bar();
                bool moved = false;
                if (slice == 0 && row < n)
                    moved = merge_row(lists, s_max, s_cnt, tpar, r, row, xn, e_dot, old_c, tex, cand, assign, flags, amb_list, amb_count, bt);
                if (bt.chg_count && slice == 0) ml.push(bt, lane, moved, (uint32_t)row, old_c);
            }
        }
        if (!MW && bt.chg_count && slice == 0) ml.flush(bt, lane);
    }

    tc_fence_before();
    __syncthreads();
    if (warp == kMmaWarp) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
    }
}

// ---- operand preparation ------------------------------------------------------------------------------
// column sums of the N x 64 feature matrix (fp64, atomics: mu only steers the error margin, it never
// changes a result, so its last bits need not be reproducible)
__global__ void __launch_bounds__(256) col_sums_kernel(const float* __restrict__ feat, size_t n, double* __restrict__ sums) {
    __shared__ double red[4][64];
    const int col = threadIdx.x & 63, sub = threadIdx.x >> 6;
    double acc = 0.0;
    for (size_t i = (size_t)blockIdx.x * 4 + sub; i < n; i += (size_t)gridDim.x * 4) acc += (double)feat[i * 64 + col];
    red[sub][col] = acc;
    __syncthreads();
    if (sub == 0) atomicAdd(sums + col, red[0][col] + red[1][col] + red[2][col] + red[3][col]);
}

__global__ void mean_kernel(const double* __restrict__ sums, double inv_n, float* __restrict__ mu) {
    const float m = (float)(sums[threadIdx.x] * inv_n);
    mu[threadIdx.x] = isfinite(m) ? m : 0.0f;
}

// A' row (256 B) = [bf16 hi of w (x - mu) x 64 | bf16 lo x 64],  xnorm = |w (x - mu)| (rounded up a little)
// rgb: row = [hi: 16 x (L a b) | 1 1 1 0..0 (16) || lo: 16 x (L a b) | 0..0 (16)]  (the alpha column is exactly zero)
__global__ void __launch_bounds__(256) prep_blocks_kernel(const float* __restrict__ feat, size_t n, const float* __restrict__ mu,
                                                          float* __restrict__ a_prime, float* __restrict__ xnorm, int rgb) {
    const size_t row = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 4;   // 16 threads per row
    const int q = threadIdx.x & 15;                                            // pixel == float4 index
    float ss = 0.0f;
    if (row < n) {
        const float4 x = ldg_nc_f4(reinterpret_cast<const float4*>(feat) + row * 16 + q);
        const float4 m = reinterpret_cast<const float4*>(mu)[q];
        float4 o;
        o.x = 2.0f * (x.x - m.x);
        o.y = x.y - m.y;
        o.z = x.z - m.z;
        o.w = x.w - m.w;
        ss = o.x * o.x + o.y * o.y + o.z * o.z + o.w * o.w;
        uint16_t h0, h1, h2, h3, l0, l1, l2, l3;
        split_bf16(o.x, h0, l0);
        split_bf16(o.y, h1, l1);
        split_bf16(o.z, h2, l2);
        split_bf16(o.w, h3, l3);
        if (rgb) {
            uint16_t* r16 = reinterpret_cast<uint16_t*>(a_prime) + row * 128;   // 128 bf16 per row
            r16[3 * q] = h0; r16[3 * q + 1] = h1; r16[3 * q + 2] = h2;
            r16[64 + 3 * q] = l0; r16[64 + 3 * q + 1] = l1; r16[64 + 3 * q + 2] = l2;
            r16[48 + q] = q < 3 ? (uint16_t)0x3f80u : (uint16_t)0u;             // bf16 1.0 multiplies the folded norm terms
            r16[112 + q] = 0;
        } else {
            uint2* dst = reinterpret_cast<uint2*>(a_prime) + row * 32 + q;      // 32 x 8 B per row
            dst[0] = make_uint2((uint32_t)h0 | ((uint32_t)h1 << 16), (uint32_t)h2 | ((uint32_t)h3 << 16));
            dst[16] = make_uint2((uint32_t)l0 | ((uint32_t)l1 << 16), (uint32_t)l2 | ((uint32_t)l3 << 16));
        }
    }
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
    if (row < n && q == 0) xnorm[row] = sqrtf(ss) * 1.0001f;
}

// One warp per centroid row (tc_prep.cuh); blockIdx.y = texture of a lock-step batch: operands stacked k_pad rows
// apart, mu / cmax small_stride bytes apart.
__global__ void __launch_bounds__(256) prep_centroids_kernel(const float* __restrict__ cent, uint32_t K, uint32_t k_pad,
                                                             const float* __restrict__ mu, float* __restrict__ b_main,
                                                             float* __restrict__ b_ext, unsigned int* __restrict__ cmax_bits,
                                                             size_t small_stride, int rgb) {
    cent += (size_t)blockIdx.y * K * 64;
    b_main += (size_t)blockIdx.y * k_pad * 64;
    b_ext += (size_t)blockIdx.y * k_pad * 8;
    mu = reinterpret_cast<const float*>(reinterpret_cast<const uint8_t*>(mu) + (size_t)blockIdx.y * small_stride);
    cmax_bits = reinterpret_cast<unsigned int*>(reinterpret_cast<uint8_t*>(cmax_bits) + (size_t)blockIdx.y * small_stride);
    const int lane = threadIdx.x & 31;
    const uint32_t c = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (c >= k_pad) return;
    const bool real = c < K;
    const float x0 = real ? cent[(size_t)c * 64 + lane] : 0.0f, x1 = real ? cent[(size_t)c * 64 + 32 + lane] : 0.0f;
    tc_prep_centroid_warp(x0, x1, real, mu, reinterpret_cast<uint16_t*>(b_main) + (size_t)c * 128,
                          reinterpret_cast<uint16_t*>(b_ext) + (size_t)c * 16, cmax_bits, rgb, lane);
}

// ---- exact re-rank ---------------------------------------------------------------------------------------
// Streaming, warp-granular: a warp owns 32 consecutive blocks.  It reads their candidate words with one
// coalesced 512-byte load, votes which blocks need arithmetic at all (more than one candidate, or
// errors[] requested), compacts those and evaluates them 16 at a time with two lanes per block: lane h
// (0/1) owns the reference's lane accumulators 4h..4h+3 (src/vq_encoder.cpp:220-229), keeps its half of
// the feature row in registers (8 x 16-byte no-allocate loads, so L1 is left to the K centroid rows) and
// walks the block's candidates.  The horizontal add tree (:231-236) is four shuffles.  No shared-memory
// staging and no CTA-wide barrier: the only cross-lane traffic is warp shuffles, and the final
// assign/errors stores are coalesced.
constexpr int kRrThreads = 256;

// the feature-row half of lane h: float4 #(2i + h), i = 0..7
struct RowHalf {
    float4 v[8];
    __device__ __forceinline__ void load_stream(const float* __restrict__ row, int h) {
#pragma unroll
        for (int i = 0; i < 8; ++i) v[i] = ldg_nc_f4(reinterpret_cast<const float4*>(row) + 2 * i + h);
    }
};

// distance of the block held by a lane pair to centroid row c (L1/L2-resident); both lanes return it
__device__ __forceinline__ float pair_distance(const RowHalf& x, const float* __restrict__ c, int h) {
    float4 cv[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) cv[i] = __ldg(reinterpret_cast<const float4*>(c) + 2 * i + h);
    float a0 = 0.0f, a1 = 0.0f, a2 = 0.0f, a3 = 0.0f;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        float d;
        d = __fmul_rn(__fsub_rn(x.v[i].x, cv[i].x), 2.0f); a0 = __fmaf_rn(d, d, a0);
        d = __fsub_rn(x.v[i].y, cv[i].y);                  a1 = __fmaf_rn(d, d, a1);
        d = __fsub_rn(x.v[i].z, cv[i].z);                  a2 = __fmaf_rn(d, d, a2);
        d = __fsub_rn(x.v[i].w, cv[i].w);                  a3 = __fmaf_rn(d, d, a3);
    }
    // s_j = acc_j + acc_{j+4}: the partner lane holds the other half (fp add is commutative)
    const float s0 = __fadd_rn(a0, __shfl_xor_sync(0xffffffffu, a0, 1));
    const float s1 = __fadd_rn(a1, __shfl_xor_sync(0xffffffffu, a1, 1));
    const float s2 = __fadd_rn(a2, __shfl_xor_sync(0xffffffffu, a2, 1));
    const float s3 = __fadd_rn(a3, __shfl_xor_sync(0xffffffffu, a3, 1));
    return __fadd_rn(__fadd_rn(s0, s1), __fadd_rn(s2, s3));
}

// stat[0] += exact distance evaluations on multi-candidate / overflow blocks, stat[1] += overflow blocks
__global__ void __launch_bounds__(kRrThreads, 2) rerank_kernel(const float* __restrict__ feat, size_t n, const float* __restrict__ cent,
                                                               uint32_t K, const uint4* __restrict__ cand, uint32_t* __restrict__ assign,
                                                               float* __restrict__ errors, uint32_t* __restrict__ flags,
                                                               unsigned int* __restrict__ stat, const uint32_t* __restrict__ stop) {
    __shared__ uint8_t s_src[kRrThreads / 32][32];
    const uint32_t stopped = stop ? __ldg(stop) : 0u;  // tested after the candidate load has been issued
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int g = lane >> 1, h = lane & 1;
    const size_t blk0 = ((size_t)blockIdx.x * (kRrThreads / 32) + wid) * 32;
    if (blk0 >= n) return;
    const size_t blk = blk0 + lane;
    const bool valid = blk < n;
    uint4 cw = make_uint4(0, 0, 0, 0);
    if (valid) cw = ldg_nc_v4(cand + blk);
    if (stopped) return;
    const uint32_t cnt_raw = valid ? (cw.w >> 16) : 0u;
    const bool overflow = cnt_raw == 0xffffu;
    // blocks that need arithmetic: none when the list has one entry and no error is wanted
    const bool need = valid && !overflow && (cnt_raw > 1 || (errors != nullptr && cnt_raw == 1));
    uint32_t best_c = cw.x & 0xffffu;
    float best = 0.0f;
    uint32_t pairs = (need && cnt_raw > 1) ? cnt_raw : 0u;

    const unsigned need_mask = __ballot_sync(0xffffffffu, need);
    const int n_need = __popc(need_mask);
    const int my_rank = __popc(need_mask & ((1u << lane) - 1u));
    if (need) s_src[wid][my_rank] = (uint8_t)lane;
    __syncwarp();
    for (int base = 0; base < n_need; base += 16) {
        const int r = base + g;
        const bool act = r < n_need;
        const int src = act ? (int)s_src[wid][r] : 0;
        const uint32_t w0 = __shfl_sync(0xffffffffu, cw.x, src), w1 = __shfl_sync(0xffffffffu, cw.y, src);
        const uint32_t w2 = __shfl_sync(0xffffffffu, cw.z, src), w3 = __shfl_sync(0xffffffffu, cw.w, src);
        const uint32_t m = act ? (w3 >> 16) : 0u;
        RowHalf x;
        x.load_stream(feat + (blk0 + (size_t)src) * 64, h);
        const uint32_t m_max = __reduce_max_sync(0xffffffffu, m);
        // the reference scans c ascending with strict '<' from FLT_MAX: smallest distance, lowest index
        float bd = FLT_MAX;
        uint32_t bc = 0;
        bool have = false;
        for (uint32_t i = 0; i < m_max; ++i) {
            const uint32_t wsel = i < 2 ? w0 : (i < 4 ? w1 : (i < 6 ? w2 : w3));
            const uint32_t c = i < m ? ((wsel >> (16 * (i & 1))) & 0xffffu) : 0u;
            const float d = pair_distance(x, cent + (size_t)c * 64, h);
            if (i < m && d < FLT_MAX && (!have || d < bd || (d == bd && c < bc))) {
                bd = d;
                bc = c;
                have = true;
            }
        }
        // hand the result back to the lane that owns the block
        const int back = (2 * (my_rank - base)) & 31;
        const uint32_t rc = __shfl_sync(0xffffffffu, bc, back);
        const float rd = __shfl_sync(0xffffffffu, bd, back);
        if (need && my_rank >= base && my_rank < base + 16) {
            best_c = rc;
            best = rd;
        }
    }
    // ---- overflow blocks (empty or overfull list: flat textures, NaN): full scan, 16 centroids per step
    unsigned ovf_mask = __ballot_sync(0xffffffffu, overflow);
    const uint32_t n_ovf = (uint32_t)__popc(ovf_mask);
    while (ovf_mask) {
        const int src = __ffs(ovf_mask) - 1;
        ovf_mask &= ovf_mask - 1;
        RowHalf x;
        x.load_stream(feat + (blk0 + (size_t)src) * 64, h);
        float bd = FLT_MAX;
        uint32_t bc = 0;
        for (uint32_t c0 = 0; c0 < K; c0 += 16) {
            const uint32_t c = c0 + g;
            const float d = pair_distance(x, cent + (size_t)(c < K ? c : 0) * 64, h);
            if (c < K && d < bd) {  // ascending c inside the group: strict '<' keeps the first minimum
                bd = d;
                bc = c;
            }
        }
        // combine the 16 groups: smallest distance, lowest index on ties (what a serial scan returns)
#pragma unroll
        for (int o = 2; o <= 16; o <<= 1) {
            const float od = __shfl_xor_sync(0xffffffffu, bd, o);
            const uint32_t oc = __shfl_xor_sync(0xffffffffu, bc, o);
            if (od < bd || (od == bd && od < FLT_MAX && oc < bc)) {
                bd = od;
                bc = oc;
            }
        }
        if (lane == src) {
            best_c = bc;
            best = bd;
            pairs = K;
        }
    }
    if (valid) {
        if (flags && assign[blk] != best_c) flags[0] = 1u;
        assign[blk] = best_c;
        if (errors) errors[blk] = best;
    }
    if (stat) {
        pairs = __reduce_add_sync(0xffffffffu, pairs);
        if (lane == 0 && pairs) atomicAdd(stat, pairs);
        if (lane == 0 && n_ovf) atomicAdd(stat + 1, n_ovf);
    }
}

// List-driven variant for the Lloyd passes: the front-end has already written the assignment of every block
// with a single candidate; `amb_list[0, *amb_count)` holds the rows that still need exact arithmetic (about
// 1 % of a texture).  Two lanes per row as above.  The last CTA to finish clears the counter for the next pass.
__global__ void __launch_bounds__(kRrThreads, 2) rerank_sparse_kernel(const float* __restrict__ feat, const float* __restrict__ cent,
                                                                      uint32_t K, const uint4* __restrict__ cand,
                                                                      const uint32_t* __restrict__ amb_list, unsigned int* amb_count,
                                                                      unsigned int* ticket, uint32_t* __restrict__ assign,
                                                                      uint32_t* __restrict__ flags, unsigned int* __restrict__ stat,
                                                                      TcBatch bt, AccPrep prep) {
    const float* const cent_base = cent;
    uint32_t* const flags_base = flags;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int g = lane >> 1, h = lane & 1;
    const uint32_t total = __ldcg(amb_count);
    const uint32_t n_warps = gridDim.x * (kRrThreads / 32);
    uint32_t pairs = 0, n_ovf = 0;
    for (uint32_t base = (blockIdx.x * (kRrThreads / 32) + wid) * 16; base < total; base += n_warps * 16) {
        const uint32_t idx = base + g;
        const bool act = idx < total;
        const size_t row = __ldcg(amb_list + (act ? idx : base));
        {   // lock-step batch: the row's texture selects the centroid table and the changed flag
            const uint32_t tex = (uint32_t)(row / bt.rows_per_tex);
            cent = cent_base + (size_t)tex * K * 64;
            flags = flags_base ? flags_base + (size_t)tex * bt.flags_stride : nullptr;
        }
        const uint4 cw = ldg_nc_v4(cand + row);
        const uint32_t m_raw = cw.w >> 16;
        const bool overflow = act && m_raw == 0xffffu;
        const uint32_t m = (act && !overflow) ? m_raw : 0u;
        RowHalf x;
        x.load_stream(feat + row * 64, h);
        const uint32_t m_max = __reduce_max_sync(0xffffffffu, m);
        float bd = FLT_MAX;
        uint32_t bc = 0;
        bool have = false;
        for (uint32_t i = 0; i < m_max; ++i) {
            const uint32_t wsel = i < 2 ? cw.x : (i < 4 ? cw.y : (i < 6 ? cw.z : cw.w));
            const uint32_t c = i < m ? ((wsel >> (16 * (i & 1))) & 0xffffu) : 0u;
            const float d = pair_distance(x, cent + (size_t)c * 64, h);
            if (i < m && d < FLT_MAX && (!have || d < bd || (d == bd && c < bc))) {
                bd = d;
                bc = c;
                have = true;
            }
        }
        if (h == 0) pairs += m;
        // rows without a usable list (flat textures, NaN): full scan, the whole warp on one row at a time
        unsigned ovf_mask = __ballot_sync(0xffffffffu, overflow && h == 0);
        while (ovf_mask) {
            const int src = __ffs(ovf_mask) - 1;
            ovf_mask &= ovf_mask - 1;
            const size_t orow = __shfl_sync(0xffffffffu, (unsigned long long)row, src);
            RowHalf ox;
            ox.load_stream(feat + orow * 64, h);
            const float* ocent = cent_base + (size_t)(orow / bt.rows_per_tex) * K * 64;
            float od = FLT_MAX;
            uint32_t oc = 0;
            for (uint32_t c0 = 0; c0 < K; c0 += 16) {
                const uint32_t c = c0 + g;
                const float d = pair_distance(ox, ocent + (size_t)(c < K ? c : 0) * 64, h);
                if (c < K && d < od) {  // ascending c inside the group: strict '<' keeps the first minimum
                    od = d;
                    oc = c;
                }
            }
#pragma unroll
            for (int o = 2; o <= 16; o <<= 1) {
                const float pd = __shfl_xor_sync(0xffffffffu, od, o);
                const uint32_t pc = __shfl_xor_sync(0xffffffffu, oc, o);
                if (pd < od || (pd == od && pd < FLT_MAX && pc < oc)) {
                    od = pd;
                    oc = pc;
                }
            }
            if (lane == src) {
                bc = oc;
                pairs += K;
                ++n_ovf;
            }
        }
        if (act && h == 0) {
            const uint32_t prev = assign[row];
            if (prev != bc) {
                if (flags) flags[0] = 1u;
                if (bt.chg_count) {  // delta update (about 1 % of the rows come through here)
                    const uint32_t slot = atomicAdd(bt.chg_count, 1u);
                    bt.chg_rows[slot] = (uint32_t)row;
                    bt.chg_old[slot] = prev;
                }
            }
            assign[row] = bc;
        }
    }
    if (stat) {
        pairs = __reduce_add_sync(0xffffffffu, pairs);
        n_ovf = __reduce_add_sync(0xffffffffu, n_ovf);
        if (lane == 0 && pairs) atomicAdd(stat, pairs);
        if (lane == 0 && n_ovf) atomicAdd(stat + 1, n_ovf);
    }
    __shared__ uint32_t s_last;
    __threadfence();   // this CTA's moved-row appends are visible before its ticket
    __syncthreads();
    if (threadIdx.x == 0) {
        const bool last = atomicAdd(ticket, 1u) == gridDim.x - 1;  // every CTA has read the count by now
        if (last) {
            *amb_count = 0;
            *ticket = 0;
        }
        s_last = last ? 1u : 0u;
    }
    if (!prep.dctl) return;
    __syncthreads();
    if (!s_last) return;
    // ---- update prologue (== acc_prepare_kernel), by the last CTA: every moved row of this pass has been listed
    __threadfence();
    const uint32_t n_chg = __ldcg(prep.dctl);
    const bool full = prep.force_full || n_chg > prep.full_above;
    __syncthreads();
    if (threadIdx.x == 0) {
        prep.dctl[32] = n_chg;
        prep.dctl[33] = full ? 1u : 0u;
        prep.dctl[0] = 0u;
    }
    for (uint32_t t = threadIdx.x; t < prep.n_textures; t += blockDim.x) prep.acc[(size_t)t * prep.acc_stride + (size_t)prep.K * 65] = 0;
    if (full)
        for (uint32_t t = 0; t < prep.n_textures; ++t)
            for (size_t i = threadIdx.x; i < (size_t)prep.K * 65; i += blockDim.x) prep.acc[(size_t)t * prep.acc_stride + i] = 0;
}

// ---- host side -------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn() {
    static const EncodeTiledFn fn = [] {   // resolved once, thread-safe (C++11 static initialisation)
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            return reinterpret_cast<EncodeTiledFn>(p);
        return (EncodeTiledFn) nullptr;
    }();
    return fn;
}

// 2-D map over rows of `cols` bf16 elements
bool make_map(CUtensorMap* m, const float* ptr, uint64_t cols, uint64_t rows, uint32_t box_cols, uint32_t box_rows,
              CUtensorMapSwizzle sw) {
    EncodeTiledFn fn = encode_fn();
    if (!fn) return false;
    const cuuint64_t dims[2] = {cols, rows};
    const cuuint64_t strides[1] = {cols * 2};
    const cuuint32_t box[2] = {box_cols, box_rows};
    const cuuint32_t estr[2] = {1, 1};
    return fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<float*>(ptr), dims, strides, box, estr,
              CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
              CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

}  // namespace

uint32_t tc_padded_k(uint32_t K) { return (K + BN - 1) / BN * BN; }

bool tc_supported(uint32_t K) { return K >= 1 && K <= 65535 && encode_fn() != nullptr; }

cudaError_t launch_tc_mean(const float* feat, size_t n, double* sums64, float* mu, cudaStream_t s) {
    cudaError_t e = cudaMemsetAsync(sums64, 0, 64 * sizeof(double), s);
    if (e != cudaSuccess) return e;
    if (n > 0) {
        size_t blocks = (n + 3) / 4;
        if (blocks > (size_t)YUPVQ_NUM_SMS * 8) blocks = (size_t)YUPVQ_NUM_SMS * 8;
        col_sums_kernel<<<(unsigned)blocks, 256, 0, s>>>(feat, n, sums64);
    }
    mean_kernel<<<1, 64, 0, s>>>(sums64, n ? 1.0 / (double)n : 0.0, mu);
    return cudaGetLastError();
}

cudaError_t launch_tc_prep_blocks(const float* feat, size_t n, const float* mu, float* a_prime, float* xnorm, cudaStream_t s, int rgb) {
    if (n == 0) return cudaSuccess;
    prep_blocks_kernel<<<(unsigned)((n * 16 + 255) / 256), 256, 0, s>>>(feat, n, mu, a_prime, xnorm, rgb);
    return cudaGetLastError();
}

cudaError_t launch_tc_prep_centroids(const float* cent, uint32_t K, const float* mu, float* b_main, float* b_ext,
                                     float* cmax, cudaStream_t s, int rgb) {
    cudaError_t e = cudaMemsetAsync(cmax, 0, 4, s);
    if (e != cudaSuccess) return e;
    const uint32_t k_pad = tc_padded_k(K);
    prep_centroids_kernel<<<(k_pad + 7) / 8, 256, 0, s>>>(cent, K, k_pad, mu, b_main, b_ext, reinterpret_cast<unsigned int*>(cmax), 0, rgb);
    return cudaGetLastError();
}

cudaError_t launch_tc_prep_centroids_batch(const float* cent, uint32_t K, uint32_t n_textures, uint8_t* small, size_t small_stride,
                                           size_t mu_off, size_t cmax_off, float* b_main, float* b_ext, cudaStream_t s, int rgb) {
    cudaError_t e = cudaMemset2DAsync(small + cmax_off, small_stride, 0, 4, n_textures, s);
    if (e != cudaSuccess) return e;
    const uint32_t k_pad = tc_padded_k(K);
    prep_centroids_kernel<<<dim3((k_pad + 7) / 8, n_textures), 256, 0, s>>>(cent, K, k_pad, reinterpret_cast<const float*>(small + mu_off),
                                                                           b_main, b_ext, reinterpret_cast<unsigned int*>(small + cmax_off),
                                                                           small_stride, rgb);
    return cudaGetLastError();
}

cudaError_t launch_assign_tc(const float* a_prime, const float* xnorm, size_t n, const float* b_main, const float* b_ext,
                             uint32_t K, const float* cmax, uint4* cand, uint32_t* assign, uint32_t* flags, uint32_t* amb_list,
                             unsigned int* amb_count, const uint32_t* stop, cudaStream_t s, const TcBatch* batch) {
    if (n == 0) return cudaSuccess;
    static std::atomic<uint64_t> attr_set{0};  // per device (a process may hold contexts on several GPUs)
    {
        const cudaError_t e = once_per_device(attr_set, [] {
            cudaError_t a = cudaFuncSetAttribute(assign_tc_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_TC);
            if (a == cudaSuccess) a = cudaFuncSetAttribute(assign_tc_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_TC);
            if (a == cudaSuccess) a = cudaFuncSetAttribute(assign_tc_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_TC);
            if (a == cudaSuccess) a = cudaFuncSetAttribute(assign_tc_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_TC);
            return a;
        });
        if (e != cudaSuccess) return e;
    }
    const uint32_t k_pad = tc_padded_k(K);
    const uint32_t n_tex = batch ? batch->n_textures : 1u;
    if (batch && n_tex > 1 && (batch->rows_per_tex % BM != 0 || (size_t)batch->rows_per_tex * n_tex != n || batch->k_pad != k_pad))
        return cudaErrorInvalidValue;  // tiles must not straddle textures
    CUtensorMap tm_a, tm_b, tm_bx;
    if (!make_map(&tm_a, a_prime, 128, n, 64, BM, CU_TENSOR_MAP_SWIZZLE_128B) ||
        !make_map(&tm_b, b_main, 128, (uint64_t)k_pad * n_tex, 64, BN, CU_TENSOR_MAP_SWIZZLE_128B) ||
        !make_map(&tm_bx, b_ext, 16, (uint64_t)k_pad * n_tex, 16, BN, CU_TENSOR_MAP_SWIZZLE_32B))
        return cudaErrorInvalidValue;
    const uint32_t num_tiles = (uint32_t)((n + BM - 1) / BM);
    const uint32_t num_chunks = k_pad / BN;
    const uint32_t grid = num_tiles < (uint32_t)YUPVQ_NUM_SMS ? num_tiles : (uint32_t)YUPVQ_NUM_SMS;
    TcBatch bt = batch ? *batch : tc_single();
    if (!batch || n_tex <= 1) bt.k_pad = k_pad;
    const uint32_t res_b = num_chunks <= 2 ? 1u : 0u;
    const bool v1 = debug_opts().tc_v1;   // A/B: the one-barrier-per-tile epilogue without merge warps
#define YUPVQ_TC_LAUNCH(B, M)                                                                                                      \
    assign_tc_kernel<B, M><<<grid, (M) ? kThreadsTc2 : kThreadsTc, SMEM_TC, s>>>(tm_a, tm_b, tm_bx, xnorm, cmax, (uint32_t)n, num_tiles, \
                                                                                num_chunks, res_b, cand, assign, flags, amb_list,       \
                                                                                amb_count, stop, bt)
    const bool force_batch = debug_opts().tc_force_batch;   // diagnostic only (codegen A/B on one texture)
    if (force_batch && !(batch && n_tex > 1)) {
        bt.tiles_per_tex = num_tiles;
        bt.n_textures = 1;
        bt.rows_per_tex = (uint32_t)n;
    }
    if ((batch && n_tex > 1) || force_batch) {
        if (v1) YUPVQ_TC_LAUNCH(true, false);
        else YUPVQ_TC_LAUNCH(true, true);
    } else {
        if (v1) YUPVQ_TC_LAUNCH(false, false);
        else YUPVQ_TC_LAUNCH(false, true);
    }
#undef YUPVQ_TC_LAUNCH
    return cudaGetLastError();
}

cudaError_t launch_rerank_sparse(const float* feat, const float* cent, uint32_t K, const uint4* cand, const uint32_t* amb_list,
                                 unsigned int* amb_count, unsigned int* ticket, uint32_t* assign, uint32_t* flags,
                                 unsigned int* stat_pairs, cudaStream_t s, const TcBatch* batch, const AccPrep* prep) {
    rerank_sparse_kernel<<<YUPVQ_NUM_SMS * 2, kRrThreads, 0, s>>>(feat, cent, K, cand, amb_list, amb_count, ticket, assign, flags,
                                                                 stat_pairs, batch ? *batch : tc_single(),
                                                                 prep ? *prep : AccPrep{nullptr, nullptr, 0u, 0u, 0u, 0u, 0});
    return cudaGetLastError();
}

cudaError_t launch_rerank(const float* feat, size_t n, const float* cent, uint32_t K, const uint4* cand, uint32_t* assign,
                          float* errors, uint32_t* flags, unsigned int* stat_pairs, const uint32_t* stop, cudaStream_t s) {
    if (n == 0) return cudaSuccess;
    rerank_kernel<<<(unsigned)((n + kRrThreads - 1) / kRrThreads), kRrThreads, 0, s>>>(feat, n, cent, K, cand, assign, errors,
                                                                                      flags, stat_pairs, stop);
    return cudaGetLastError();
}

}  // namespace yupvq
