// C ABI (include/yupvq.h) and host orchestration of the VQ hot path.
// Replaces, behind plain C entry points, the three starred lines of the reference's
// VQBCnCompressor::Compress (src/vq_bcn_compressor.cpp:142-150) / CompressHDR (:230-256) and the
// four public VQEncoder calls (include/vq_encoder.h:76-103).  All arithmetic runs in CUDA kernels
// on the context's device; there is no CPU path.
#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <numeric>
#include <random>
#include <string>
#include <vector>

#include "../../include/yupvq.h"
#include "host_util.h"
#include "kernels.h"
#include "nccl_dyn.h"

using namespace yupvq;

namespace yupvq {
const DebugOpts& debug_opts() {
    static const DebugOpts opts = [] {   // parsed once (thread-safe static initialisation)
        DebugOpts o;
        auto on = [](const char* name) { return getenv(name) != nullptr; };
        o.seed_per_round = on("YUPVQ_SEED_PER_ROUND");
        o.seed_tri = on("YUPVQ_SEED_TRI");
        o.seed_no_tri = on("YUPVQ_SEED_NO_TRI");
        o.seed_no_pal = on("YUPVQ_SEED_NO_PAL");
        o.seed_no_smem = on("YUPVQ_SEED_NO_SMEM");
        o.lloyd_sync = on("YUPVQ_LLOYD_SYNC");
        o.no_delta = on("YUPVQ_NO_DELTA");
        o.no_rgb48 = on("YUPVQ_NO_RGB48");
        o.no_split_decode = on("YUPVQ_NO_SPLIT_DECODE");
        o.no_decode_fuse = on("YUPVQ_NO_DECODE_FUSE");
        o.batch_per_texture = on("YUPVQ_BATCH_PER_TEXTURE");
        o.no_tail_fuse = on("YUPVQ_NO_TAIL_FUSE");
        o.no_peer_xchg = on("YUPVQ_NO_PEER_XCHG");
        o.tc_force_batch = on("YUPVQ_TC_FORCE_BATCH");
        if (const char* v = getenv("YUPVQ_XCHG_MODE")) o.xchg_mode = atoi(v);
        o.tc_v1 = on("YUPVQ_TC_V1");   // v2 (merge warps) is the default since r02c: 0.143 against 0.177 ms per pass, bit-identical
        if (const char* v = getenv("YUPVQ_XCHG_TIMEOUT_MS")) o.xchg_timeout_ms = std::max(1, atoi(v));
        if (const char* v = getenv("YUPVQ_XCHG_DESYNC_RANK")) o.xchg_desync_rank = atoi(v);
        if (const char* v = getenv("YUPVQ_ASYNC_DBG")) o.async_dbg = atoi(v);
        return o;
    }();
    return opts;
}
}  // namespace yupvq

namespace {

thread_local std::string g_create_error;

struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
    cudaError_t ensure(size_t bytes) {
        if (bytes <= cap) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
        // round up so repeated slightly-larger requests do not thrash the allocator
        size_t want = (bytes + (1u << 20) - 1) & ~(size_t)((1u << 20) - 1);
        cudaError_t e = cudaMalloc(&p, want);
        if (e == cudaSuccess) cap = want;
        return e;
    }
    void release() {
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
    }
    template <typename T>
    T* as() const { return reinterpret_cast<T*>(p); }
};

}  // namespace

struct yupvq_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    std::string err;
    cudaEvent_t ev[12] = {};
    cudaEvent_t ev_user[2] = {};
    cudaEvent_t ev_poll = nullptr;   // lagged convergence poll of the fused Lloyd loops
    uint32_t launches = 0;
    uint32_t reseeds_sharded = 0;

    // device arena (grow-only, reused across calls)
    DevBuf in_bytes;    // staged input (BCn bytes / pixel blocks / float image or blocks)
    DevBuf img_in, bc_enc;  // N4: staged source image and the BCn blocks the GPU encoder writes
    DevBuf planes;      // uint8 pixel planes (ch planes of uint4)
    DevBuf pal5;        // BC1 input, 3 channels: palette form of the blocks for the seeding (5 planes of uint32)
    DevBuf feat;        // N x 64 fp32 features
    DevBuf s_planes, s_feat, s_rows;  // fast-mode sampled subset
    DevBuf min_dist, partial, seed_idx, canon, small;  // seeding
    DevBuf assign, errors, cent, acc;                  // Lloyd
    DevBuf cent_px, cent_feat, out_idx;                // finalize / quantize
    // tensor engine (assign_tc.cu): per block set A', |x'|, mu; per pass B', folded norms, candidates
    DevBuf tc_a, tc_xnorm, tc_small, tc_b, tc_bx, tc_cand, tc_amb;
    DevBuf tail_scratch;       // fused pass tail: one (max error, index) slot per CTA for the grid-wide reseed arg-max
    DevBuf tc_chg, delta_ctl;  // delta update: (row, previous centroid) lists; counter / mode words
    // lock-step batch (yupvq_encode_bcn_batch_sharded): per-texture slices of one buffer each
    DevBuf bt_feat, bt_tca, bt_xnorm, bt_tcsmall, bt_assign, bt_errors, bt_cent, bt_acc, bt_misc, bt_centpx, bt_cent0, bt_acc_local;
    DevBuf bt_planes, bt_in, bt_rows;   // general lock-step batch: SAD pixel planes, device-resident inputs, row lists
    const float* tc_feat = nullptr;                    // block set the A' buffer was built from
    const float* maxabs_feat = nullptr;                // block set whose max |x| already sits in small[F_MAXABS] (fused feature pass)
    int tc_rgb = 0;                                    // 1: channels are 1 or 3 (alpha feature constant): 48-column tensor operands
    size_t tc_n = 0;
    uint64_t rerank_pairs = 0;
    uint32_t* h_small = nullptr;                       // pinned: flags + status readback
    // optional per-launch timing of the tensor front-end (yupvq_profile): event pairs around each launch
    bool profile = false;
    std::vector<cudaEvent_t> pev;
    size_t pev_used = 0;
    // stage marks of the fused Lloyd pass while profiling (yupvq_profile_stages): event + the stage that ended at it
    std::vector<cudaEvent_t> sev;
    std::vector<uint8_t> sev_stage;
    size_t sev_used = 0;
    void stage_mark(int stage) {   // 255 = start of a pass; 0 front-end, 1 re-rank, 2 update, 3 exchange + tail
        if (!profile || sev_used >= sev.size()) return;
        cudaEventRecord(sev[sev_used], stream);
        sev_stage[sev_used++] = (uint8_t)stage;
    }

    // multi-GPU
    NcclApi nccl;
    void* comm = nullptr;
    int n_ranks = 1, rank = 0;
    // peer exchange (XchgArgs, kernels.h): one device buffer per rank -- [flags 4 KB | 2 x staging | 2 x result] -- mapped into
    // every peer with CUDA IPC; xchg_state: 0 = not tried, 1 = usable, -1 = unavailable (the NCCL path is used)
    void* xchg_base = nullptr;
    void* xchg_peer[kXchgMaxRanks] = {};
    size_t xchg_cap = 0;              // capacity in int64 words
    uint32_t xchg_epoch = 0;
    int xchg_state = 0;
    int xchg_path = YUPVQ_XCHG_NONE;  // how the last sharded call exchanged its partial sums (yupvq_exchange_path)

    int fail(const char* fmt, ...) {
        char buf[512];
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(buf, sizeof(buf), fmt, ap);
        va_end(ap);
        err = buf;
        return 1;
    }
};

#define CK(call)                                                                                        \
    do {                                                                                                \
        cudaError_t e_ = (call);                                                                        \
        if (e_ != cudaSuccess) return ctx->fail("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), \
                                                __FILE__, __LINE__);                                    \
    } while (0)
#define LAUNCH(call)     \
    do {                 \
        CK(call);        \
        ++ctx->launches; \
    } while (0)

namespace {

// small device scratch layout (uint32 words)
enum { F_CHANGED = 0, F_EMPTY = 1, F_RESEEDS = 2, F_SPARE = 3, F_SEED_STATUS = 4 /*2 words*/, F_COUNTER = 6, F_MAXABS = 7,
       F_PAIRS = 8 /*2 words*/, F_CTL = 10 /*stop, passes, empty clusters*/, F_AMB = 13 /*count, ticket*/,
       F_WORDS = 16 };
// tc_small layout (bytes): [0,512) column sums fp64 | [512,768) mu fp32 | [768,776) cmax (two words: double-buffered by
// pass parity in the fused Lloyd loop, word 0 otherwise)
enum { TC_SUMS = 0, TC_MU = 512, TC_CMAX = 768, TC_SMALL_BYTES = 1024 };

bool is_device_ptr(const void* p) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}

// Returns a device pointer for `src` (staging host memory through `stage`).
int to_device(yupvq_ctx* ctx, const void* src, size_t bytes, DevBuf& stage, const void** out) {
    if (is_device_ptr(src)) {
        *out = src;
        return 0;
    }
    CK(stage.ensure(bytes));
    CK(cudaMemcpyAsync(stage.p, src, bytes, cudaMemcpyHostToDevice, ctx->stream));
    *out = stage.p;
    return 0;
}

int from_device(yupvq_ctx* ctx, void* dst, const void* src_dev, size_t bytes) {
    if (!dst || bytes == 0) return 0;
    CK(cudaMemcpyAsync(dst, src_dev, bytes, is_device_ptr(dst) ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost,
                       ctx->stream));
    return 0;
}

size_t bc_block_bytes(int fmt) { return (fmt == YUPVQ_BC1 || fmt == YUPVQ_BC4) ? 8 : 16; }

bool bc_supported(int fmt) {
    return fmt == YUPVQ_BC1 || fmt == YUPVQ_BC3 || fmt == YUPVQ_BC4 || fmt == YUPVQ_BC5 || fmt == YUPVQ_BC7;
}

struct Timer {
    yupvq_ctx* ctx;
    int next = 0;
    explicit Timer(yupvq_ctx* c) : ctx(c) {}
    int mark() {
        cudaEventRecord(ctx->ev[next], ctx->stream);
        return next++;
    }
    float ms(int a, int b) {
        float v = 0.f;
        cudaEventElapsedTime(&v, ctx->ev[a], ctx->ev[b]);
        return v;
    }
};

// The RNG stream of BuildCodebook: optional shuffle, one uniform_int draw,
// then one canonical double per seeding round.  Host code uses the same libstdc++ <random> as the
// reference, so the draws are bit-identical for the same seed.
struct Draws {
    std::vector<uint32_t> sample;  // empty -> all blocks in order
    uint32_t first = 0;
    std::vector<double> canon;
};

Draws make_draws(size_t n_all, uint32_t K, float ratio, uint64_t seed, size_t* n_used) {
    Draws d;
    std::mt19937 rng;
    if (seed == 0) rng.seed(std::random_device{}());
    else rng.seed((uint32_t)seed);
    size_t n = n_all;
    if (ratio < 1.0f && ratio > 0.0f) {  // src/vq_encoder.cpp:263-276
        size_t m = (size_t)(n_all * ratio);
        m = std::max((size_t)K, m);
        m = std::min(m, n_all);
        std::vector<size_t> idx(n_all);
        std::iota(idx.begin(), idx.end(), (size_t)0);
        std::shuffle(idx.begin(), idx.end(), rng);
        d.sample.resize(m);
        for (size_t i = 0; i < m; ++i) d.sample[i] = (uint32_t)idx[i];
        n = m;
    }
    std::uniform_int_distribution<size_t> distrib(0, n - 1);  // :289-290
    d.first = (uint32_t)distrib(rng);
    d.canon.resize(K > 1 ? K - 1 : 0);
    std::uniform_real_distribution<double> unit(0.0, 1.0);    // p = u * S + 0.0 (:303-304)
    for (auto& u : d.canon) u = unit(rng);
    *n_used = n;
    return d;
}

struct Blocks {        // device-resident block set
    size_t n = 0;
    int ch = 0;
    const uint4* planes = nullptr;  // null for HDR
    size_t stride = 0;
    const float* feat = nullptr;
    const uint32_t* pal5 = nullptr;  // optional: the same blocks in BC1 palette form (seeding only), 5 planes
    size_t pal_stride = 0;
};

int fixed_shift(size_t n_total, float max_abs) {
    // sum of n_total values of magnitude <= max_abs must stay below 2^62
    int bits_n = 1;
    while (((size_t)1 << bits_n) < n_total && bits_n < 40) ++bits_n;
    int e = 0;
    if (max_abs > 0.f && std::isfinite(max_abs)) {
        (void)std::frexp(max_abs, &e);  // max_abs < 2^e
    } else if (!std::isfinite(max_abs)) {
        e = 128;
    }
    int shift = 62 - bits_n - std::max(e, 0);
    return std::max(-60, std::min(shift, 40));
}

// ---- K2: seeding -----------------------------------------------------------------------------------
// BC1 input with 3-channel output: the seeding works from the 20-byte palette form of the blocks, which fits in
// shared memory whole (seed.cu, PAL variant).  `all` must describe exactly the n blocks at bc_dev.
int attach_palette_form(yupvq_ctx* ctx, Blocks& all, const void* bc_dev, int bc_format, int channels) {
    if (bc_format != YUPVQ_BC1 || channels != 3 || all.n == 0) return 0;
    CK(ctx->pal5.ensure(all.n * 20));
    LAUNCH(launch_bc1_palette_planes(reinterpret_cast<const uint8_t*>(bc_dev), all.n, ctx->pal5.as<uint32_t>(), all.n, ctx->stream));
    all.pal5 = ctx->pal5.as<uint32_t>();
    all.pal_stride = all.n;
    return 0;
}

int run_seeding(yupvq_ctx* ctx, const Blocks& b, uint32_t K, const Draws& dr, bool hdr, uint32_t* out_rounds) {
    const size_t n = b.n;
    const int grid = (int)seed_grid_for(n);
    CK(ctx->min_dist.ensure(n * sizeof(float)));
    CK(ctx->partial.ensure(seed_partial_bytes(grid)));
    CK(ctx->seed_idx.ensure((size_t)K * 4));
    CK(ctx->canon.ensure(std::max<size_t>(1, dr.canon.size()) * 8));
    if (!dr.canon.empty())
        CK(cudaMemcpyAsync(ctx->canon.p, dr.canon.data(), dr.canon.size() * 8, cudaMemcpyHostToDevice, ctx->stream));
    SeedState st;
    st.min_dist = ctx->min_dist.as<float>();
    st.partial = ctx->partial.as<unsigned long long>();
    st.counter = ctx->small.as<unsigned int>() + F_COUNTER;
    st.seed_idx = ctx->seed_idx.as<uint32_t>();
    st.canon = ctx->canon.as<double>();
    st.status = ctx->small.as<uint32_t>() + F_SEED_STATUS;
    st.grid = grid;
    LAUNCH(launch_seed_init(st, n, dr.first, ctx->stream));
    // one persistent cooperative kernel for all K-1 rounds; per-round launches only where that is not possible
    const bool per_round = debug_opts().seed_per_round;  // A/B timing and tests
    cudaError_t pe = per_round ? cudaErrorNotSupported
                               : launch_seed_persistent(st, b.planes, b.stride, b.feat, n, b.ch, hdr, K, ctx->stream, b.pal5, b.pal_stride);
    if (pe == cudaSuccess) {
        if (K > 1) ++ctx->launches;
    } else if (pe == cudaErrorNotSupported || pe == cudaErrorCooperativeLaunchTooLarge) {
        cudaGetLastError();  // not all CTAs can be co-resident (shared GPU, MPS limits): one launch per round instead
        for (uint32_t i = 1; i < K; ++i) {
            if (hdr) LAUNCH(launch_seed_round_hdr(st, b.feat, n, i, K, ctx->stream));
            else LAUNCH(launch_seed_round_ldr(st, b.planes, b.stride, n, b.ch, i, K, ctx->stream));
        }
    } else {
        CK(pe);
    }
    if (out_rounds) {
        CK(cudaMemcpyAsync(ctx->h_small, st.status, 8, cudaMemcpyDeviceToHost, ctx->stream));
        CK(cudaStreamSynchronize(ctx->stream));
        *out_rounds = ctx->h_small[1];
    }
    return 0;
}

// Empty-cluster reseed when blocks are sharded (rare path, host-driven): for every empty cluster in
// ascending order each rank proposes its local (max error, first index); the global winner is the
// largest error with the lowest global index -- exactly the single-GPU std::max_element result --
// and its owner broadcasts the feature row and zeroes that error.
int reseed_sharded(yupvq_ctx* ctx, const Blocks& b, size_t first_block, uint32_t K, float* cent, float* errors,
                   const long long* acc, uint32_t* flags) {
    CK(cudaMemcpyAsync(ctx->h_small, flags + F_EMPTY, 4, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    if (ctx->h_small[0] == 0) return 0;
    std::vector<long long> counts(K);
    CK(cudaMemcpyAsync(counts.data(), acc + (size_t)K * 64, (size_t)K * 8, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    const int R = ctx->n_ranks;
    CK(ctx->s_rows.ensure((size_t)(R + 1) * 16 + 256));
    uint32_t* prop = ctx->s_rows.as<uint32_t>();          // [R+1][4]: slot R is this rank's proposal
    float* row = reinterpret_cast<float*>(prop + (size_t)(R + 1) * 4);
    std::vector<uint32_t> h((size_t)R * 4);
    for (uint32_t c = 0; c < K; ++c) {
        if (counts[c] != 0) continue;
        uint32_t mine[4] = {0, 0xffffffffu, 0, 0};        // value bits, global index, valid, pad
        if (b.n > 0) {
            LAUNCH(launch_argmax_error(errors, b.n, prop + (size_t)R * 4, ctx->stream));
            CK(cudaMemcpyAsync(mine, prop + (size_t)R * 4, 8, cudaMemcpyDeviceToHost, ctx->stream));
            CK(cudaStreamSynchronize(ctx->stream));
            mine[1] += (uint32_t)first_block;
            mine[2] = 1;
        }
        CK(cudaMemcpyAsync(prop + (size_t)R * 4, mine, 16, cudaMemcpyHostToDevice, ctx->stream));
        if (ctx->nccl.all_gather_bytes(ctx->comm, prop + (size_t)R * 4, prop, 16, ctx->stream)) return ctx->fail("ncclAllGather failed");
        CK(cudaMemcpyAsync(h.data(), prop, (size_t)R * 16, cudaMemcpyDeviceToHost, ctx->stream));
        CK(cudaStreamSynchronize(ctx->stream));
        int owner = -1;
        float bv = 0.f;
        uint32_t bi = 0;
        for (int r = 0; r < R; ++r) {
            if (!h[(size_t)r * 4 + 2]) continue;
            float v;
            std::memcpy(&v, &h[(size_t)r * 4], 4);
            const uint32_t gi = h[(size_t)r * 4 + 1];
            if (owner < 0 || v > bv || (v == bv && gi < bi)) {
                owner = r;
                bv = v;
                bi = gi;
            }
        }
        if (owner < 0) return ctx->fail("reseed: no rank holds any block");
        if (owner == ctx->rank) {
            const size_t local = (size_t)bi - first_block;
            CK(cudaMemcpyAsync(row, b.feat + local * 64, 256, cudaMemcpyDeviceToDevice, ctx->stream));
            CK(cudaMemsetAsync(errors + local, 0, 4, ctx->stream));  // errors[worst] = 0.0f
        }
        if (ctx->nccl.broadcast_bytes(ctx->comm, row, 256, owner, ctx->stream)) return ctx->fail("ncclBroadcast failed");
        CK(cudaMemcpyAsync(cent + (size_t)c * 64, row, 256, cudaMemcpyDeviceToDevice, ctx->stream));
        ctx->reseeds_sharded++;
    }
    return 0;
}

// ---- peer exchange buffers ------------------------------------------------------------------------
constexpr size_t kXchgFlagBytes = 4096;
size_t xchg_slot_stride(size_t cap, int) { return cap; }   // a slot holds a whole vector (one-shot mode); a slice in two-shot mode
size_t xchg_bytes(size_t cap, int R) { return kXchgFlagBytes + 2 * (size_t)R * xchg_slot_stride(cap, R) * 8 + 2 * cap * 8; }

void xchg_release(yupvq_ctx* ctx) {
    for (int r = 0; r < kXchgMaxRanks; ++r) {
        if (ctx->xchg_peer[r] && r != ctx->rank) cudaIpcCloseMemHandle(ctx->xchg_peer[r]);
        ctx->xchg_peer[r] = nullptr;
    }
    if (ctx->xchg_base) cudaFree(ctx->xchg_base);
    ctx->xchg_base = nullptr;
    ctx->xchg_cap = 0;
}

// Collective over the communicator (every rank calls it with the same `words`): makes sure each rank owns an exchange
// buffer of at least `words` int64 and has every peer's buffer mapped.  Returns 0 and sets *ok; *ok == false (on all ranks
// alike) means peer mapping is not available here and the caller keeps the NCCL all-reduce.
int xchg_ensure(yupvq_ctx* ctx, size_t words, bool* ok) {
    *ok = false;
    const int R = ctx->n_ranks;
    if (!ctx->comm || R < 2 || R > kXchgMaxRanks || ctx->xchg_state < 0 || debug_opts().no_peer_xchg) return 0;
    if (ctx->xchg_state == 1 && words <= ctx->xchg_cap) {
        *ok = true;
        return 0;
    }
    // (re)allocate: all ranks decide alike (same words, same history).  The collectives below double as the barrier that
    // makes it safe to drop the old mapping: nobody is still inside a kernel that touches it.
    CK(cudaStreamSynchronize(ctx->stream));
    size_t cap = 1;
    while (cap < words) cap <<= 1;
    uint32_t status = 0;   // 1 = this rank failed
    void* base = nullptr;
    // The zeroing (counters included) must be complete before any peer can learn the handle: cudaMemset runs on the legacy
    // default stream, which the context's non-blocking stream and the collectives below do not wait for, hence the
    // device-wide synchronize.
    if (cudaMalloc(&base, xchg_bytes(cap, R)) != cudaSuccess || cudaMemset(base, 0, xchg_bytes(cap, R)) != cudaSuccess ||
        cudaDeviceSynchronize() != cudaSuccess) {
        cudaGetLastError();
        status = 1;
    }
    // what every rank tells the others: the handle of its buffer, the grid its tail kernel runs with (the counters reach
    // epoch * grid, so all ranks must use the same grid) and where its GPU sits (to ask about native peer atomics)
    struct XchgHello {
        cudaIpcMemHandle_t handle;
        int32_t tail_grid;
        char pci[28];
    };
    static_assert(sizeof(cudaIpcMemHandle_t) == 64 && sizeof(XchgHello) == 96, "record size");
    XchgHello mine;
    std::memset(&mine, 0, sizeof(mine));
    if (!status && cudaIpcGetMemHandle(&mine.handle, base) != cudaSuccess) {
        cudaGetLastError();
        status = 1;
    }
    {
        int sms = 0;
        if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, ctx->device) != cudaSuccess) cudaGetLastError();
        mine.tail_grid = sms;   // launch_lloyd_tail: one CTA per SM
        if (cudaDeviceGetPCIBusId(mine.pci, (int)sizeof(mine.pci), ctx->device) != cudaSuccess) {
            cudaGetLastError();
            mine.pci[0] = 0;
        }
    }
    constexpr size_t kRec = sizeof(XchgHello);
    CK(ctx->s_rows.ensure((size_t)(R + 1) * kRec + 64));
    uint8_t* hbuf = ctx->s_rows.as<uint8_t>();          // [R records | mine | status word]
    CK(cudaMemcpyAsync(hbuf + (size_t)R * kRec, &mine, kRec, cudaMemcpyHostToDevice, ctx->stream));
    if (ctx->nccl.all_gather_bytes(ctx->comm, hbuf + (size_t)R * kRec, hbuf, kRec, ctx->stream)) return ctx->fail("ncclAllGather failed");
    std::vector<XchgHello> all(R);
    CK(cudaMemcpyAsync(all.data(), hbuf, (size_t)R * kRec, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    xchg_release(ctx);   // every rank has passed the all-gather: the old buffers are idle everywhere
    void* peer[kXchgMaxRanks] = {};
    for (int r = 0; r < R && !status; ++r) {
        if (r == ctx->rank) {
            peer[r] = base;
            continue;
        }
        if (all[r].tail_grid != mine.tail_grid || mine.tail_grid <= 0) {   // different SM counts: the counters would never meet
            status = 1;
            break;
        }
        // red.release.sys on a peer's counters needs native atomics over the link; ask when the peer's GPU is visible to this
        // process (it is under torchrun; when it is not, the mapping below is the only test there is)
        all[r].pci[sizeof(all[r].pci) - 1] = 0;
        int peer_dev = -1, native = 1;
        if (all[r].pci[0] && cudaDeviceGetByPCIBusId(&peer_dev, all[r].pci) == cudaSuccess && peer_dev != ctx->device) {
            if (cudaDeviceGetP2PAttribute(&native, cudaDevP2PAttrNativeAtomicSupported, ctx->device, peer_dev) != cudaSuccess) native = 1;
        }
        cudaGetLastError();
        if (!native) {
            status = 1;
            break;
        }
        if (cudaIpcOpenMemHandle(&peer[r], all[r].handle, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
            cudaGetLastError();
            peer[r] = nullptr;
            status = 1;
        }
    }
    // agree on the outcome
    uint32_t* st_d = reinterpret_cast<uint32_t*>(hbuf + (size_t)(R + 1) * kRec);
    CK(cudaMemcpyAsync(st_d, &status, 4, cudaMemcpyHostToDevice, ctx->stream));
    if (ctx->nccl.all_reduce_max_u32(ctx->comm, st_d, 1, ctx->stream)) return ctx->fail("ncclAllReduce(max) failed");
    uint32_t any = 0;
    CK(cudaMemcpyAsync(&any, st_d, 4, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    if (any) {
        for (int r = 0; r < R; ++r)
            if (peer[r] && r != ctx->rank) cudaIpcCloseMemHandle(peer[r]);
        if (base) cudaFree(base);
        ctx->xchg_state = -1;
        return 0;
    }
    ctx->xchg_base = base;
    for (int r = 0; r < R; ++r) ctx->xchg_peer[r] = peer[r];
    ctx->xchg_cap = cap;
    ctx->xchg_epoch = 0;   // fresh counters on every rank
    ctx->xchg_state = 1;
    *ok = true;
    return 0;
}

// XchgArgs of the next exchange of `words` int64 from `src`; *result receives this rank's result buffer (the tail's acc)
void xchg_args(yupvq_ctx* ctx, const long long* src, size_t words, XchgArgs* x, const long long** result) {
    const int R = ctx->n_ranks;
    const size_t stride = xchg_slot_stride(ctx->xchg_cap, R);
    x->n_ranks = R;
    x->rank = ctx->rank;
    x->epoch = ++ctx->xchg_epoch;
    const uint32_t par = x->epoch & 1u;
    x->words = words;
    // one shot while the (R - 1) copies of the vector are a few MB (their NVLink time stays below a second phase's latency)
    const int forced = debug_opts().xchg_mode;
    x->one_shot = forced ? (forced == 1) : ((size_t)(R - 1) * words * 8 <= (4u << 20));
    x->slice = (words + R - 1) / R;
    x->slot_stride = stride;
    x->src = src;
    for (int r = 0; r < R; ++r) {
        uint8_t* b = static_cast<uint8_t*>(ctx->xchg_peer[r]);
        x->peer_flags[r] = reinterpret_cast<uint32_t*>(b);
        x->peer_stage[r] = reinterpret_cast<long long*>(b + kXchgFlagBytes) + (size_t)par * R * stride;
        x->peer_result[r] = reinterpret_cast<long long*>(b + kXchgFlagBytes) + 2 * (size_t)R * stride + (size_t)par * ctx->xchg_cap;
    }
    x->err = x->peer_flags[ctx->rank] + kXchgErrWord;
    x->timeout_ns = (unsigned long long)debug_opts().xchg_timeout_ms * 1000000ull;
    *result = x->peer_result[R > 0 ? ctx->rank : 0];
}

// After the passes of a call have been synchronised: did one of its exchanges give up waiting?  Then the sums it read were
// not the peers' and the call fails; the peer path stays off for this communicator (a new yupvq_comm_init starts afresh).
int xchg_check(yupvq_ctx* ctx) {
    if (ctx->xchg_state != 1 || !ctx->xchg_base) return 0;
    uint32_t err = 0;
    CK(cudaMemcpyAsync(&err, static_cast<uint32_t*>(ctx->xchg_base) + kXchgErrWord, 4, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    if (!err) return 0;
    ctx->xchg_state = -1;
    return ctx->fail("peer exchange timed out after %d ms waiting for rank %u (phase %u, exchange %u): a peer failed, exited or "
                     "launched a different sequence of passes",
                     debug_opts().xchg_timeout_ms, (err - 1u) % kXchgMaxRanks, (err - 1u) / kXchgMaxRanks, ctx->xchg_epoch);
}

// ---- K3 dispatch: exact engine or tcgen05 front-end + exact re-rank -----------------------------------
bool use_tensor(int mode, size_t n, uint32_t K) {
    if (mode == YUPVQ_ASSIGN_EXACT) return false;
    if (!tc_supported(K)) return false;
    if (mode == YUPVQ_ASSIGN_TENSOR) return true;
    return K >= 64 && n >= 8192;  // below that the exact kernel is already launch-latency sized
}

// Builds mu / A' / |x'| for a block set (cached per call: the Lloyd passes and the final quantise
// share them).
int tc_prepare_blocks(yupvq_ctx* ctx, const float* feat, size_t n) {
    if (ctx->tc_feat == feat && ctx->tc_n == n) return 0;
    CK(ctx->tc_small.ensure(TC_SMALL_BYTES));
    CK(ctx->tc_a.ensure(std::max<size_t>(1, n) * 256));
    CK(ctx->tc_xnorm.ensure(std::max<size_t>(1, n) * 4));
    uint8_t* sm = ctx->tc_small.as<uint8_t>();
    LAUNCH(launch_tc_mean(feat, n, reinterpret_cast<double*>(sm + TC_SUMS), reinterpret_cast<float*>(sm + TC_MU), ctx->stream));
    ++ctx->launches;  // launch_tc_mean is two kernels
    LAUNCH(launch_tc_prep_blocks(feat, n, reinterpret_cast<const float*>(sm + TC_MU), ctx->tc_a.as<float>(),
                                 ctx->tc_xnorm.as<float>(), ctx->stream, ctx->tc_rgb));
    ctx->tc_feat = feat;
    ctx->tc_n = n;
    return 0;
}

// One assignment pass.  errors may be null (tensor mode then skips the exact distance of blocks with
// a single candidate); flags[0] |= 1 when an assignment changed.
// `view` (lock-step batch): the per-texture A' / |x'| / mu buffers built by tc_prepare_view; null = the
// context's own, built on demand for `feat`.
struct TcView {
    float* a = nullptr;
    float* xnorm = nullptr;
    uint8_t* small = nullptr;  // TC_SMALL_BYTES
};

int tc_prepare_view(yupvq_ctx* ctx, const float* feat, size_t n, const TcView& v) {
    LAUNCH(launch_tc_mean(feat, n, reinterpret_cast<double*>(v.small + TC_SUMS), reinterpret_cast<float*>(v.small + TC_MU), ctx->stream));
    ++ctx->launches;
    LAUNCH(launch_tc_prep_blocks(feat, n, reinterpret_cast<const float*>(v.small + TC_MU), v.a, v.xnorm, ctx->stream, ctx->tc_rgb));
    return 0;
}

int assign_blocks(yupvq_ctx* ctx, const float* feat, size_t n, const float* cent, uint32_t K, int mode, uint32_t* assign,
                  float* errors, uint32_t* flags, const uint32_t* stop = nullptr, const TcView* view = nullptr,
                  const TcBatch* delta = nullptr) {
    if (mode == YUPVQ_ASSIGN_TENSOR && !tc_supported(K))
        return ctx->fail("tensor assignment needs 1 <= K <= 65535 and a driver with cuTensorMapEncodeTiled (K = %u)", K);
    if (!use_tensor(mode, n, K)) {
        LAUNCH(launch_assign_exact(feat, n, cent, K, assign, errors, flags, stop, ctx->stream));
        return 0;
    }
    if (!view && tc_prepare_blocks(ctx, feat, n)) return 1;
    const uint32_t k_pad = tc_padded_k(K);
    CK(ctx->tc_b.ensure((size_t)k_pad * 256));
    CK(ctx->tc_bx.ensure((size_t)k_pad * 32));
    CK(ctx->tc_cand.ensure(std::max<size_t>(1, n) * 16));
    uint8_t* sm = view ? view->small : ctx->tc_small.as<uint8_t>();
    float* tc_a = view ? view->a : ctx->tc_a.as<float>();
    float* tc_xnorm = view ? view->xnorm : ctx->tc_xnorm.as<float>();
    LAUNCH(launch_tc_prep_centroids(cent, K, reinterpret_cast<const float*>(sm + TC_MU), ctx->tc_b.as<float>(),
                                    ctx->tc_bx.as<float>(), reinterpret_cast<float*>(sm + TC_CMAX), ctx->stream, ctx->tc_rgb));
    TcBatch tcb = delta ? *delta : tc_single();
    tcb.rgb = (uint32_t)ctx->tc_rgb;
    uint32_t* small = ctx->small.as<uint32_t>();
    if (errors) {
        // errors[] wanted for every block: dense candidate lists + dense re-rank
        LAUNCH(launch_assign_tc(tc_a, tc_xnorm, n, ctx->tc_b.as<float>(), ctx->tc_bx.as<float>(),
                                K, reinterpret_cast<const float*>(sm + TC_CMAX), ctx->tc_cand.as<uint4>(), nullptr, nullptr, nullptr,
                                nullptr, stop, ctx->stream, &tcb));
        LAUNCH(launch_rerank(feat, n, cent, K, ctx->tc_cand.as<uint4>(), assign, errors, flags, small + F_PAIRS, stop, ctx->stream));
    } else {
        // the front-end finishes single-candidate blocks itself; the exact re-rank sees only the ambiguous ones
        CK(ctx->tc_amb.ensure(std::max<size_t>(1, n) * 4));
        if (!stop) CK(cudaMemsetAsync(small + F_AMB, 0, 8, ctx->stream));  // inside the Lloyd loop the re-rank leaves it cleared
        const bool prof = ctx->profile && ctx->pev_used + 2 <= ctx->pev.size();
        if (prof) cudaEventRecord(ctx->pev[ctx->pev_used++], ctx->stream);
        LAUNCH(launch_assign_tc(tc_a, tc_xnorm, n, ctx->tc_b.as<float>(), ctx->tc_bx.as<float>(),
                                K, reinterpret_cast<const float*>(sm + TC_CMAX), ctx->tc_cand.as<uint4>(), assign, flags,
                                ctx->tc_amb.as<uint32_t>(), small + F_AMB, stop, ctx->stream, &tcb));
        if (prof) cudaEventRecord(ctx->pev[ctx->pev_used++], ctx->stream);
        LAUNCH(launch_rerank_sparse(feat, cent, K, ctx->tc_cand.as<uint4>(), ctx->tc_amb.as<uint32_t>(), small + F_AMB,
                                    small + F_AMB + 1, assign, flags, small + F_PAIRS, ctx->stream, &tcb));
    }
    return 0;
}

// ---- K3..K5: Lloyd iterations ----------------------------------------------------------------------
// break_after: LDR breaks on (!changed && iter > 0), HDR on (!changed && iter > 1)
// (src/vq_encoder.cpp:341 vs :572).  cent (device, K x 64) holds the seeds on entry.
// Always rank-local: ctx->comm is used by the *_sharded entry points only (the lock-step batch), so a context with a
// communicator can still encode independent textures.
int run_lloyd(yupvq_ctx* ctx, const Blocks& b, size_t n_total, uint32_t K, uint32_t max_iter, uint32_t break_after, int assign_mode,
              float* cent, yupvq_stats* stats) {
    const size_t n = b.n;
    const bool tensor = use_tensor(assign_mode, n, K);
    CK(ctx->assign.ensure(std::max<size_t>(1, n) * 4));
    CK(ctx->errors.ensure(std::max<size_t>(1, n) * 4));
    const size_t acc_words = (size_t)K * 65 + 1;  // [K*64 sums | K counts | changed]
    CK(ctx->acc.ensure(acc_words * 8));
    uint32_t* flags = ctx->small.as<uint32_t>();
    uint32_t* assign = ctx->assign.as<uint32_t>();
    float* errors = ctx->errors.as<float>();
    long long* acc = ctx->acc.as<long long>();
    CK(cudaMemsetAsync(assign, 0, n * 4, ctx->stream));  // assignments(numBlocks, 0), :316
    CK(cudaMemsetAsync(flags + F_RESEEDS, 0, 4, ctx->stream));
    CK(cudaMemsetAsync(flags + F_PAIRS, 0, 8, ctx->stream));
    ctx->reseeds_sharded = 0;

    // fixed-point scale from the data range (deterministic function of the data)
    if (ctx->maxabs_feat != b.feat) {   // else: the fused feature pass of this call already reduced it (same maximum, same bits)
        CK(cudaMemsetAsync(flags + F_MAXABS, 0, 4, ctx->stream));
        LAUNCH(launch_max_abs(b.feat, n * 64, flags + F_MAXABS, ctx->stream));
    }
    ctx->maxabs_feat = nullptr;
    CK(cudaMemcpyAsync(ctx->h_small, flags + F_MAXABS, 4, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    float max_abs;
    std::memcpy(&max_abs, ctx->h_small, 4);
    const int shift = fixed_shift(n_total, max_abs);

    uint32_t passes = 0;
    const DebugOpts& dbg = debug_opts();
    const bool async_loop = !dbg.lloyd_sync;
    if (async_loop) {
        // Single GPU: every pass is enqueued without a host round trip.  The reference's convergence test
        // (between assign and update) runs on the device (launch_lloyd_control); once it fires, the kernels
        // of that pass' update and of all later passes return immediately.  The host looks at the stop word
        // only a few times, to avoid enqueueing dozens of no-op passes for textures that converge at once.
        uint32_t* ctl = flags + F_CTL;
        const uint32_t* stopk = ctl;
        CK(cudaMemsetAsync(ctl, 0, 12 + 8, ctx->stream));  // stop, passes, empty | ambiguous-row count, ticket
        CK(cudaMemsetAsync(flags, 0, 8, ctx->stream));  // changed, empty
        // Delta update (tensor engine): the assignment kernels list the rows that moved; from the second pass on the
        // update touches only those (exact: integer sums), unless more than a quarter of the blocks moved.
        const bool delta = tensor && !dbg.no_delta && n < 0xffffffffull;
        TcBatch dl = tc_single();
        uint32_t* dctl = nullptr;
        if (delta) {
            CK(ctx->tc_chg.ensure(std::max<size_t>(1, n) * 8));
            CK(ctx->delta_ctl.ensure(256));
            dctl = ctx->delta_ctl.as<uint32_t>();
            CK(cudaMemsetAsync(dctl, 0, 256, ctx->stream));
            dl.chg_rows = ctx->tc_chg.as<uint32_t>();
            dl.chg_old = dl.chg_rows + n;
            dl.chg_count = dctl;
        }
        // Fused pass: four launches per pass -- front-end, re-rank (its last CTA prepares the update), one
        // update kernel (full or delta, chosen on the device) and one tail kernel (convergence test, finalize, reseed,
        // operands of the next pass).  Same arithmetic, same results as the separate kernels below (YUPVQ_NO_TAIL_FUSE=1).
        const bool fused_pass = delta && !dbg.no_tail_fuse && accumulate_auto_supported(n, K, 1u) && K <= 65535u;
        if (fused_pass) {
            if (tc_prepare_blocks(ctx, b.feat, n)) return 1;
            const uint32_t k_pad = tc_padded_k(K);
            CK(ctx->tc_b.ensure((size_t)k_pad * 256));
            CK(ctx->tc_bx.ensure((size_t)k_pad * 32));
            CK(ctx->tc_cand.ensure(std::max<size_t>(1, n) * 16));
            CK(ctx->tc_amb.ensure(std::max<size_t>(1, n) * 4));
            CK(ctx->tail_scratch.ensure(4096));
            uint8_t* sm = ctx->tc_small.as<uint8_t>();
            CK(cudaMemsetAsync(sm + TC_CMAX, 0, 8, ctx->stream));
            LAUNCH(launch_tc_prep_centroids(cent, K, reinterpret_cast<const float*>(sm + TC_MU), ctx->tc_b.as<float>(), ctx->tc_bx.as<float>(),
                                            reinterpret_cast<float*>(sm + TC_CMAX), ctx->stream, ctx->tc_rgb));
            dl.rgb = (uint32_t)ctx->tc_rgb;
            AccPrep prep{acc, dctl, K, (uint32_t)(n / 4), 0u, 1u, 0};
            TailArgs ta{};
            ta.acc = acc;
            ta.n_textures = 1;
            ta.K = K;
            ta.k_pad = k_pad;
            ta.break_after = break_after;
            ta.flags = flags;
            ta.flags_stride = 0;
            ta.ctl = ctl;
            ta.sticky = nullptr;
            ta.inv_scale = std::ldexp(1.0, -shift);
            ta.cent = cent;
            ta.small = sm;
            ta.mu_off = TC_MU;
            ta.cmax_off = TC_CMAX;
            ta.b_main = ctx->tc_b.as<float>();
            ta.b_ext = ctx->tc_bx.as<float>();
            ta.rgb = ctx->tc_rgb;
            ta.feat = b.feat;
            ta.n_rows = n;
            ta.assign = assign;
            ta.errors = errors;
            ta.scratch_val = ctx->tail_scratch.as<float>();
            ta.scratch_idx = reinterpret_cast<unsigned long long*>(ctx->tail_scratch.as<uint8_t>() + 2048);
            for (uint32_t iter = 0; iter < max_iter; ++iter) {
                const float* cmax = reinterpret_cast<const float*>(sm + TC_CMAX) + (iter & 1u);
                const bool prof = ctx->profile && ctx->pev_used + 2 <= ctx->pev.size();
                if (prof) cudaEventRecord(ctx->pev[ctx->pev_used++], ctx->stream);
                ctx->stage_mark(255);
                LAUNCH(launch_assign_tc(ctx->tc_a.as<float>(), ctx->tc_xnorm.as<float>(), n, ctx->tc_b.as<float>(), ctx->tc_bx.as<float>(), K, cmax,
                                        ctx->tc_cand.as<uint4>(), assign, flags, ctx->tc_amb.as<uint32_t>(), flags + F_AMB, stopk, ctx->stream, &dl));
                if (prof) cudaEventRecord(ctx->pev[ctx->pev_used++], ctx->stream);
                ctx->stage_mark(0);
                prep.force_full = iter == 0 ? 1u : 0u;
                LAUNCH(launch_rerank_sparse(b.feat, cent, K, ctx->tc_cand.as<uint4>(), ctx->tc_amb.as<uint32_t>(), flags + F_AMB, flags + F_AMB + 1,
                                            assign, flags, flags + F_PAIRS, ctx->stream, &dl, &prep));
                ctx->stage_mark(1);
                LAUNCH(launch_accumulate_auto(b.feat, n, assign, K, shift, acc, dl.chg_rows, dl.chg_old, dctl, flags, stopk, ctx->stream));
                ctx->stage_mark(2);
                ta.iter = iter;
                LAUNCH(launch_lloyd_tail(ta, ctx->stream));
                ctx->stage_mark(3);
                if (dbg.async_dbg & 1) CK(cudaStreamSynchronize(ctx->stream));
                // Lagged poll: the stop word after passes 3 / 7 / 15 is read back while the NEXT pass is already queued, so the
                // GPU never drains for it (a blocking read cost three pipeline bubbles per texture); a texture that has
                // converged runs one more no-op pass than it would need to.
                if ((iter == 3 || iter == 7 || iter == 15) && iter + 1 < max_iter) {
                    CK(cudaMemcpyAsync(ctx->h_small, ctl, 4, cudaMemcpyDeviceToHost, ctx->stream));
                    CK(cudaEventRecord(ctx->ev_poll, ctx->stream));
                }
                if (iter == 4 || iter == 8 || iter == 16) {
                    CK(cudaEventSynchronize(ctx->ev_poll));
                    if (ctx->h_small[0]) break;
                }
            }
        }
        for (uint32_t iter = 0; iter < max_iter && !fused_pass; ++iter) {
            // tensor mode computes errors[] lazily (they only feed the empty-cluster reseed)
            if (assign_blocks(ctx, b.feat, n, cent, K, assign_mode, assign, tensor ? nullptr : errors, flags, stopk, nullptr,
                              delta ? &dl : nullptr))
                return 1;
            if (delta) {
                LAUNCH(launch_acc_prepare(acc, K, n, iter == 0 ? 1u : 0u, dctl, stopk, ctx->stream));
                LAUNCH(launch_accumulate(b.feat, n, assign, K, shift, acc, flags, stopk, ctx->stream, dctl));
                LAUNCH(launch_accumulate_delta(b.feat, assign, K, shift, acc, dl.chg_rows, dl.chg_old, dctl, flags, stopk, ctx->stream));
            } else {
                CK(cudaMemsetAsync(acc, 0, acc_words * 8, ctx->stream));
                LAUNCH(launch_accumulate(b.feat, n, assign, K, shift, acc, flags, stopk, ctx->stream));
            }
            LAUNCH(launch_lloyd_control(acc, K, iter, break_after, flags, ctl, ctx->stream));
            if (tensor)  // errors[i] = distance to the assigned centroid, before the update (:336); only if a reseed needs them
                LAUNCH(launch_exact_errors(b.feat, n, cent, assign, errors, stopk, ctl + 2, ctx->stream));
            LAUNCH(launch_finalize_centroids(acc, K, shift, cent, flags, stopk, ctx->stream));
            LAUNCH(launch_reseed_empty(b.feat, n, errors, acc, K, cent, flags, stopk, ctx->stream));
            if (dbg.async_dbg & 1) CK(cudaStreamSynchronize(ctx->stream));
            if ((dbg.async_dbg & 8) && delta) {  // per-pass delta statistics (debug)
                uint32_t w[2];
                CK(cudaMemcpyAsync(w, dctl + 32, 8, cudaMemcpyDeviceToHost, ctx->stream));
                CK(cudaStreamSynchronize(ctx->stream));
                fprintf(stderr, "pass %u: moved %u of %zu rows, mode %s\n", iter, w[0], n, w[1] ? "full" : "delta");
            }
            if ((iter == 3 || iter == 7 || iter == 15) && iter + 1 < max_iter) {
                CK(cudaMemcpyAsync(ctx->h_small, ctl, 4, cudaMemcpyDeviceToHost, ctx->stream));
                CK(cudaStreamSynchronize(ctx->stream));
                if (ctx->h_small[0]) break;
            }
        }
        CK(cudaMemcpyAsync(ctx->h_small + 4, ctl + 1, 4, cudaMemcpyDeviceToHost, ctx->stream));
    } else {
        for (uint32_t iter = 0; iter < max_iter; ++iter) {
            CK(cudaMemsetAsync(flags, 0, 8, ctx->stream));  // changed, empty
            if (assign_blocks(ctx, b.feat, n, cent, K, assign_mode, assign, tensor ? nullptr : errors, flags)) return 1;
            ++passes;
            // The reference checks hasChanged between assign and update; the update is launched
            // speculatively so one sync and one all-reduce per pass carry both.
            CK(cudaMemsetAsync(acc, 0, acc_words * 8, ctx->stream));
            LAUNCH(launch_accumulate(b.feat, n, assign, K, shift, acc, flags, nullptr, ctx->stream));
            LAUNCH(launch_count_empty(acc, K, flags + F_SPARE, ctx->stream));
            CK(cudaMemcpyAsync(ctx->h_small, acc + (size_t)K * 65, 8, cudaMemcpyDeviceToHost, ctx->stream));
            CK(cudaMemcpyAsync(ctx->h_small + 2, flags + F_SPARE, 4, cudaMemcpyDeviceToHost, ctx->stream));
            CK(cudaStreamSynchronize(ctx->stream));
            long long changed;
            std::memcpy(&changed, ctx->h_small, 8);
            if (!changed && iter > break_after) break;
            if (tensor && ctx->h_small[2] != 0)  // errors[i] = distance to the assigned centroid, before the update (:336)
                LAUNCH(launch_exact_errors(b.feat, n, cent, assign, errors, nullptr, nullptr, ctx->stream));
            LAUNCH(launch_finalize_centroids(acc, K, shift, cent, flags, nullptr, ctx->stream));
            LAUNCH(launch_reseed_empty(b.feat, n, errors, acc, K, cent, flags, nullptr, ctx->stream));
        }
    }
    CK(cudaMemcpyAsync(ctx->h_small, flags + F_RESEEDS, 4, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaMemcpyAsync(ctx->h_small + 1, flags + F_PAIRS, 8, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    if (async_loop) passes = ctx->h_small[4];
    if (stats) {
        stats->lloyd_passes = passes;
        stats->reseeds = ctx->h_small[0] + ctx->reseeds_sharded;
        stats->rerank_blocks = ctx->h_small[1];    // exact distance evaluations spent on multi-candidate blocks
        stats->rerank_overflow = ctx->h_small[2];  // blocks whose candidate list was empty/overfull (full scan)
    }
    return 0;
}

// ---- A21: SAD-metric Lloyd (src/vq_encoder.cpp:380-430) ---------------------------------------------------------
// b.planes = the pixel planes, b.feat = the same bytes as integer-valued fp32 rows (launch_sad_expand); cent (K x 64
// fp32, integer-valued) holds the seed blocks on entry and the final pixel centroids on exit.  Every pass is
// enqueued without a host round trip, exactly like the Oklab loop; all arithmetic is integer, so the result is
// bit-identical to the reference.
int run_lloyd_sad(yupvq_ctx* ctx, const Blocks& b, uint32_t K, uint32_t max_iter, float* cent, yupvq_stats* stats) {
    const size_t n = b.n;
    CK(ctx->assign.ensure(std::max<size_t>(1, n) * 4));
    CK(ctx->errors.ensure(std::max<size_t>(1, n) * 4));
    const size_t acc_words = (size_t)K * 65 + 1;
    CK(ctx->acc.ensure(acc_words * 8));
    CK(ctx->cent_px.ensure((size_t)K * 16 * b.ch));
    uint32_t* flags = ctx->small.as<uint32_t>();
    uint32_t* ctl = flags + F_CTL;
    uint32_t* assign = ctx->assign.as<uint32_t>();
    float* errors = ctx->errors.as<float>();
    long long* acc = ctx->acc.as<long long>();
    CK(cudaMemsetAsync(assign, 0, n * 4, ctx->stream));  // assignments(numBlocks, 0), :316
    CK(cudaMemsetAsync(flags, 0, F_WORDS * 4, ctx->stream));
    for (uint32_t iter = 0; iter < max_iter; ++iter) {
        LAUNCH(launch_sad_pack(cent, K, b.ch, ctx->cent_px.as<uint8_t>(), ctx->stream));
        LAUNCH(launch_sad_assign(b.planes, b.stride, n, b.ch, ctx->cent_px.as<uint8_t>(), K, assign, errors, flags, ctl, ctx->stream));
        CK(cudaMemsetAsync(acc, 0, acc_words * 8, ctx->stream));
        LAUNCH(launch_accumulate(b.feat, n, assign, K, 0, acc, flags, ctl, ctx->stream));   // shift 0: exact integer sums
        LAUNCH(launch_lloyd_control(acc, K, iter, 0u, flags, ctl, ctx->stream));            // !hasChanged && iter > 0 (:395)
        LAUNCH(launch_sad_finalize(acc, K, cent, flags, ctl, ctx->stream));
        LAUNCH(launch_reseed_empty(b.feat, n, errors, acc, K, cent, flags, ctl, ctx->stream));  // :423-427, ascending c
        if ((iter == 3 || iter == 7 || iter == 15) && iter + 1 < max_iter) {
            CK(cudaMemcpyAsync(ctx->h_small, ctl, 4, cudaMemcpyDeviceToHost, ctx->stream));
            CK(cudaStreamSynchronize(ctx->stream));
            if (ctx->h_small[0]) break;
        }
    }
    LAUNCH(launch_sad_pack(cent, K, b.ch, ctx->cent_px.as<uint8_t>(), ctx->stream));
    CK(cudaMemcpyAsync(ctx->h_small, ctl + 1, 4, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaMemcpyAsync(ctx->h_small + 1, flags + F_RESEEDS, 4, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    if (stats) {
        stats->lloyd_passes = ctx->h_small[0];
        stats->reseeds = ctx->h_small[1];
    }
    return 0;
}

void init_stats(yupvq_stats* s) {
    if (s) std::memset(s, 0, sizeof(*s));
}

// Shared tail of the LDR paths: seeding + Lloyd + centroid -> pixel conversion on a device block set.
// `all` is the full block set; when sampling is active the subset is gathered first.
// Leaves K x 16*ch uint8 pixel centroids in ctx->cent_px.
// `late_features` (optional): builds all.feat AFTER the seeding -- LDR seeding reads only the pixel planes, and it runs
// faster while they are the only thing in L2 (the 256 MB feature write would evict them); used when no sampling is active.
int build_codebook_device(yupvq_ctx* ctx, const Blocks& all, size_t n_total, const yupvq_config* cfg, bool hdr,
                          yupvq_stats* stats, Timer& tm, int* t_seed0, int* t_seed1, int* t_lloyd1, int* t_fin1,
                          const std::function<int()>& late_features = nullptr) {
    const uint32_t K = cfg->codebook_size;
    size_t n_used = 0;
    Draws dr = make_draws(all.n, K, cfg->sample_ratio, cfg->seed, &n_used);
    Blocks b = all;
    if (!dr.sample.empty()) {
        CK(ctx->s_rows.ensure(n_used * 4));
        CK(cudaMemcpyAsync(ctx->s_rows.p, dr.sample.data(), n_used * 4, cudaMemcpyHostToDevice, ctx->stream));
        CK(ctx->s_feat.ensure(n_used * 256));
        if (!hdr) CK(ctx->s_planes.ensure(n_used * 16 * all.ch));
        LAUNCH(launch_gather_blocks(all.planes, all.stride, all.feat, ctx->s_rows.as<uint32_t>(), n_used, all.ch,
                                    hdr ? nullptr : ctx->s_planes.as<uint4>(), n_used, ctx->s_feat.as<float>(),
                                    ctx->stream));
        b.n = n_used;
        b.planes = hdr ? nullptr : ctx->s_planes.as<uint4>();
        b.stride = n_used;
        b.feat = ctx->s_feat.as<float>();
        b.pal5 = nullptr;  // the palette form describes the full set, not the gathered subset
        b.pal_stride = 0;
        n_total = n_used;
    }
    if (stats) {
        stats->n_blocks = (uint32_t)b.n;
        stats->codebook_size = K;
    }
    *t_seed0 = tm.mark();
    uint32_t rounds = 0;
    if (run_seeding(ctx, b, K, dr, hdr, stats ? &rounds : nullptr)) return 1;
    if (stats) stats->seed_rounds = rounds;
    if (late_features && late_features()) return 1;
    // seeds -> Oklab centroids: PixelBlockToOklabBlock(pixelCentroids[i]) (:324-326) equals the
    // feature row of the chosen block (same pixels, same conversion).
    CK(ctx->cent.ensure((size_t)K * 256));
    LAUNCH(launch_gather_rows_f32(b.feat, ctx->seed_idx.as<uint32_t>(), K, 64, ctx->cent.as<float>(), ctx->stream));
    *t_seed1 = tm.mark();
    const int metric = hdr ? YUPVQ_METRIC_OKLAB : cfg->metric;
    if (metric == YUPVQ_METRIC_SAD) {
        // the caller built b.feat with launch_sad_expand: the gathered seed rows are the seed pixel blocks
        if (run_lloyd_sad(ctx, b, K, cfg->max_iterations, ctx->cent.as<float>(), stats)) return 1;
        *t_lloyd1 = tm.mark();
        *t_fin1 = tm.mark();   // the centroids already are pixel blocks (no colour conversion on this path)
        return 0;
    }
    if (run_lloyd(ctx, b, n_total, K, cfg->max_iterations, hdr ? 1u : 0u, cfg->assign_mode, ctx->cent.as<float>(), stats))
        return 1;
    *t_lloyd1 = tm.mark();
    if (hdr) {
        CK(ctx->cent_px.ensure((size_t)K * 16 * all.ch * 4));
        LAUNCH(launch_oklab_to_floats(ctx->cent.as<float>(), K, all.ch, ctx->cent_px.as<float>(), ctx->stream));
    } else {
        CK(ctx->cent_px.ensure((size_t)K * 16 * all.ch));
        LAUNCH(launch_oklab_to_pixels(ctx->cent.as<float>(), K, all.ch, ctx->cent_px.as<uint8_t>(), ctx->stream));
    }
    *t_fin1 = tm.mark();
    return 0;
}

// QuantizeBlocks on device data: centroids (pixel/float, device) -> Oklab -> one assignment pass.
int quantize_device(yupvq_ctx* ctx, const Blocks& all, const void* cent_px_dev, uint32_t K, bool hdr, int assign_mode,
                    int metric = YUPVQ_METRIC_OKLAB) {
    if (!hdr && metric == YUPVQ_METRIC_SAD) {  // src/vq_encoder.cpp:473-484
        CK(ctx->out_idx.ensure(std::max<size_t>(1, all.n) * 4));
        LAUNCH(launch_sad_assign(all.planes, all.stride, all.n, all.ch, reinterpret_cast<const uint8_t*>(cent_px_dev), K,
                                 ctx->out_idx.as<uint32_t>(), nullptr, nullptr, nullptr, ctx->stream));
        return 0;
    }
    CK(ctx->cent_feat.ensure((size_t)K * 256));
    if (hdr) LAUNCH(launch_floats_to_features(reinterpret_cast<const float*>(cent_px_dev), K, all.ch, ctx->cent_feat.as<float>(), ctx->stream));
    else LAUNCH(launch_pixels_to_features(reinterpret_cast<const uint8_t*>(cent_px_dev), K, all.ch, nullptr, 0, ctx->cent_feat.as<float>(), ctx->stream));
    CK(ctx->out_idx.ensure(std::max<size_t>(1, all.n) * 4));
    return assign_blocks(ctx, all.feat, all.n, ctx->cent_feat.as<float>(), K, assign_mode, ctx->out_idx.as<uint32_t>(), nullptr,
                         nullptr);
}

int check_cfg(yupvq_ctx* ctx, const yupvq_config* cfg, size_t n, int ch) {
    if (!cfg) return ctx->fail("config is null");
    if (ch < 1 || ch > 4) return ctx->fail("channels must be 1..4 (got %d)", ch);
    if (n == 0) return ctx->fail("no blocks (the reference's LDR BuildCodebook has no N == 0 guard; rejected here)");
    if (n > 0xffffffffull) return ctx->fail("too many blocks");
    if (cfg->codebook_size == 0 || cfg->codebook_size > 65536) return ctx->fail("codebook_size out of range");
    return 0;
}

// Features built by our own decode / conversion kernels carry alpha == 1.0f exactly when the input has 1 or 3 channels
// (convert.cu block_features / float path), so the 16 alpha columns vanish after centring: 48-column tensor operands.
int rgb_layout(int channels) {
    return (!debug_opts().no_rgb48 && (channels == 1 || channels == 3)) ? 1 : 0;
}

struct Scope {  // selects the context's device for the duration of a call
    int prev = 0;
    explicit Scope(int dev) {
        cudaGetDevice(&prev);
        cudaSetDevice(dev);
    }
    ~Scope() { cudaSetDevice(prev); }
};

}  // namespace

// =====================================================================================================
// C ABI
// =====================================================================================================
extern "C" {

const char* yupvq_version(void) { return "yupvq 0.1 (sm_100a)"; }

uint32_t yupvq_codebook_size(float quality) {
    // CompressionConfig::SetQuality, src/vq_encoder.cpp:12-18; powers are the struct defaults 4..10
    // because Compress copies the params' powers only after calling SetQuality.
    const float q = std::min(std::max(quality, 0.0f), 1.0f);
    const uint32_t power = 4u + (uint32_t)roundf(q * (float)(10 - 4));
    return 1u << power;
}

void yupvq_config_init(yupvq_config* cfg, float quality, uint32_t max_iterations, int metric, float sample_ratio,
                       uint64_t seed) {
    std::memset(cfg, 0, sizeof(*cfg));
    cfg->codebook_size = yupvq_codebook_size(quality);
    cfg->max_iterations = max_iterations;
    cfg->metric = metric;
    cfg->sample_ratio = sample_ratio;
    cfg->seed = seed;
    cfg->assign_mode = YUPVQ_ASSIGN_AUTO;
}

int yupvq_create(int device, yupvq_ctx** out) {
    if (!out) return 1;
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) {
        g_create_error = std::string("no CUDA device: ") + cudaGetErrorString(e) + " (yupvq has no CPU fallback)";
        cudaGetLastError();
        return 1;
    }
    if (device < 0 || device >= count) {
        g_create_error = "device index out of range";
        return 1;
    }
    cudaDeviceProp prop;
    cudaGetDeviceProperties(&prop, device);
    if (prop.major != 10) {
        g_create_error = std::string("device ") + prop.name + " is not sm_100a (kernels are built for B200 only)";
        return 1;
    }
    Scope scope(device);
    (void)debug_opts();  // the only place the environment is read
    yupvq_ctx* ctx = new yupvq_ctx();
    ctx->device = device;
    bool ok = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) == cudaSuccess;
    for (auto& ev : ctx->ev) ok = ok && cudaEventCreate(&ev) == cudaSuccess;
    for (auto& ev : ctx->ev_user) ok = ok && cudaEventCreate(&ev) == cudaSuccess;
    ok = ok && cudaEventCreateWithFlags(&ctx->ev_poll, cudaEventDisableTiming) == cudaSuccess;
    ok = ok && cudaMallocHost(reinterpret_cast<void**>(&ctx->h_small), 65536) == cudaSuccess;
    ok = ok && ctx->small.ensure(F_WORDS * 4) == cudaSuccess;
    ok = ok && cudaMemset(ctx->small.p, 0, F_WORDS * 4) == cudaSuccess;
    ok = ok && colour_tables_init() == cudaSuccess;
    if (!ok) {
        g_create_error = std::string("context creation failed: ") + cudaGetErrorString(cudaGetLastError());
        yupvq_destroy(ctx);
        return 1;
    }
    *out = ctx;
    return 0;
}

void yupvq_destroy(yupvq_ctx* ctx) {
    if (!ctx) return;
    Scope scope(ctx->device);
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    xchg_release(ctx);
    if (ctx->comm) ctx->nccl.destroy(ctx->comm);
    DevBuf* bufs[] = {&ctx->in_bytes, &ctx->planes, &ctx->feat, &ctx->s_planes, &ctx->s_feat, &ctx->s_rows,
                      &ctx->min_dist, &ctx->partial, &ctx->seed_idx, &ctx->canon, &ctx->small, &ctx->assign,
                      &ctx->errors, &ctx->cent, &ctx->acc, &ctx->cent_px, &ctx->cent_feat, &ctx->out_idx,
                      &ctx->tc_a, &ctx->tc_xnorm, &ctx->tc_small, &ctx->tc_b, &ctx->tc_bx, &ctx->tc_cand, &ctx->tc_amb,
                      &ctx->bt_feat, &ctx->bt_tca, &ctx->bt_xnorm, &ctx->bt_tcsmall, &ctx->bt_assign, &ctx->bt_errors,
                      &ctx->bt_cent, &ctx->bt_acc, &ctx->bt_misc, &ctx->bt_centpx, &ctx->bt_cent0, &ctx->tc_chg, &ctx->delta_ctl, &ctx->bt_acc_local, &ctx->pal5, &ctx->tail_scratch, &ctx->bt_planes, &ctx->bt_in, &ctx->bt_rows, &ctx->img_in, &ctx->bc_enc};
    for (DevBuf* b : bufs) b->release();
    if (ctx->h_small) cudaFreeHost(ctx->h_small);
    for (auto& ev : ctx->ev)
        if (ev) cudaEventDestroy(ev);
    for (auto& ev : ctx->ev_user)
        if (ev) cudaEventDestroy(ev);
    if (ctx->ev_poll) cudaEventDestroy(ctx->ev_poll);
    for (auto& ev : ctx->pev)
        if (ev) cudaEventDestroy(ev);
    for (auto& ev : ctx->sev)
        if (ev) cudaEventDestroy(ev);
    if (ctx->stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

int yupvq_profile(yupvq_ctx* ctx, int enable) {
    if (!ctx) return 1;
    Scope scope(ctx->device);
    if (enable && ctx->pev.empty()) {
        ctx->pev.resize(256);
        for (auto& e : ctx->pev) CK(cudaEventCreate(&e));
        ctx->sev.resize(512);
        ctx->sev_stage.assign(512, 255);
        for (auto& e : ctx->sev) CK(cudaEventCreate(&e));
    }
    ctx->profile = enable != 0;
    ctx->pev_used = 0;
    ctx->sev_used = 0;
    return 0;
}

int yupvq_profile_read(yupvq_ctx* ctx, float* out_ms_front_end, uint32_t* out_launches) {
    if (!ctx) return 1;
    Scope scope(ctx->device);
    CK(cudaStreamSynchronize(ctx->stream));
    float total = 0.f;
    uint32_t n = 0;
    for (size_t i = 0; i + 1 < ctx->pev_used; i += 2) {
        float ms = 0.f;
        CK(cudaEventElapsedTime(&ms, ctx->pev[i], ctx->pev[i + 1]));
        total += ms;
        ++n;
    }
    ctx->pev_used = 0;
    if (out_ms_front_end) *out_ms_front_end = total;
    if (out_launches) *out_launches = n;
    return 0;
}

int yupvq_exchange_path(const yupvq_ctx* ctx) { return ctx ? ctx->xchg_path : YUPVQ_XCHG_NONE; }

int yupvq_profile_stages(yupvq_ctx* ctx, float out_ms[4], uint32_t* out_passes) {
    if (!ctx) return 1;
    Scope scope(ctx->device);
    CK(cudaStreamSynchronize(ctx->stream));
    float ms[4] = {0.f, 0.f, 0.f, 0.f};
    uint32_t passes = 0;
    for (size_t i = 0; i < ctx->sev_used; ++i) {
        const int st = ctx->sev_stage[i];
        if (st == 255) {
            ++passes;
            continue;
        }
        if (i == 0 || st > 3) continue;
        float v = 0.f;
        CK(cudaEventElapsedTime(&v, ctx->sev[i - 1], ctx->sev[i]));
        ms[st] += v;
    }
    ctx->sev_used = 0;
    for (int i = 0; i < 4; ++i)
        if (out_ms) out_ms[i] = ms[i];
    if (out_passes) *out_passes = passes;
    return 0;
}

int yupvq_timer_start(yupvq_ctx* ctx) {
    if (!ctx) return 1;
    Scope scope(ctx->device);
    CK(cudaEventRecord(ctx->ev_user[0], ctx->stream));
    return 0;
}

int yupvq_timer_stop(yupvq_ctx* ctx, float* out_ms) {
    if (!ctx) return 1;
    Scope scope(ctx->device);
    CK(cudaEventRecord(ctx->ev_user[1], ctx->stream));
    CK(cudaEventSynchronize(ctx->ev_user[1]));
    float ms = 0.f;
    CK(cudaEventElapsedTime(&ms, ctx->ev_user[0], ctx->ev_user[1]));
    if (out_ms) *out_ms = ms;
    return 0;
}

const char* yupvq_last_error(const yupvq_ctx* ctx) { return ctx ? ctx->err.c_str() : g_create_error.c_str(); }

// ---- whole path: LDR ---------------------------------------------------------------------------------
int yupvq_encode_bcn(yupvq_ctx* ctx, const uint8_t* bc_blocks, size_t n_blocks, int bc_format, int channels,
                     const yupvq_config* cfg, uint8_t* out_pixel_centroids, uint32_t* out_indices, yupvq_stats* stats) {
    if (!ctx) return 1;
    Scope scope(ctx->device);
    init_stats(stats);
    ctx->launches = 0;
    ctx->tc_feat = nullptr;
    ctx->maxabs_feat = nullptr;
    ctx->tc_rgb = rgb_layout(channels);
    if (check_cfg(ctx, cfg, n_blocks, channels)) return 1;
    if (!bc_supported(bc_format)) return ctx->fail("BC format %d is not supported on the LDR VQ path", bc_format);
    Timer tm(ctx);
    const int t0 = tm.mark();
    const void* bc_dev = nullptr;
    if (to_device(ctx, bc_blocks, n_blocks * bc_block_bytes(bc_format), ctx->in_bytes, &bc_dev)) return 1;
    const int t1 = tm.mark();
    CK(ctx->planes.ensure(n_blocks * 16 * channels));
    CK(ctx->feat.ensure(n_blocks * 256));
    // Without sampling the decode is split: pixel planes now (all the seeding reads), Oklab features after the seeding.
    const bool no_split = debug_opts().no_split_decode;  // A/B timing
    const bool sampling = cfg->sample_ratio < 1.0f && cfg->sample_ratio > 0.0f;
    const bool split = !sampling && !no_split;
    const bool sad = cfg->metric == YUPVQ_METRIC_SAD;
    LAUNCH(launch_decode_bcn(reinterpret_cast<const uint8_t*>(bc_dev), n_blocks, bc_format, channels, ctx->planes.as<uint4>(), n_blocks,
                             (split || sad) ? nullptr : ctx->feat.as<float>(), nullptr, ctx->stream));
    if (sad && !split)  // SAD path: the "features" are the pixel bytes themselves (sad.cu)
        LAUNCH(launch_sad_expand(ctx->planes.as<uint4>(), n_blocks, n_blocks, channels, ctx->feat.as<float>(), ctx->stream));
    float ms_late = 0.f;
    int tl0 = -1, tl1e = -1;
    std::function<int()> late;
    // Fused feature pass (tensor engine): the decode that writes the Oklab features also writes the front-end's A' operand,
    // |x'| and max |x| -- otherwise three more passes over the 256 MB feature matrix (col_sums, max_abs, prep_blocks).  The
    // centring row mu must exist before that pass, so it is the mean of the K seed blocks' features instead of the mean of
    // all blocks: any mu gives the same distances (translation invariance); it only conditions the bf16 split.
    const bool fuse_operands = split && !sad && !debug_opts().no_decode_fuse && use_tensor(cfg->assign_mode, n_blocks, cfg->codebook_size) &&
                               (reinterpret_cast<uintptr_t>(bc_dev) & 15u) == 0;
    if (split)
        late = [&]() -> int {
            tl0 = tm.mark();
            if (sad) {
                LAUNCH(launch_sad_expand(ctx->planes.as<uint4>(), n_blocks, n_blocks, channels, ctx->feat.as<float>(), ctx->stream));
            } else if (fuse_operands) {
                const uint32_t K = cfg->codebook_size;
                const size_t bs = bc_block_bytes(bc_format);
                CK(ctx->s_rows.ensure((size_t)K * bs));
                CK(ctx->cent.ensure((size_t)K * 256));
                CK(ctx->tc_small.ensure(TC_SMALL_BYTES));
                CK(ctx->tc_a.ensure(n_blocks * 256));
                CK(ctx->tc_xnorm.ensure(n_blocks * 4));
                uint8_t* sm = ctx->tc_small.as<uint8_t>();
                uint32_t* flags = ctx->small.as<uint32_t>();
                // the K seed blocks -> their feature rows -> mu
                LAUNCH(launch_gather_codebook(reinterpret_cast<const uint8_t*>(bc_dev), (uint32_t)n_blocks, (uint32_t)bs, ctx->seed_idx.as<uint32_t>(), K,
                                              ctx->s_rows.as<uint8_t>(), ctx->stream));
                LAUNCH(launch_decode_bcn(ctx->s_rows.as<uint8_t>(), K, bc_format, channels, nullptr, 0, ctx->cent.as<float>(), nullptr, ctx->stream));
                LAUNCH(launch_tc_mean(ctx->cent.as<float>(), K, reinterpret_cast<double*>(sm + TC_SUMS), reinterpret_cast<float*>(sm + TC_MU), ctx->stream));
                ++ctx->launches;
                CK(cudaMemsetAsync(flags + F_MAXABS, 0, 4, ctx->stream));
                DecodeFuse fu;
                fu.mu = reinterpret_cast<const float*>(sm + TC_MU);
                fu.a_prime = ctx->tc_a.as<float>();
                fu.xnorm = ctx->tc_xnorm.as<float>();
                fu.max_abs_bits = flags + F_MAXABS;
                fu.rgb = ctx->tc_rgb;
                LAUNCH(launch_decode_bcn(reinterpret_cast<const uint8_t*>(bc_dev), n_blocks, bc_format, channels, nullptr, 0,
                                         ctx->feat.as<float>(), nullptr, ctx->stream, &fu));
                ctx->tc_feat = ctx->feat.as<float>();     // tc_prepare_blocks and run_lloyd's max_abs find their work done
                ctx->tc_n = n_blocks;
                ctx->maxabs_feat = ctx->feat.as<float>();
            } else {
                LAUNCH(launch_decode_bcn(reinterpret_cast<const uint8_t*>(bc_dev), n_blocks, bc_format, channels, nullptr, 0,
                                         ctx->feat.as<float>(), nullptr, ctx->stream));
            }
            tl1e = tm.mark();
            return 0;
        };
    const int t2 = tm.mark();
    Blocks all;
    all.n = n_blocks;
    all.ch = channels;
    all.planes = ctx->planes.as<uint4>();
    all.stride = n_blocks;
    all.feat = ctx->feat.as<float>();
    if (!sampling && attach_palette_form(ctx, all, bc_dev, bc_format, channels)) return 1;
    int ts0, ts1, tl1, tf1;
    if (build_codebook_device(ctx, all, n_blocks, cfg, false, stats, tm, &ts0, &ts1, &tl1, &tf1, late)) return 1;
    if (quantize_device(ctx, all, ctx->cent_px.p, cfg->codebook_size, false, cfg->assign_mode, cfg->metric)) return 1;
    const int t3 = tm.mark();
    if (from_device(ctx, out_pixel_centroids, ctx->cent_px.p, (size_t)cfg->codebook_size * 16 * channels)) return 1;
    if (from_device(ctx, out_indices, ctx->out_idx.p, n_blocks * 4)) return 1;
    const int t4 = tm.mark();
    CK(cudaStreamSynchronize(ctx->stream));
    if (stats) {
        if (tl0 >= 0) ms_late = tm.ms(tl0, tl1e);   // the feature half of the split decode ran inside the seeding window
        stats->ms_h2d = tm.ms(t0, t1);
        stats->ms_decode = tm.ms(t1, t2) + ms_late;
        stats->ms_seed = tm.ms(ts0, ts1) - ms_late;
        stats->ms_lloyd = tm.ms(ts1, tl1);
        stats->ms_finalize = tm.ms(tl1, tf1);
        stats->ms_quantize = tm.ms(tf1, t3);
        stats->ms_d2h = tm.ms(t3, t4);
        stats->ms_total = tm.ms(t0, t4);
        stats->gpu_launches = ctx->launches;
    }
    return 0;
}

// ---- whole path: HDR ---------------------------------------------------------------------------------
int yupvq_encode_hdr(yupvq_ctx* ctx, const float* image, uint32_t width, uint32_t height, int channels,
                     const yupvq_config* cfg, float* out_float_centroids, uint32_t* out_indices, yupvq_stats* stats) {
    if (!ctx) return 1;
    Scope scope(ctx->device);
    init_stats(stats);
    ctx->launches = 0;
    ctx->tc_feat = nullptr;
    ctx->maxabs_feat = nullptr;
    ctx->tc_rgb = rgb_layout(channels);
    const size_t n_blocks = (size_t)((width + 3) / 4) * ((height + 3) / 4);
    if (check_cfg(ctx, cfg, n_blocks, channels)) return 1;
    Timer tm(ctx);
    const int t0 = tm.mark();
    const void* img_dev = nullptr;
    if (to_device(ctx, image, (size_t)width * height * channels * 4, ctx->in_bytes, &img_dev)) return 1;
    const int t1 = tm.mark();
    CK(ctx->feat.ensure(n_blocks * 256));
    LAUNCH(launch_image_to_features(reinterpret_cast<const float*>(img_dev), width, height, channels,
                                    ctx->feat.as<float>(), nullptr, ctx->stream));
    const int t2 = tm.mark();
    Blocks all;
    all.n = n_blocks;
    all.ch = channels;
    all.feat = ctx->feat.as<float>();
    int ts0, ts1, tl1, tf1;
    if (build_codebook_device(ctx, all, n_blocks, cfg, true, stats, tm, &ts0, &ts1, &tl1, &tf1)) return 1;
    if (quantize_device(ctx, all, ctx->cent_px.p, cfg->codebook_size, true, cfg->assign_mode)) return 1;
    const int t3 = tm.mark();
    if (from_device(ctx, out_float_centroids, ctx->cent_px.p, (size_t)cfg->codebook_size * 16 * channels * 4)) return 1;
    if (from_device(ctx, out_indices, ctx->out_idx.p, n_blocks * 4)) return 1;
    const int t4 = tm.mark();
    CK(cudaStreamSynchronize(ctx->stream));
    if (stats) {
        stats->ms_h2d = tm.ms(t0, t1);
        stats->ms_decode = tm.ms(t1, t2);
        stats->ms_seed = tm.ms(ts0, ts1);
        stats->ms_lloyd = tm.ms(ts1, tl1);
        stats->ms_finalize = tm.ms(tl1, tf1);
        stats->ms_quantize = tm.ms(tf1, t3);
        stats->ms_d2h = tm.ms(t3, t4);
        stats->ms_total = tm.ms(t0, t4);
        stats->gpu_launches = ctx->launches;
    }
    return 0;
}

int yupvq_encode_bcn_batch(yupvq_ctx* ctx, const uint8_t* const* bc_blocks, size_t n_textures, size_t n_blocks,
                           int bc_format, int channels, const yupvq_config* cfg, uint8_t* const* out_pixel_centroids,
                           uint32_t* const* out_indices, yupvq_stats* stats) {
    if (!ctx) return 1;
    if (!cfg) return ctx->fail("config is null");
    yupvq_stats total;
    std::memset(&total, 0, sizeof(total));
    for (size_t t = 0; t < n_textures; ++t) {
        yupvq_config c = *cfg;
        if (c.seed != 0) c.seed = cfg->seed + t;
        yupvq_stats s;
        if (yupvq_encode_bcn(ctx, bc_blocks[t], n_blocks, bc_format, channels, &c, out_pixel_centroids[t],
                             out_indices[t], &s))
            return 1;
        total.n_blocks += s.n_blocks;
        total.codebook_size = s.codebook_size;
        total.lloyd_passes += s.lloyd_passes;
        total.seed_rounds += s.seed_rounds;
        total.reseeds += s.reseeds;
        total.rerank_blocks += s.rerank_blocks;
        total.rerank_overflow += s.rerank_overflow;
        total.ms_h2d += s.ms_h2d; total.ms_decode += s.ms_decode; total.ms_seed += s.ms_seed;
        total.ms_lloyd += s.ms_lloyd; total.ms_finalize += s.ms_finalize; total.ms_quantize += s.ms_quantize;
        total.ms_d2h += s.ms_d2h; total.ms_total += s.ms_total;
        total.gpu_launches += s.gpu_launches;
    }
    if (stats) *stats = total;
    return 0;
}

// ---- the four VQEncoder calls -----------------------------------------------------------------------
int yupvq_build_codebook(yupvq_ctx* ctx, const uint8_t* pixel_blocks, size_t n_blocks, int channels,
                         const yupvq_config* cfg, uint8_t* out_pixel_centroids, yupvq_stats* stats) {
    if (!ctx) return 1;
    Scope scope(ctx->device);
    init_stats(stats);
    ctx->launches = 0;
    ctx->tc_feat = nullptr;
    ctx->maxabs_feat = nullptr;
    ctx->tc_rgb = rgb_layout(channels);
    if (check_cfg(ctx, cfg, n_blocks, channels)) return 1;
    Timer tm(ctx);
    const int t0 = tm.mark();
    const void* px_dev = nullptr;
    if (to_device(ctx, pixel_blocks, n_blocks * 16 * channels, ctx->in_bytes, &px_dev)) return 1;
    CK(ctx->planes.ensure(n_blocks * 16 * channels));
    CK(ctx->feat.ensure(n_blocks * 256));
    // as in yupvq_encode_bcn: pixel planes before the seeding, features after it (no sampling)
    const bool no_split = debug_opts().no_split_decode;
    const bool sampling = cfg->sample_ratio < 1.0f && cfg->sample_ratio > 0.0f;
    const bool split = !sampling && !no_split;
    const bool sad = cfg->metric == YUPVQ_METRIC_SAD;
    LAUNCH(launch_pixels_to_features(reinterpret_cast<const uint8_t*>(px_dev), n_blocks, channels, ctx->planes.as<uint4>(), n_blocks,
                                     (split || sad) ? nullptr : ctx->feat.as<float>(), ctx->stream));
    if (sad && !split)
        LAUNCH(launch_sad_expand(ctx->planes.as<uint4>(), n_blocks, n_blocks, channels, ctx->feat.as<float>(), ctx->stream));
    std::function<int()> late;
    if (split)
        late = [&]() -> int {
            if (sad) LAUNCH(launch_sad_expand(ctx->planes.as<uint4>(), n_blocks, n_blocks, channels, ctx->feat.as<float>(), ctx->stream));
            else LAUNCH(launch_pixels_to_features(reinterpret_cast<const uint8_t*>(px_dev), n_blocks, channels, nullptr, 0,
                                                  ctx->feat.as<float>(), ctx->stream));
            return 0;
        };
    Blocks all;
    all.n = n_blocks;
    all.ch = channels;
    all.planes = ctx->planes.as<uint4>();
    all.stride = n_blocks;
    all.feat = ctx->feat.as<float>();
    int ts0, ts1, tl1, tf1;
    if (build_codebook_device(ctx, all, n_blocks, cfg, false, stats, tm, &ts0, &ts1, &tl1, &tf1, late)) return 1;
    if (from_device(ctx, out_pixel_centroids, ctx->cent_px.p, (size_t)cfg->codebook_size * 16 * channels)) return 1;
    const int t4 = tm.mark();
    CK(cudaStreamSynchronize(ctx->stream));
    if (stats) {
        stats->ms_decode = tm.ms(t0, ts0);
        stats->ms_seed = tm.ms(ts0, ts1);
        stats->ms_lloyd = tm.ms(ts1, tl1);
        stats->ms_finalize = tm.ms(tl1, tf1);
        stats->ms_total = tm.ms(t0, t4);
        stats->gpu_launches = ctx->launches;
    }
    return 0;
}

int yupvq_quantize_blocks(yupvq_ctx* ctx, const uint8_t* pixel_blocks, size_t n_blocks, int channels,
                          const uint8_t* pixel_centroids, uint32_t codebook_size, int metric, int assign_mode,
                          uint32_t* out_indices, yupvq_stats* stats) {
    if (!ctx) return 1;
    Scope scope(ctx->device);
    init_stats(stats);
    ctx->launches = 0;
    ctx->tc_feat = nullptr;
    ctx->maxabs_feat = nullptr;
    ctx->tc_rgb = rgb_layout(channels);
    if (n_blocks == 0) return 0;  // reference returns {} (:449)
    if (channels < 1 || channels > 4) return ctx->fail("channels must be 1..4");
    if (metric != YUPVQ_METRIC_OKLAB && metric != YUPVQ_METRIC_SAD) return ctx->fail("unknown metric %d", metric);
    Timer tm(ctx);
    const int t0 = tm.mark();
    const void* px_dev = nullptr;
    if (to_device(ctx, pixel_blocks, n_blocks * 16 * channels, ctx->in_bytes, &px_dev)) return 1;
    CK(ctx->feat.ensure(n_blocks * 256));
    const bool sad = metric == YUPVQ_METRIC_SAD;
    if (sad) CK(ctx->planes.ensure(n_blocks * 16 * channels));
    LAUNCH(launch_pixels_to_features(reinterpret_cast<const uint8_t*>(px_dev), n_blocks, channels,
                                     sad ? ctx->planes.as<uint4>() : nullptr, sad ? n_blocks : 0, ctx->feat.as<float>(), ctx->stream));
    const void* c_dev = nullptr;
    if (to_device(ctx, pixel_centroids, (size_t)codebook_size * 16 * channels, ctx->cent_px, &c_dev)) return 1;
    if (sad && ((uintptr_t)c_dev & 15)) {  // the SAD kernel reads centroids as uint4
        CK(ctx->cent_px.ensure((size_t)codebook_size * 16 * channels));
        CK(cudaMemcpyAsync(ctx->cent_px.p, c_dev, (size_t)codebook_size * 16 * channels, cudaMemcpyDeviceToDevice, ctx->stream));
        c_dev = ctx->cent_px.p;
    }
    Blocks all;
    all.n = n_blocks;
    all.ch = channels;
    all.planes = sad ? ctx->planes.as<uint4>() : nullptr;
    all.stride = n_blocks;
    all.feat = ctx->feat.as<float>();
    if (quantize_device(ctx, all, c_dev, codebook_size, false, assign_mode, metric)) return 1;
    if (from_device(ctx, out_indices, ctx->out_idx.p, n_blocks * 4)) return 1;
    const int t1 = tm.mark();
    CK(cudaStreamSynchronize(ctx->stream));
    if (stats) {
        stats->n_blocks = (uint32_t)n_blocks;
        stats->codebook_size = codebook_size;
        stats->ms_quantize = stats->ms_total = tm.ms(t0, t1);
        stats->gpu_launches = ctx->launches;
    }
    return 0;
}

int yupvq_build_codebook_hdr(yupvq_ctx* ctx, const float* float_blocks, size_t n_blocks, int channels,
                             const yupvq_config* cfg, float* out_float_centroids, yupvq_stats* stats) {
    if (!ctx) return 1;
    Scope scope(ctx->device);
    init_stats(stats);
    ctx->launches = 0;
    ctx->tc_feat = nullptr;
    ctx->maxabs_feat = nullptr;
    ctx->tc_rgb = rgb_layout(channels);
    if (n_blocks == 0) return 0;  // reference returns {} (:517)
    if (check_cfg(ctx, cfg, n_blocks, channels)) return 1;
    Timer tm(ctx);
    const int t0 = tm.mark();
    const void* px_dev = nullptr;
    if (to_device(ctx, float_blocks, n_blocks * 16 * channels * 4, ctx->in_bytes, &px_dev)) return 1;
    CK(ctx->feat.ensure(n_blocks * 256));
    LAUNCH(launch_floats_to_features(reinterpret_cast<const float*>(px_dev), n_blocks, channels, ctx->feat.as<float>(),
                                     ctx->stream));
    Blocks all;
    all.n = n_blocks;
    all.ch = channels;
    all.feat = ctx->feat.as<float>();
    int ts0, ts1, tl1, tf1;
    if (build_codebook_device(ctx, all, n_blocks, cfg, true, stats, tm, &ts0, &ts1, &tl1, &tf1)) return 1;
    if (from_device(ctx, out_float_centroids, ctx->cent_px.p, (size_t)cfg->codebook_size * 16 * channels * 4)) return 1;
    const int t4 = tm.mark();
    CK(cudaStreamSynchronize(ctx->stream));
    if (stats) {
        stats->ms_decode = tm.ms(t0, ts0);
        stats->ms_seed = tm.ms(ts0, ts1);
        stats->ms_lloyd = tm.ms(ts1, tl1);
        stats->ms_finalize = tm.ms(tl1, tf1);
        stats->ms_total = tm.ms(t0, t4);
        stats->gpu_launches = ctx->launches;
    }
    return 0;
}

int yupvq_quantize_blocks_hdr(yupvq_ctx* ctx, const float* float_blocks, size_t n_blocks, int channels,
                              const float* float_centroids, uint32_t codebook_size, int assign_mode,
                              uint32_t* out_indices, yupvq_stats* stats) {
    if (!ctx) return 1;
    Scope scope(ctx->device);
    init_stats(stats);
    ctx->launches = 0;
    ctx->tc_feat = nullptr;
    ctx->maxabs_feat = nullptr;
    ctx->tc_rgb = rgb_layout(channels);
    if (n_blocks == 0) return 0;  // :631
    if (channels < 1 || channels > 4) return ctx->fail("channels must be 1..4");
    Timer tm(ctx);
    const int t0 = tm.mark();
    const void* px_dev = nullptr;
    if (to_device(ctx, float_blocks, n_blocks * 16 * channels * 4, ctx->in_bytes, &px_dev)) return 1;
    CK(ctx->feat.ensure(n_blocks * 256));
    LAUNCH(launch_floats_to_features(reinterpret_cast<const float*>(px_dev), n_blocks, channels, ctx->feat.as<float>(),
                                     ctx->stream));
    const void* c_dev = nullptr;
    if (to_device(ctx, float_centroids, (size_t)codebook_size * 16 * channels * 4, ctx->cent_px, &c_dev)) return 1;
    Blocks all;
    all.n = n_blocks;
    all.ch = channels;
    all.feat = ctx->feat.as<float>();
    if (quantize_device(ctx, all, c_dev, codebook_size, true, assign_mode)) return 1;
    if (from_device(ctx, out_indices, ctx->out_idx.p, n_blocks * 4)) return 1;
    const int t1 = tm.mark();
    CK(cudaStreamSynchronize(ctx->stream));
    if (stats) {
        stats->n_blocks = (uint32_t)n_blocks;
        stats->codebook_size = codebook_size;
        stats->ms_quantize = stats->ms_total = tm.ms(t0, t1);
        stats->gpu_launches = ctx->launches;
    }
    return 0;
}

// ---- kernel-level entry points ------------------------------------------------------------------------
int yupvq_decode_bcn(yupvq_ctx* ctx, const uint8_t* bc_blocks, size_t n_blocks, int bc_format, int channels,
                     uint8_t* out_pixel_blocks) {
    if (!ctx) return 1;
    Scope scope(ctx->device);
    if (!bc_supported(bc_format)) return ctx->fail("BC format %d is not supported", bc_format);
    if (channels < 1 || channels > 4) return ctx->fail("channels must be 1..4");
    if (n_blocks == 0) return 0;
    const void* bc_dev = nullptr;
    if (to_device(ctx, bc_blocks, n_blocks * bc_block_bytes(bc_format), ctx->in_bytes, &bc_dev)) return 1;
    CK(ctx->planes.ensure(n_blocks * 16 * channels));
    LAUNCH(launch_decode_bcn(reinterpret_cast<const uint8_t*>(bc_dev), n_blocks, bc_format, channels, nullptr, 0,
                             nullptr, ctx->planes.as<uint8_t>(), ctx->stream));
    if (from_device(ctx, out_pixel_blocks, ctx->planes.p, n_blocks * 16 * channels)) return 1;
    CK(cudaStreamSynchronize(ctx->stream));
    return 0;
}

int yupvq_pixels_to_oklab(yupvq_ctx* ctx, const uint8_t* pixel_blocks, size_t n_blocks, int channels, float* out_oklab) {
    if (!ctx) return 1;
    Scope scope(ctx->device);
    if (channels < 1 || channels > 4) return ctx->fail("channels must be 1..4");
    if (n_blocks == 0) return 0;
    const void* px_dev = nullptr;
    if (to_device(ctx, pixel_blocks, n_blocks * 16 * channels, ctx->in_bytes, &px_dev)) return 1;
    CK(ctx->feat.ensure(n_blocks * 256));
    LAUNCH(launch_pixels_to_features(reinterpret_cast<const uint8_t*>(px_dev), n_blocks, channels, nullptr, 0,
                                     ctx->feat.as<float>(), ctx->stream));
    if (from_device(ctx, out_oklab, ctx->feat.p, n_blocks * 256)) return 1;
    CK(cudaStreamSynchronize(ctx->stream));
    return 0;
}

int yupvq_oklab_to_pixels(yupvq_ctx* ctx, const float* oklab, size_t n_blocks, int channels, uint8_t* out_pixel_blocks) {
    if (!ctx) return 1;
    Scope scope(ctx->device);
    if (channels < 1 || channels > 4) return ctx->fail("channels must be 1..4");
    if (n_blocks == 0) return 0;
    const void* ok_dev = nullptr;
    if (to_device(ctx, oklab, n_blocks * 256, ctx->feat, &ok_dev)) return 1;
    CK(ctx->cent_px.ensure(n_blocks * 16 * channels));
    LAUNCH(launch_oklab_to_pixels(reinterpret_cast<const float*>(ok_dev), n_blocks, channels, ctx->cent_px.as<uint8_t>(), ctx->stream));
    if (from_device(ctx, out_pixel_blocks, ctx->cent_px.p, n_blocks * 16 * channels)) return 1;
    CK(cudaStreamSynchronize(ctx->stream));
    return 0;
}

int yupvq_floats_to_oklab(yupvq_ctx* ctx, const float* float_blocks, size_t n_blocks, int channels, float* out_oklab) {
    if (!ctx) return 1;
    Scope scope(ctx->device);
    if (channels < 1 || channels > 4) return ctx->fail("channels must be 1..4");
    if (n_blocks == 0) return 0;
    const void* px_dev = nullptr;
    if (to_device(ctx, float_blocks, n_blocks * 16 * channels * 4, ctx->in_bytes, &px_dev)) return 1;
    CK(ctx->feat.ensure(n_blocks * 256));
    LAUNCH(launch_floats_to_features(reinterpret_cast<const float*>(px_dev), n_blocks, channels, ctx->feat.as<float>(), ctx->stream));
    if (from_device(ctx, out_oklab, ctx->feat.p, n_blocks * 256)) return 1;
    CK(cudaStreamSynchronize(ctx->stream));
    return 0;
}

int yupvq_oklab_to_floats(yupvq_ctx* ctx, const float* oklab, size_t n_blocks, int channels, float* out_float_blocks) {
    if (!ctx) return 1;
    Scope scope(ctx->device);
    if (channels < 1 || channels > 4) return ctx->fail("channels must be 1..4");
    if (n_blocks == 0) return 0;
    const void* ok_dev = nullptr;
    if (to_device(ctx, oklab, n_blocks * 256, ctx->feat, &ok_dev)) return 1;
    CK(ctx->cent_px.ensure(n_blocks * 16 * channels * 4));
    LAUNCH(launch_oklab_to_floats(reinterpret_cast<const float*>(ok_dev), n_blocks, channels, ctx->cent_px.as<float>(), ctx->stream));
    if (from_device(ctx, out_float_blocks, ctx->cent_px.p, n_blocks * 16 * channels * 4)) return 1;
    CK(cudaStreamSynchronize(ctx->stream));
    return 0;
}

int yupvq_gather_tiles(yupvq_ctx* ctx, const float* image, uint32_t width, uint32_t height, int channels,
                       float* out_float_blocks) {
    if (!ctx) return 1;
    Scope scope(ctx->device);
    if (channels < 1 || channels > 4) return ctx->fail("channels must be 1..4");
    const size_t n = (size_t)((width + 3) / 4) * ((height + 3) / 4);
    if (n == 0) return 0;
    const void* img_dev = nullptr;
    if (to_device(ctx, image, (size_t)width * height * channels * 4, ctx->in_bytes, &img_dev)) return 1;
    CK(ctx->cent_px.ensure(n * 16 * channels * 4));
    LAUNCH(launch_image_to_features(reinterpret_cast<const float*>(img_dev), width, height, channels, nullptr,
                                    ctx->cent_px.as<float>(), ctx->stream));
    if (from_device(ctx, out_float_blocks, ctx->cent_px.p, n * 16 * channels * 4)) return 1;
    CK(cudaStreamSynchronize(ctx->stream));
    return 0;
}

int yupvq_assign(yupvq_ctx* ctx, const float* features, size_t n_blocks, const float* centroids, uint32_t codebook_size,
                 int assign_mode, uint32_t* out_indices, float* out_errors, yupvq_stats* stats) {
    if (!ctx) return 1;
    Scope scope(ctx->device);
    init_stats(stats);
    ctx->launches = 0;
    ctx->tc_feat = nullptr;
    ctx->maxabs_feat = nullptr;
    ctx->tc_rgb = 0;
    if (n_blocks == 0) return 0;
    const void* f_dev = nullptr;
    const void* c_dev = nullptr;
    if (to_device(ctx, features, n_blocks * 256, ctx->feat, &f_dev)) return 1;
    if (to_device(ctx, centroids, (size_t)codebook_size * 256, ctx->cent, &c_dev)) return 1;
    CK(ctx->out_idx.ensure(n_blocks * 4));
    CK(ctx->errors.ensure(n_blocks * 4));
    CK(cudaMemsetAsync(ctx->small.as<uint32_t>() + F_PAIRS, 0, 8, ctx->stream));
    Timer tm(ctx);
    const int t0 = tm.mark();
    if (assign_blocks(ctx, reinterpret_cast<const float*>(f_dev), n_blocks, reinterpret_cast<const float*>(c_dev), codebook_size,
                      assign_mode, ctx->out_idx.as<uint32_t>(), out_errors ? ctx->errors.as<float>() : nullptr, nullptr))
        return 1;
    const int t1 = tm.mark();
    if (from_device(ctx, out_indices, ctx->out_idx.p, n_blocks * 4)) return 1;
    if (from_device(ctx, out_errors, ctx->errors.p, n_blocks * 4)) return 1;
    CK(cudaMemcpyAsync(ctx->h_small, ctx->small.as<uint32_t>() + F_PAIRS, 8, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    if (stats) {
        stats->n_blocks = (uint32_t)n_blocks;
        stats->codebook_size = codebook_size;
        stats->rerank_blocks = ctx->h_small[0];
        stats->rerank_overflow = ctx->h_small[1];
        stats->ms_quantize = stats->ms_total = tm.ms(t0, t1);
        stats->gpu_launches = ctx->launches;
    }
    return 0;
}

int yupvq_seed_ldr(yupvq_ctx* ctx, const uint8_t* pixel_blocks, size_t n_blocks, int channels, uint32_t codebook_size,
                   uint64_t seed, uint32_t* out_seed_blocks, uint32_t* out_rounds) {
    if (!ctx) return 1;
    Scope scope(ctx->device);
    ctx->launches = 0;
    ctx->tc_feat = nullptr;
    ctx->maxabs_feat = nullptr;
    ctx->tc_rgb = 0;
    if (channels < 1 || channels > 4) return ctx->fail("channels must be 1..4");
    if (n_blocks == 0 || codebook_size == 0) return ctx->fail("empty input");
    const void* px_dev = nullptr;
    if (to_device(ctx, pixel_blocks, n_blocks * 16 * channels, ctx->in_bytes, &px_dev)) return 1;
    CK(ctx->planes.ensure(n_blocks * 16 * channels));
    LAUNCH(launch_pixels_to_features(reinterpret_cast<const uint8_t*>(px_dev), n_blocks, channels,
                                     ctx->planes.as<uint4>(), n_blocks, nullptr, ctx->stream));
    size_t n_used;
    Draws dr = make_draws(n_blocks, codebook_size, 1.0f, seed, &n_used);
    Blocks b;
    b.n = n_blocks;
    b.ch = channels;
    b.planes = ctx->planes.as<uint4>();
    b.stride = n_blocks;
    uint32_t rounds = 0;
    if (run_seeding(ctx, b, codebook_size, dr, false, &rounds)) return 1;
    if (out_rounds) *out_rounds = rounds;
    if (from_device(ctx, out_seed_blocks, ctx->seed_idx.p, (size_t)codebook_size * 4)) return 1;
    CK(cudaStreamSynchronize(ctx->stream));
    return 0;
}

int yupvq_seed_hdr(yupvq_ctx* ctx, const float* features, size_t n_blocks, uint32_t codebook_size, uint64_t seed,
                   uint32_t* out_seed_blocks, uint32_t* out_rounds) {
    if (!ctx) return 1;
    Scope scope(ctx->device);
    ctx->launches = 0;
    ctx->tc_feat = nullptr;
    ctx->maxabs_feat = nullptr;
    ctx->tc_rgb = 0;
    if (n_blocks == 0 || codebook_size == 0) return ctx->fail("empty input");
    const void* f_dev = nullptr;
    if (to_device(ctx, features, n_blocks * 256, ctx->feat, &f_dev)) return 1;
    size_t n_used;
    Draws dr = make_draws(n_blocks, codebook_size, 1.0f, seed, &n_used);
    Blocks b;
    b.n = n_blocks;
    b.ch = 4;
    b.feat = reinterpret_cast<const float*>(f_dev);
    uint32_t rounds = 0;
    if (run_seeding(ctx, b, codebook_size, dr, true, &rounds)) return 1;
    if (out_rounds) *out_rounds = rounds;
    if (from_device(ctx, out_seed_blocks, ctx->seed_idx.p, (size_t)codebook_size * 4)) return 1;
    CK(cudaStreamSynchronize(ctx->stream));
    return 0;
}

int yupvq_update_centroids(yupvq_ctx* ctx, const float* features, size_t n_blocks, const uint32_t* assignments,
                           const float* errors, uint32_t codebook_size, float* inout_centroids, uint32_t* out_counts) {
    if (!ctx) return 1;
    Scope scope(ctx->device);
    ctx->launches = 0;
    ctx->tc_feat = nullptr;
    ctx->maxabs_feat = nullptr;
    ctx->tc_rgb = 0;
    const uint32_t K = codebook_size;
    if (n_blocks == 0 || K == 0) return ctx->fail("empty input");
    const void* f_dev = nullptr;
    const void* a_dev = nullptr;
    const void* e_dev = nullptr;
    if (to_device(ctx, features, n_blocks * 256, ctx->feat, &f_dev)) return 1;
    if (to_device(ctx, assignments, n_blocks * 4, ctx->assign, &a_dev)) return 1;
    // errors are modified by the reseed: always work on the context's copy
    CK(ctx->errors.ensure(n_blocks * 4));
    CK(cudaMemcpyAsync(ctx->errors.p, errors, n_blocks * 4, cudaMemcpyDefault, ctx->stream));
    e_dev = ctx->errors.p;
    CK(ctx->cent.ensure((size_t)K * 256));
    CK(cudaMemcpyAsync(ctx->cent.p, inout_centroids, (size_t)K * 256, cudaMemcpyDefault, ctx->stream));
    const size_t acc_words = (size_t)K * 65 + 1;
    CK(ctx->acc.ensure(acc_words * 8));
    uint32_t* flags = ctx->small.as<uint32_t>();
    CK(cudaMemsetAsync(flags, 0, F_WORDS * 4, ctx->stream));
    LAUNCH(launch_max_abs(reinterpret_cast<const float*>(f_dev), n_blocks * 64, flags + F_MAXABS, ctx->stream));
    CK(cudaMemcpyAsync(ctx->h_small, flags + F_MAXABS, 4, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    float max_abs;
    std::memcpy(&max_abs, ctx->h_small, 4);
    const int shift = fixed_shift(n_blocks, max_abs);
    CK(cudaMemsetAsync(ctx->acc.p, 0, acc_words * 8, ctx->stream));
    LAUNCH(launch_accumulate(reinterpret_cast<const float*>(f_dev), n_blocks, reinterpret_cast<const uint32_t*>(a_dev), K,
                             shift, ctx->acc.as<long long>(), flags, nullptr, ctx->stream));
    LAUNCH(launch_finalize_centroids(ctx->acc.as<long long>(), K, shift, ctx->cent.as<float>(), flags, nullptr, ctx->stream));
    LAUNCH(launch_reseed_empty(reinterpret_cast<const float*>(f_dev), n_blocks, reinterpret_cast<float*>(const_cast<void*>(e_dev)),
                               ctx->acc.as<long long>(), K, ctx->cent.as<float>(), flags, nullptr, ctx->stream));
    if (from_device(ctx, inout_centroids, ctx->cent.p, (size_t)K * 256)) return 1;
    if (out_counts) {
        // counts live as 64-bit words behind the sums
        std::vector<long long> h(K);
        CK(cudaMemcpyAsync(h.data(), ctx->acc.as<long long>() + (size_t)K * 64, (size_t)K * 8, cudaMemcpyDeviceToHost, ctx->stream));
        CK(cudaStreamSynchronize(ctx->stream));
        if (is_device_ptr(out_counts)) {
            std::vector<uint32_t> h32(K);
            for (uint32_t c = 0; c < K; ++c) h32[c] = (uint32_t)h[c];
            CK(cudaMemcpy(out_counts, h32.data(), (size_t)K * 4, cudaMemcpyHostToDevice));
        } else {
            for (uint32_t c = 0; c < K; ++c) out_counts[c] = (uint32_t)h[c];
        }
    }
    CK(cudaStreamSynchronize(ctx->stream));
    return 0;
}

int yupvq_gather_codebook(yupvq_ctx* ctx, const uint8_t* codebook, uint32_t n_entries, uint32_t block_size,
                          const uint32_t* indices, size_t n_blocks, uint8_t* out_bc_blocks) {
    if (!ctx) return 1;
    Scope scope(ctx->device);
    ctx->launches = 0;
    ctx->tc_feat = nullptr;
    ctx->maxabs_feat = nullptr;
    ctx->tc_rgb = 0;
    if (block_size != 8 && block_size != 16) return ctx->fail("block_size must be 8 or 16");
    if (n_blocks == 0) return 0;
    const void* cb_dev = nullptr;
    const void* idx_dev = nullptr;
    if (to_device(ctx, codebook, (size_t)n_entries * block_size, ctx->cent_px, &cb_dev)) return 1;
    if (to_device(ctx, indices, n_blocks * 4, ctx->assign, &idx_dev)) return 1;
    uint8_t* out_dev = out_bc_blocks;
    const bool out_is_dev = is_device_ptr(out_bc_blocks);
    if (!out_is_dev) {
        CK(ctx->in_bytes.ensure(n_blocks * block_size));
        out_dev = ctx->in_bytes.as<uint8_t>();
    }
    LAUNCH(launch_gather_codebook(reinterpret_cast<const uint8_t*>(cb_dev), n_entries, block_size,
                                  reinterpret_cast<const uint32_t*>(idx_dev), n_blocks, out_dev, ctx->stream));
    if (!out_is_dev && from_device(ctx, out_bc_blocks, out_dev, n_blocks * block_size)) return 1;
    CK(cudaStreamSynchronize(ctx->stream));
    return 0;
}

// ---- N4: whole-texture BCn encode on the GPU (bcn_encode.cu) ---------------------------------------------------
int yupvq_bcn_encode_supported(int bc_format) { return bcn_encode_supported(bc_format) ? 1 : 0; }

namespace {
// image (host or device) -> BCn blocks in `out_dev` (device); CUDA-event time of the kernel alone in *out_ms
int bcn_encode_to_device(yupvq_ctx* ctx, const uint8_t* image, uint32_t width, uint32_t height, int channels, int bc_format,
                         int swap_rb, int alpha_threshold, uint8_t* out_dev, float* out_ms) {
    if (!image) return ctx->fail("image is null");
    if (width == 0 || height == 0) return ctx->fail("empty image");
    if (channels < 1 || channels > 4) return ctx->fail("channels must be 1..4 (got %d)", channels);
    if (!bcn_encode_supported(bc_format)) return ctx->fail("BC format %d has no GPU encoder (BC1, BC3, BC4, BC5 do)", bc_format);
    const void* img_dev = nullptr;
    if (to_device(ctx, image, (size_t)width * height * channels, ctx->img_in, &img_dev)) return 1;
    Timer tm(ctx);
    const int t0 = tm.mark();
    LAUNCH(launch_bcn_encode(reinterpret_cast<const uint8_t*>(img_dev), width, height, (uint32_t)channels, bc_format, swap_rb ? 1 : 0,
                             (uint32_t)std::min(std::max(alpha_threshold, 0), 256), out_dev, ctx->stream));
    const int t1 = tm.mark();
    if (out_ms) {
        CK(cudaStreamSynchronize(ctx->stream));
        *out_ms = tm.ms(t0, t1);
    }
    return 0;
}
}  // namespace

int yupvq_bcn_encode_image(yupvq_ctx* ctx, const uint8_t* image, uint32_t width, uint32_t height, int channels, int bc_format,
                           int swap_rb, int alpha_threshold, uint8_t* out_bc_blocks, float* out_ms_kernel) {
    if (!ctx) return 1;
    Scope scope(ctx->device);
    ctx->launches = 0;
    if (!out_bc_blocks) return ctx->fail("out_bc_blocks is null");
    const size_t bytes = (size_t)((width + 3) / 4) * ((height + 3) / 4) * bc_block_bytes(bc_format);
    uint8_t* out_dev = out_bc_blocks;
    // the encoder stores whole blocks as uint2 / uint4: it writes in place only into device memory aligned to the block size;
    // host buffers and misaligned device pointers (a slice of a larger buffer) are staged and copied
    const bool in_place = is_device_ptr(out_bc_blocks) && reinterpret_cast<uintptr_t>(out_bc_blocks) % bc_block_bytes(bc_format) == 0;
    if (!in_place) {
        CK(ctx->bc_enc.ensure(std::max<size_t>(bytes, 16)));
        out_dev = ctx->bc_enc.as<uint8_t>();
    }
    if (bcn_encode_to_device(ctx, image, width, height, channels, bc_format, swap_rb, alpha_threshold, out_dev, out_ms_kernel)) return 1;
    if (!in_place && from_device(ctx, out_bc_blocks, out_dev, bytes)) return 1;
    CK(cudaStreamSynchronize(ctx->stream));
    return 0;
}

int yupvq_compress_image(yupvq_ctx* ctx, const uint8_t* image, uint32_t width, uint32_t height, int channels, int bc_format,
                         int swap_rb, int alpha_threshold, const yupvq_config* cfg, uint8_t* out_bc_blocks,
                         uint8_t* out_pixel_centroids, uint32_t* out_indices, yupvq_stats* stats, float* out_ms_bcn_encode) {
    if (!ctx) return 1;
    {
        Scope scope(ctx->device);
        const size_t bytes = (size_t)((width + 3) / 4) * ((height + 3) / 4) * bc_block_bytes(bc_format);
        CK(ctx->bc_enc.ensure(std::max<size_t>(bytes, 16)));
        ctx->launches = 0;
        if (bcn_encode_to_device(ctx, image, width, height, channels, bc_format, swap_rb, alpha_threshold, ctx->bc_enc.as<uint8_t>(),
                                 out_ms_bcn_encode))
            return 1;
        if (out_bc_blocks && from_device(ctx, out_bc_blocks, ctx->bc_enc.p, bytes)) return 1;
    }
    // the blocks stay in HBM: the VQ stage reads them where the encoder wrote them
    const size_t n_blocks = (size_t)((width + 3) / 4) * ((height + 3) / 4);
    const int rc = yupvq_encode_bcn(ctx, ctx->bc_enc.as<uint8_t>(), n_blocks, bc_format, channels, cfg, out_pixel_centroids, out_indices, stats);
    if (rc == 0 && stats) stats->gpu_launches += 1;
    return rc;
}

// ---- multi-GPU -------------------------------------------------------------------------------------------
int yupvq_comm_unique_id(yupvq_ctx* ctx, uint8_t out_id[128]) {
    if (!ctx) return 1;
    if (ctx->nccl.load()) return ctx->fail("cannot load NCCL: %s", ctx->nccl.error.c_str());
    if (ctx->nccl.unique_id(out_id)) return ctx->fail("ncclGetUniqueId failed");
    return 0;
}

int yupvq_comm_init(yupvq_ctx* ctx, const uint8_t id[128], int n_ranks, int rank) {
    if (!ctx) return 1;
    Scope scope(ctx->device);
    if (ctx->nccl.load()) return ctx->fail("cannot load NCCL: %s", ctx->nccl.error.c_str());
    if (ctx->comm) {
        CK(cudaStreamSynchronize(ctx->stream));
        xchg_release(ctx);
        ctx->xchg_state = 0;
        ctx->nccl.destroy(ctx->comm);
        ctx->comm = nullptr;
    }
    if (ctx->nccl.init_rank(&ctx->comm, n_ranks, id, rank)) return ctx->fail("ncclCommInitRank failed");
    ctx->n_ranks = n_ranks;
    ctx->rank = rank;
    return 0;
}

// One texture over the ranks of the communicator: the lock-step batch below with a single texture.  Rank 0 seeds
// (K-Means++ is K-1 dependent rounds: replicating it on every rank would not make it shorter), the seed rows go out with
// one all-reduce, every rank runs the device-side Lloyd loop on its shard (no host round trip per pass) and ONE all-reduce
// per pass carries the fixed-point sums and counts.  With seed == 0 only rank 0 draws from std::random_device.
int yupvq_encode_bcn_sharded(yupvq_ctx* ctx, const uint8_t* bc_blocks, size_t n_total, size_t first_block,
                             size_t n_local, int bc_format, int channels, const yupvq_config* cfg,
                             uint8_t* out_pixel_centroids, uint32_t* out_indices, yupvq_stats* stats) {
    if (!ctx) return 1;
    const uint8_t* tex[1] = {bc_blocks};
    uint8_t* cents[1] = {out_pixel_centroids};
    uint32_t* idx[1] = {out_indices};
    return yupvq_encode_bcn_batch_sharded(ctx, tex, 1, n_total, first_block, n_local, bc_format, channels, cfg, cents, idx, stats, nullptr);
}


}  // extern "C" (reopened below)

namespace {

// Input of a lock-step batch: T BCn block arrays (LDR) or T float images (HDR), host or device pointers.
struct BatchSrc {
    bool hdr = false;
    const void* const* data = nullptr;
    int bc_format = 0;
    uint32_t width = 0, height = 0;
};

// The lock-step batch for the variants of the path: fast-mode sampling (src/vq_encoder.cpp:262-284), the SAD metric
// (:380-430) and HDR input (:493-622).  Same protocol as the Oklab LDR batch below -- texture t is seeded on rank
// t % R, the K seed rows of all textures go out with one all-reduce, all textures advance one Lloyd pass at a time and
// ONE all-reduce per pass carries the [K x 64 sums | K counts | changed] words of the whole batch; the convergence test
// runs on the device, a rank-sharded empty-cluster reseed falls back to the host-driven loop -- but written with the
// per-texture kernels (these variants are not the headline configuration):
//   * every rank needs every texture WHOLE (a sampled Lloyd row can be any block; an HDR tile any part of the image);
//   * with sampling the ranks split the SAMPLE list (in sample order, so "first maximum" keeps its meaning for the
//     reseed), not the texture; the list is drawn on the seeding rank and broadcast; the final index map is still this
//     rank's block range [first_block, first_block + n_local);
//   * SAD: integer sums with shift 0 (exact), sad_assign / sad_finalize instead of the Oklab kernels;
//   * HDR: the reference's `iter > 1` break rule, float centroids out.
// Results equal the single-GPU yupvq_encode_bcn / yupvq_encode_hdr with seed + t bit for bit (integer or fixed-point
// sums are order independent), HDR seeding as stated in DESIGN.md 4.3.
int batch_sharded_general(yupvq_ctx* ctx, const BatchSrc& src, size_t T, size_t n_total, size_t first_block, size_t n_local, int channels,
                          const yupvq_config* cfg, void* const* out_centroids, uint32_t* const* out_indices, yupvq_stats* stats) {
    const bool hdr = src.hdr;
    const bool sad = !hdr && cfg->metric == YUPVQ_METRIC_SAD;
    const bool sampling = cfg->sample_ratio < 1.0f && cfg->sample_ratio > 0.0f;
    const uint32_t break_after = hdr ? 1u : 0u;
    const uint32_t K = cfg->codebook_size;
    const int R = ctx->comm ? ctx->n_ranks : 1, rank = ctx->comm ? ctx->rank : 0;
    const size_t in_bytes = hdr ? (size_t)src.width * src.height * channels * 4 : n_total * bc_block_bytes(src.bc_format);
    const size_t cpx_bytes = (size_t)K * 16 * channels * (hdr ? 4 : 1);
    const size_t acc_words = (size_t)K * 65 + 1;
    // the Lloyd row set of this rank: its block range, or its slice of the sample list
    size_t n_used = n_total;
    if (sampling) n_used = std::min(std::max((size_t)K, (size_t)(n_total * cfg->sample_ratio)), n_total);   // make_draws
    size_t row_first = first_block, n_rows = n_local;
    if (sampling) {
        const size_t per = (n_used + (size_t)R - 1) / (size_t)R;
        row_first = std::min((size_t)rank * per, n_used);
        n_rows = std::min(per, n_used - row_first);
    }
    const size_t nl = std::max<size_t>(1, n_rows);
    const bool tensor = !sad && use_tensor(cfg->assign_mode, n_rows, K);
    Timer tm(ctx);
    const int t0 = tm.mark();

    CK(ctx->bt_feat.ensure(T * nl * 256));
    CK(ctx->bt_assign.ensure(T * nl * 4));
    CK(ctx->bt_errors.ensure(T * nl * 4));
    CK(ctx->bt_cent.ensure(T * (size_t)K * 256));
    CK(ctx->bt_acc.ensure(T * acc_words * 8));
    CK(ctx->bt_misc.ensure(T * 56 + 512));
    CK(ctx->bt_centpx.ensure(T * cpx_bytes));
    if (sad) CK(ctx->bt_planes.ensure(T * nl * 16 * channels));
    if (tensor) {
        CK(ctx->bt_tca.ensure(T * nl * 256));
        CK(ctx->bt_xnorm.ensure(T * nl * 4));
        CK(ctx->bt_tcsmall.ensure(T * TC_SMALL_BYTES));
    }
    auto feat_t = [&](size_t t) { return ctx->bt_feat.as<float>() + t * nl * 64; };
    auto planes_t = [&](size_t t) { return sad ? ctx->bt_planes.as<uint4>() + t * nl * channels : nullptr; };
    auto assign_t = [&](size_t t) { return ctx->bt_assign.as<uint32_t>() + t * nl; };
    auto errors_t = [&](size_t t) { return ctx->bt_errors.as<float>() + t * nl; };
    auto cent_t = [&](size_t t) { return ctx->bt_cent.as<float>() + t * (size_t)K * 64; };
    auto acc_t = [&](size_t t) { return ctx->bt_acc.as<long long>() + t * acc_words; };
    auto centpx_t = [&](size_t t) { return ctx->bt_centpx.as<uint8_t>() + t * cpx_bytes; };
    uint32_t* maxabs_d = ctx->bt_misc.as<uint32_t>();
    uint32_t* empty_d = maxabs_d + T;
    std::vector<TcView> views(T);
    for (size_t t = 0; t < T && tensor; ++t) {
        views[t].a = ctx->bt_tca.as<float>() + t * nl * 64;
        views[t].xnorm = ctx->bt_xnorm.as<float>() + t * nl;
        views[t].small = ctx->bt_tcsmall.as<uint8_t>() + t * TC_SMALL_BYTES;
    }

    // ---- inputs: device resident for the whole call (the final index map decodes them a second time when sampling) ----
    std::vector<const void*> in_dev(T);
    {
        size_t host_count = 0;
        for (size_t t = 0; t < T; ++t) host_count += is_device_ptr(src.data[t]) ? 0 : 1;
        if (host_count) CK(ctx->bt_in.ensure(host_count * in_bytes));
        size_t slot = 0;
        for (size_t t = 0; t < T; ++t) {
            if (is_device_ptr(src.data[t])) {
                in_dev[t] = src.data[t];
            } else {
                uint8_t* d = ctx->bt_in.as<uint8_t>() + slot++ * in_bytes;
                CK(cudaMemcpyAsync(d, src.data[t], in_bytes, cudaMemcpyHostToDevice, ctx->stream));
                in_dev[t] = d;
            }
        }
    }
    // whole texture -> ctx->planes (LDR) and ctx->feat (Oklab features; SAD: the pixel bytes as fp32 rows)
    auto decode_whole = [&](size_t t) -> int {
        CK(ctx->feat.ensure(n_total * 256));
        if (hdr) {
            LAUNCH(launch_image_to_features(reinterpret_cast<const float*>(in_dev[t]), src.width, src.height, channels, ctx->feat.as<float>(),
                                            nullptr, ctx->stream));
        } else {
            CK(ctx->planes.ensure(n_total * 16 * channels));
            LAUNCH(launch_decode_bcn(reinterpret_cast<const uint8_t*>(in_dev[t]), n_total, src.bc_format, channels, ctx->planes.as<uint4>(),
                                     n_total, sad ? nullptr : ctx->feat.as<float>(), nullptr, ctx->stream));
            if (sad) LAUNCH(launch_sad_expand(ctx->planes.as<uint4>(), n_total, n_total, channels, ctx->feat.as<float>(), ctx->stream));
        }
        return 0;
    };
    // row lists: [0, n_local) -> first_block + i (shared by all textures); with sampling each texture's own list in s_rows
    CK(ctx->bt_rows.ensure(std::max<size_t>(1, n_local) * 4));
    {
        std::vector<uint32_t> iota(n_local);
        for (size_t i = 0; i < n_local; ++i) iota[i] = (uint32_t)(first_block + i);
        if (n_local) CK(cudaMemcpyAsync(ctx->bt_rows.p, iota.data(), n_local * 4, cudaMemcpyHostToDevice, ctx->stream));
        CK(cudaStreamSynchronize(ctx->stream));  // iota is stack-owned
    }
    if (sampling) CK(ctx->s_rows.ensure(n_used * 4));

    // ---- phases A + B, texture by texture: decode, (sample,) seed on the owning rank, gather this rank's Lloyd rows -------
    uint32_t seed_rounds = 0;
    CK(cudaMemsetAsync(ctx->bt_cent.p, 0, T * (size_t)K * 256, ctx->stream));
    CK(cudaMemsetAsync(maxabs_d, 0, T * 8, ctx->stream));
    for (size_t t = 0; t < T; ++t) {
        const int owner = (int)(t % (size_t)R);
        if (decode_whole(t)) return 1;
        Blocks all;
        all.n = n_total;
        all.ch = channels;
        all.planes = hdr ? nullptr : ctx->planes.as<uint4>();
        all.stride = hdr ? 0 : n_total;
        all.feat = ctx->feat.as<float>();
        Draws dr;
        if (owner == rank) {
            size_t used = 0;
            dr = make_draws(n_total, K, cfg->sample_ratio, cfg->seed ? cfg->seed + t : 0, &used);
            if (sampling) CK(cudaMemcpyAsync(ctx->s_rows.p, dr.sample.data(), n_used * 4, cudaMemcpyHostToDevice, ctx->stream));
        }
        if (sampling && ctx->comm && ctx->nccl.broadcast_bytes(ctx->comm, ctx->s_rows.p, n_used * 4, owner, ctx->stream))
            return ctx->fail("ncclBroadcast(sample list) failed");
        if (owner == rank) {
            Blocks b = all;
            if (sampling) {
                CK(ctx->s_feat.ensure(n_used * 256));
                if (!hdr) CK(ctx->s_planes.ensure(n_used * 16 * channels));
                LAUNCH(launch_gather_blocks(all.planes, all.stride, all.feat, ctx->s_rows.as<uint32_t>(), n_used, channels,
                                            hdr ? nullptr : ctx->s_planes.as<uint4>(), n_used, ctx->s_feat.as<float>(), ctx->stream));
                b.n = n_used;
                b.planes = hdr ? nullptr : ctx->s_planes.as<uint4>();
                b.stride = hdr ? 0 : n_used;
                b.feat = ctx->s_feat.as<float>();
            } else if (!hdr && attach_palette_form(ctx, b, in_dev[t], src.bc_format, channels)) {
                return 1;
            }
            uint32_t rounds = 0;
            if (run_seeding(ctx, b, K, dr, hdr, stats ? &rounds : nullptr)) return 1;
            seed_rounds += rounds;
            // seeds -> centroid rows: the feature row (SAD: the pixel row) of the chosen block (:324-326, :388-389, :556-558)
            LAUNCH(launch_gather_rows_f32(b.feat, ctx->seed_idx.as<uint32_t>(), K, 64, cent_t(t), ctx->stream));
        }
        if (n_rows) {
            const uint32_t* rows = sampling ? ctx->s_rows.as<uint32_t>() + row_first : ctx->bt_rows.as<uint32_t>();
            LAUNCH(launch_gather_blocks(all.planes, all.stride, all.feat, rows, n_rows, channels, planes_t(t), nl, feat_t(t), ctx->stream));
            if (!sad) LAUNCH(launch_max_abs(feat_t(t), n_rows * 64, maxabs_d + t, ctx->stream));
        }
    }
    if (ctx->comm && ctx->nccl.all_reduce_sum_i64(ctx->comm, ctx->bt_cent.as<long long>(), T * (size_t)K * 32, ctx->stream))
        return ctx->fail("ncclAllReduce(seed rows) failed");
    if (ctx->comm && ctx->nccl.all_reduce_max_u32(ctx->comm, maxabs_d, T, ctx->stream)) return ctx->fail("ncclAllReduce(max) failed");
    uint32_t* h_words = ctx->h_small + 64;
    CK(cudaMemcpyAsync(h_words, maxabs_d, T * 4, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    std::vector<int> shift(T, 0);   // SAD: exact integer sums
    for (size_t t = 0; t < T && !sad; ++t) {
        float m;
        std::memcpy(&m, h_words + t, 4);
        shift[t] = fixed_shift(n_used, m);
    }
    for (size_t t = 0; t < T && tensor; ++t)
        if (tc_prepare_view(ctx, feat_t(t), n_rows, views[t])) return 1;
    const int t1 = tm.mark();

    // ---- Lloyd, all textures in lock step ------------------------------------------------------------------------
    uint32_t* flags = ctx->small.as<uint32_t>();
    uint32_t* ctl_d = maxabs_d + ((2 * T + 31) & ~(size_t)31);   // [T][3]: stop, passes, empty clusters
    uint32_t* tflags_d = ctl_d + 3 * T;                          // [T][4]: changed, empty, reseeds, spare
    uint32_t* sticky_d = tflags_d + 4 * T;
    std::vector<uint32_t> passes(T, 0);
    uint32_t reseeds_async = 0;
    if (ctx->comm) {  // the seeds, should the pass loop have to be redone synchronously
        CK(ctx->bt_cent0.ensure(T * (size_t)K * 256));
        CK(cudaMemcpyAsync(ctx->bt_cent0.p, ctx->bt_cent.p, T * (size_t)K * 256, cudaMemcpyDeviceToDevice, ctx->stream));
    }
    auto reset_state = [&]() -> int {
        CK(cudaMemsetAsync(ctx->bt_assign.p, 0, T * nl * 4, ctx->stream));  // assignments(numBlocks, 0)
        CK(cudaMemsetAsync(flags + F_RESEEDS, 0, 4, ctx->stream));
        CK(cudaMemsetAsync(flags + F_PAIRS, 0, 8, ctx->stream));
        CK(cudaMemsetAsync(ctl_d, 0, (7 * T + 1) * 4, ctx->stream));
        CK(cudaMemsetAsync(flags + F_AMB, 0, 8, ctx->stream));
        ctx->reseeds_sharded = 0;
        return 0;
    };
    // one texture's assignment + partial sums of a pass
    auto assign_and_sum = [&](size_t t, uint32_t* fl, const uint32_t* stop) -> int {
        if (sad) {
            LAUNCH(launch_sad_pack(cent_t(t), K, channels, centpx_t(t), ctx->stream));
            LAUNCH(launch_sad_assign(planes_t(t), nl, n_rows, channels, centpx_t(t), K, assign_t(t), errors_t(t), fl, stop, ctx->stream));
        } else if (assign_blocks(ctx, feat_t(t), n_rows, cent_t(t), K, cfg->assign_mode, assign_t(t), tensor ? nullptr : errors_t(t), fl, stop,
                                 tensor ? &views[t] : nullptr)) {
            return 1;
        }
        LAUNCH(launch_accumulate(feat_t(t), n_rows, assign_t(t), K, shift[t], acc_t(t), fl, stop, ctx->stream));
        return 0;
    };
    auto finalize = [&](size_t t, uint32_t* fl, const uint32_t* stop) -> int {
        if (sad) LAUNCH(launch_sad_finalize(acc_t(t), K, cent_t(t), fl, stop, ctx->stream));
        else LAUNCH(launch_finalize_centroids(acc_t(t), K, shift[t], cent_t(t), fl, stop, ctx->stream));
        return 0;
    };
    auto run_async = [&](bool* redo) -> int {
        *redo = false;
        if (reset_state()) return 1;
        for (uint32_t iter = 0; iter < cfg->max_iterations; ++iter) {
            CK(cudaMemsetAsync(ctx->bt_acc.p, 0, T * acc_words * 8, ctx->stream));
            for (size_t t = 0; t < T; ++t)
                if (assign_and_sum(t, tflags_d + 4 * t, ctl_d + 3 * t)) return 1;
            if (ctx->comm && ctx->nccl.all_reduce_sum_i64(ctx->comm, ctx->bt_acc.as<long long>(), T * acc_words, ctx->stream))
                return ctx->fail("ncclAllReduce(sum) failed");
            LAUNCH(launch_lloyd_control_batch(ctx->bt_acc.as<long long>(), acc_words, (uint32_t)T, K, iter, break_after, tflags_d, 4u, ctl_d,
                                              ctx->comm ? sticky_d : nullptr, ctx->stream));
            for (size_t t = 0; t < T; ++t) {
                const uint32_t* stop = ctl_d + 3 * t;
                if (tensor && !ctx->comm)
                    LAUNCH(launch_exact_errors(feat_t(t), n_rows, cent_t(t), assign_t(t), errors_t(t), stop, ctl_d + 3 * t + 2, ctx->stream));
                if (finalize(t, tflags_d + 4 * t, stop)) return 1;
                if (!ctx->comm)
                    LAUNCH(launch_reseed_empty(feat_t(t), n_rows, errors_t(t), acc_t(t), K, cent_t(t), tflags_d + 4 * t, stop, ctx->stream));
            }
            if ((iter == 3 || iter == 7 || iter == 15) && iter + 1 < cfg->max_iterations) {  // all converged / redo needed?
                CK(cudaMemcpyAsync(h_words, ctl_d, (7 * T + 1) * 4, cudaMemcpyDeviceToHost, ctx->stream));
                CK(cudaStreamSynchronize(ctx->stream));
                bool all_stopped = true;
                for (size_t t = 0; t < T; ++t) all_stopped = all_stopped && h_words[3 * t] != 0;
                if (all_stopped || h_words[7 * T]) break;
            }
        }
        CK(cudaMemcpyAsync(h_words, ctl_d, (7 * T + 1) * 4, cudaMemcpyDeviceToHost, ctx->stream));
        CK(cudaStreamSynchronize(ctx->stream));
        if (h_words[7 * T]) {
            *redo = true;
            return 0;
        }
        for (size_t t = 0; t < T; ++t) {
            passes[t] = h_words[3 * t + 1];
            reseeds_async += h_words[3 * T + 4 * t + 2];
        }
        return 0;
    };
    // host-driven loop (an empty cluster on a rank-sharded row set: the reseed needs the global first maximum)
    auto run_sync = [&]() -> int {
        if (reset_state()) return 1;
        if (ctx->comm) CK(cudaMemcpyAsync(ctx->bt_cent.p, ctx->bt_cent0.p, T * (size_t)K * 256, cudaMemcpyDeviceToDevice, ctx->stream));
        CK(cudaMemsetAsync(ctx->bt_acc.p, 0, T * acc_words * 8, ctx->stream));
        std::vector<char> active(T, 1);
        std::fill(passes.begin(), passes.end(), 0u);
        size_t n_active = T;
        for (uint32_t iter = 0; iter < cfg->max_iterations && n_active; ++iter) {
            for (size_t t = 0; t < T; ++t) {
                if (!active[t]) continue;
                CK(cudaMemsetAsync(flags, 0, 8, ctx->stream));  // changed, empty
                CK(cudaMemsetAsync(acc_t(t), 0, acc_words * 8, ctx->stream));
                if (assign_and_sum(t, flags, nullptr)) return 1;
                ++passes[t];
            }
            if (ctx->comm && ctx->nccl.all_reduce_sum_i64(ctx->comm, ctx->bt_acc.as<long long>(), T * acc_words, ctx->stream))
                return ctx->fail("ncclAllReduce(sum) failed");
            for (size_t t = 0; t < T; ++t)
                if (active[t]) LAUNCH(launch_count_empty(acc_t(t), K, empty_d + t, ctx->stream));
            CK(cudaMemcpy2DAsync(h_words, 8, ctx->bt_acc.as<long long>() + (size_t)K * 65, acc_words * 8, 8, T, cudaMemcpyDeviceToHost,
                                 ctx->stream));
            CK(cudaMemcpyAsync(h_words + 2 * T, empty_d, T * 4, cudaMemcpyDeviceToHost, ctx->stream));
            CK(cudaStreamSynchronize(ctx->stream));
            for (size_t t = 0; t < T; ++t) {
                if (!active[t]) continue;
                long long changed;
                std::memcpy(&changed, h_words + 2 * t, 8);
                const uint32_t n_empty = h_words[2 * T + t];
                if (!changed && iter > break_after) {  // src/vq_encoder.cpp:341, :395, :572
                    active[t] = 0;
                    --n_active;
                    continue;
                }
                if (tensor && n_empty)  // errors[i] = distance to the assigned centroid, before the update
                    LAUNCH(launch_exact_errors(feat_t(t), n_rows, cent_t(t), assign_t(t), errors_t(t), nullptr, nullptr, ctx->stream));
                CK(cudaMemsetAsync(flags + F_EMPTY, 0, 4, ctx->stream));
                if (finalize(t, flags, nullptr)) return 1;
                if (!n_empty) continue;
                if (!ctx->comm) {
                    LAUNCH(launch_reseed_empty(feat_t(t), n_rows, errors_t(t), acc_t(t), K, cent_t(t), flags, nullptr, ctx->stream));
                } else {
                    Blocks mine;
                    mine.n = n_rows;
                    mine.ch = channels;
                    mine.feat = feat_t(t);
                    if (reseed_sharded(ctx, mine, row_first, K, cent_t(t), errors_t(t), acc_t(t), flags)) return 1;
                }
            }
        }
        return 0;
    };
    bool redo = debug_opts().lloyd_sync;
    if (!redo && run_async(&redo)) return 1;
    if (redo && run_sync()) return 1;
    const int t2 = tm.mark();

    // ---- centroids -> pixels / floats, final index map of this rank's block range (QuantizeBlocks[HDR]) ----------------
    for (size_t t = 0; t < T; ++t) {
        if (sad) LAUNCH(launch_sad_pack(cent_t(t), K, channels, centpx_t(t), ctx->stream));   // the centroids are pixel blocks already
        else if (hdr) LAUNCH(launch_oklab_to_floats(cent_t(t), K, channels, reinterpret_cast<float*>(centpx_t(t)), ctx->stream));
        else LAUNCH(launch_oklab_to_pixels(cent_t(t), K, channels, centpx_t(t), ctx->stream));
        Blocks shard;
        shard.n = n_local;
        shard.ch = channels;
        if (!sampling) {   // the Lloyd rows are this rank's blocks
            shard.planes = planes_t(t);
            shard.stride = nl;
            shard.feat = feat_t(t);
        } else if (n_local) {
            if (decode_whole(t)) return 1;
            CK(ctx->s_feat.ensure(n_local * 256));
            if (sad) CK(ctx->s_planes.ensure(n_local * 16 * channels));
            LAUNCH(launch_gather_blocks(hdr ? nullptr : ctx->planes.as<uint4>(), n_total, ctx->feat.as<float>(), ctx->bt_rows.as<uint32_t>(), n_local,
                                        channels, sad ? ctx->s_planes.as<uint4>() : nullptr, n_local, ctx->s_feat.as<float>(), ctx->stream));
            shard.planes = sad ? ctx->s_planes.as<uint4>() : nullptr;
            shard.stride = n_local;
            shard.feat = ctx->s_feat.as<float>();
        }
        ctx->tc_feat = nullptr;   // the operand cache is keyed on the feature pointer: these rows are new
        if (n_local && quantize_device(ctx, shard, centpx_t(t), K, hdr, cfg->assign_mode, sad ? YUPVQ_METRIC_SAD : YUPVQ_METRIC_OKLAB)) return 1;
        if (out_centroids && from_device(ctx, out_centroids[t], centpx_t(t), cpx_bytes)) return 1;
        if (out_indices && n_local && from_device(ctx, out_indices[t], ctx->out_idx.p, n_local * 4)) return 1;
    }
    const int t3 = tm.mark();
    CK(cudaMemcpyAsync(ctx->h_small, flags + F_RESEEDS, 4, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaMemcpyAsync(ctx->h_small + 1, flags + F_PAIRS, 8, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    if (stats) {
        stats->n_blocks = (uint32_t)(n_rows * T);
        stats->codebook_size = K;
        stats->seed_rounds = seed_rounds;
        for (size_t t = 0; t < T; ++t) stats->lloyd_passes += passes[t];
        stats->reseeds = ctx->h_small[0] + ctx->reseeds_sharded + reseeds_async;
        stats->rerank_blocks = ctx->h_small[1];
        stats->rerank_overflow = ctx->h_small[2];
        stats->ms_seed = tm.ms(t0, t1);      // upload, whole decode, seeding of the owned textures, row gather
        stats->ms_lloyd = tm.ms(t1, t2);
        stats->ms_quantize = tm.ms(t2, t3);  // incl. centroid conversion and the D2H of the results
        stats->ms_total = tm.ms(t0, t3);
        stats->gpu_launches = ctx->launches;
    }
    return 0;
}

}  // namespace

extern "C" {

// Lock-step batch over block shards: T textures of n_total blocks each, this rank
// holds blocks [first_block, first_block + n_local) of EVERY texture.
//   phase A  texture-parallel seeding: texture t is decoded whole and seeded on rank t % R only (K-Means++
//            is K-1 dependent rounds: latency, not bandwidth, so it is spread over textures rather than
//            replicated); the K seed rows of all textures are then combined with one all-reduce (every
//            row is non-zero on exactly one rank: an integer sum with zeros is exact).
//   phase B  every rank decodes its shard of every texture.
//   Lloyd    all textures advance one pass at a time; ONE all-reduce per pass carries the fixed-point
//            [K x 64 sums | K counts | changed] of all T textures, and one host readback carries the T
//            convergence / empty-cluster words.  A texture leaves the active set when the reference's
//            break rule fires for it (src/vq_encoder.cpp:341).
// Results per texture are identical to yupvq_encode_bcn with seed + t (fixed-point sums).
int yupvq_encode_bcn_batch_sharded(yupvq_ctx* ctx, const uint8_t* const* bc_blocks, size_t n_textures, size_t n_total,
                                   size_t first_block, size_t n_local, int bc_format, int channels,
                                   const yupvq_config* cfg, uint8_t* const* out_pixel_centroids,
                                   uint32_t* const* out_indices, yupvq_stats* stats, const uint8_t* shard_only) {
    if (!ctx) return 1;
    Scope scope(ctx->device);
    init_stats(stats);
    ctx->launches = 0;
    ctx->tc_feat = nullptr;
    ctx->maxabs_feat = nullptr;
    ctx->tc_rgb = rgb_layout(channels);
    ctx->xchg_path = ctx->comm ? YUPVQ_XCHG_NCCL : YUPVQ_XCHG_NONE;   // the peer path overwrites it where it runs
    if (check_cfg(ctx, cfg, n_total, channels)) return 1;
    if (!bc_blocks || n_textures == 0) return ctx->fail("no textures");
    if (n_textures > 4096) return ctx->fail("too many textures in one batch (%zu > 4096)", n_textures);
    if (!bc_supported(bc_format)) return ctx->fail("BC format %d is not supported on the LDR VQ path", bc_format);
    if (first_block + n_local > n_total) return ctx->fail("shard [%zu, %zu) exceeds %zu blocks", first_block, first_block + n_local, n_total);
    if ((cfg->sample_ratio < 1.0f && cfg->sample_ratio > 0.0f) || cfg->metric != YUPVQ_METRIC_OKLAB) {
        // fast-mode sampling / SAD metric: the general lock-step batch (whole textures on every rank)
        for (size_t t = 0; t < n_textures && shard_only; ++t)
            if (shard_only[t]) return ctx->fail("fast-mode sampling and the SAD metric need the whole texture on every rank (shard_only[%zu] is set)", t);
        BatchSrc src;
        src.data = reinterpret_cast<const void* const*>(bc_blocks);
        src.bc_format = bc_format;
        return batch_sharded_general(ctx, src, n_textures, n_total, first_block, n_local, channels, cfg,
                                     reinterpret_cast<void* const*>(out_pixel_centroids), out_indices, stats);
    }
    const uint32_t K = cfg->codebook_size;
    const size_t T = n_textures;
    const int R = ctx->comm ? ctx->n_ranks : 1, rank = ctx->comm ? ctx->rank : 0;
    const size_t bb = bc_block_bytes(bc_format);
    const size_t nl = std::max<size_t>(1, n_local);
    const size_t acc_words = (size_t)K * 65 + 1;
    const bool tensor = use_tensor(cfg->assign_mode, n_local, K);
    Timer tm(ctx);
    const int t0 = tm.mark();

    CK(ctx->bt_feat.ensure(T * nl * 256));
    CK(ctx->bt_assign.ensure(T * nl * 4));
    CK(ctx->bt_errors.ensure(T * nl * 4));
    CK(ctx->bt_cent.ensure(T * (size_t)K * 256));
    CK(ctx->bt_acc.ensure(T * acc_words * 8));
    CK(ctx->bt_misc.ensure(T * 56 + 512));  // maxabs | empty | (pad) ctl[3] | flags[4] | sticky | scales f32 | inv_scales f64
    CK(ctx->bt_centpx.ensure(T * (size_t)K * 16 * channels));
    if (tensor) {
        CK(ctx->bt_tca.ensure(T * nl * 256));
        CK(ctx->bt_xnorm.ensure(T * nl * 4));
        CK(ctx->bt_tcsmall.ensure(T * TC_SMALL_BYTES));
    }
    auto feat_t = [&](size_t t) { return ctx->bt_feat.as<float>() + t * nl * 64; };
    auto assign_t = [&](size_t t) { return ctx->bt_assign.as<uint32_t>() + t * nl; };
    auto errors_t = [&](size_t t) { return ctx->bt_errors.as<float>() + t * nl; };
    auto cent_t = [&](size_t t) { return ctx->bt_cent.as<float>() + t * (size_t)K * 64; };
    auto acc_t = [&](size_t t) { return ctx->bt_acc.as<long long>() + t * acc_words; };
    auto centpx_t = [&](size_t t) { return ctx->bt_centpx.as<uint8_t>() + t * (size_t)K * 16 * channels; };
    uint32_t* maxabs_d = ctx->bt_misc.as<uint32_t>();
    uint32_t* empty_d = maxabs_d + T;
    std::vector<TcView> views(T);
    for (size_t t = 0; t < T && tensor; ++t) {
        views[t].a = ctx->bt_tca.as<float>() + t * nl * 64;
        views[t].xnorm = ctx->bt_xnorm.as<float>() + t * nl;
        views[t].small = ctx->bt_tcsmall.as<uint8_t>() + t * TC_SMALL_BYTES;
    }

    // ---- phase A: seeding, texture-parallel --------------------------------------------------------------
    uint32_t seed_rounds = 0;
    CK(cudaMemsetAsync(ctx->bt_cent.p, 0, T * (size_t)K * 256, ctx->stream));
    for (size_t t = 0; t < T; ++t) {
        if ((int)(t % (size_t)R) != rank) continue;
        if (shard_only && shard_only[t])
            return ctx->fail("texture %zu is seeded on this rank (t %% n_ranks == rank): its whole BCn data is needed, not a shard", t);
        const void* bc_dev = nullptr;
        if (to_device(ctx, bc_blocks[t], n_total * bb, ctx->in_bytes, &bc_dev)) return 1;
        // pixel planes only: the seeding reads nothing else, and the fp64 colour conversion (the expensive part of the
        // decode) is then done for the K chosen blocks instead of the whole texture
        CK(ctx->planes.ensure(n_total * 16 * channels));
        LAUNCH(launch_decode_bcn(reinterpret_cast<const uint8_t*>(bc_dev), n_total, bc_format, channels,
                                 ctx->planes.as<uint4>(), n_total, nullptr, nullptr, ctx->stream));
        Blocks all;
        all.n = n_total;
        all.ch = channels;
        all.planes = ctx->planes.as<uint4>();
        all.stride = n_total;
        all.feat = nullptr;
        if (attach_palette_form(ctx, all, bc_dev, bc_format, channels)) return 1;
        size_t n_used;
        Draws dr = make_draws(n_total, K, 1.0f, cfg->seed ? cfg->seed + t : 0, &n_used);
        uint32_t rounds = 0;
        if (run_seeding(ctx, all, K, dr, false, stats ? &rounds : nullptr)) return 1;
        seed_rounds += rounds;
        CK(ctx->cent_px.ensure((size_t)K * 16 * channels));
        LAUNCH(launch_planes_to_blocks(all.planes, all.stride, ctx->seed_idx.as<uint32_t>(), K, channels, ctx->cent_px.as<uint8_t>(),
                                       ctx->stream));
        LAUNCH(launch_pixels_to_features(ctx->cent_px.as<uint8_t>(), K, channels, nullptr, 0, cent_t(t), ctx->stream));
    }
    if (ctx->comm && ctx->nccl.all_reduce_sum_i64(ctx->comm, ctx->bt_cent.as<long long>(), T * (size_t)K * 32, ctx->stream))
        return ctx->fail("ncclAllReduce(seed rows) failed");
    const int t1 = tm.mark();

    // ---- phase B: this rank's shard of every texture -> features, fixed-point scale, tensor operands --------
    const bool fuse_operands = tensor && !debug_opts().no_decode_fuse;
    CK(cudaMemsetAsync(maxabs_d, 0, T * 8, ctx->stream));
    for (size_t t = 0; t < T; ++t) {
        const uint8_t* shard_dev = nullptr;
        // shard_only[t]: the caller passed just blocks [first_block, first_block + n_local) of this texture
        const uint8_t* shard_src = bc_blocks[t] + ((shard_only && shard_only[t]) ? 0 : first_block * bb);
        if (is_device_ptr(bc_blocks[t])) {
            shard_dev = shard_src;
        } else {
            const void* p = nullptr;
            if (to_device(ctx, shard_src, nl * bb, ctx->in_bytes, &p)) return 1;
            shard_dev = reinterpret_cast<const uint8_t*>(p);
        }
        if (n_local && fuse_operands) {
            // fused feature pass (see yupvq_encode_bcn): mu = the mean of this texture's K seed rows (identical on every rank)
            uint8_t* sm = views[t].small;
            LAUNCH(launch_tc_mean(cent_t(t), K, reinterpret_cast<double*>(sm + TC_SUMS), reinterpret_cast<float*>(sm + TC_MU), ctx->stream));
            ++ctx->launches;
            DecodeFuse fu;
            fu.mu = reinterpret_cast<const float*>(sm + TC_MU);
            fu.a_prime = views[t].a;
            fu.xnorm = views[t].xnorm;
            fu.max_abs_bits = maxabs_d + t;
            fu.rgb = ctx->tc_rgb;
            LAUNCH(launch_decode_bcn(shard_dev, n_local, bc_format, channels, nullptr, 0, feat_t(t), nullptr, ctx->stream, &fu));
        } else if (n_local) {
            LAUNCH(launch_decode_bcn(shard_dev, n_local, bc_format, channels, nullptr, 0, feat_t(t), nullptr, ctx->stream));
            LAUNCH(launch_max_abs(feat_t(t), n_local * 64, maxabs_d + t, ctx->stream));
        }
    }
    if (ctx->comm && ctx->nccl.all_reduce_max_u32(ctx->comm, maxabs_d, T, ctx->stream)) return ctx->fail("ncclAllReduce(max) failed");
    uint32_t* h_words = ctx->h_small + 64;
    CK(cudaMemcpyAsync(h_words, maxabs_d, T * 4, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    std::vector<int> shift(T);
    for (size_t t = 0; t < T; ++t) {
        float m;
        std::memcpy(&m, h_words + t, 4);
        shift[t] = fixed_shift(n_total, m);
    }
    for (size_t t = 0; t < T && tensor && !fuse_operands; ++t)
        if (tc_prepare_view(ctx, feat_t(t), n_local, views[t])) return 1;
    const int t2 = tm.mark();

    // ---- Lloyd, all textures in lock step ------------------------------------------------------------------------
    uint32_t* flags = ctx->small.as<uint32_t>();
    // control words | per-texture flags | sticky: one contiguous run (cleared and read back together); the flags are
    // written by the kernels while the stop words are polled, so the run starts on its own 128-byte line
    uint32_t* ctl_d = maxabs_d + ((2 * T + 31) & ~(size_t)31);   // [T][3]: stop, passes, empty clusters
    uint32_t* tflags_d = ctl_d + 3 * T;   // [T][4]: changed, empty, reseeds, spare
    uint32_t* sticky_d = tflags_d + 4 * T;
    std::vector<uint32_t> passes(T, 0);
    uint32_t reseeds_async = 0;
    const bool force_sync = debug_opts().lloyd_sync;  // A/B timing and tests
    if (ctx->comm) {  // the seeds, should the pass loop have to be redone synchronously
        CK(ctx->bt_cent0.ensure(T * (size_t)K * 256));
        CK(cudaMemcpyAsync(ctx->bt_cent0.p, ctx->bt_cent.p, T * (size_t)K * 256, cudaMemcpyDeviceToDevice, ctx->stream));
    }
    auto reset_state = [&]() -> int {
        CK(cudaMemsetAsync(ctx->bt_assign.p, 0, T * nl * 4, ctx->stream));  // assignments(numBlocks, 0), :316
        CK(cudaMemsetAsync(flags + F_RESEEDS, 0, 4, ctx->stream));
        CK(cudaMemsetAsync(flags + F_PAIRS, 0, 8, ctx->stream));
        CK(cudaMemsetAsync(ctl_d, 0, (7 * T + 1) * 4, ctx->stream));
        CK(cudaMemsetAsync(flags + F_AMB, 0, 8, ctx->stream));  // ambiguous-row count, ticket (the re-rank leaves them cleared)
        ctx->reseeds_sharded = 0;
        return 0;
    };
    // Asynchronous loop: no host round trip per pass.  The reference's convergence test runs on the device for
    // every texture (launch_lloyd_control_batch); kernels of a stopped texture return at once.  Block-sharded
    // runs cannot reseed an empty cluster on the device: the first one raises `sticky` and the host redoes the
    // batch with the synchronous loop below (rare: flat or few-colour textures).
    auto run_async = [&](bool* redo) -> int {
        *redo = false;
        if (reset_state()) return 1;
        for (uint32_t iter = 0; iter < cfg->max_iterations; ++iter) {
            CK(cudaMemsetAsync(ctx->bt_acc.p, 0, T * acc_words * 8, ctx->stream));
            for (size_t t = 0; t < T; ++t) {
                const uint32_t* stop = ctl_d + 3 * t;
                if (assign_blocks(ctx, feat_t(t), n_local, cent_t(t), K, cfg->assign_mode, assign_t(t), tensor ? nullptr : errors_t(t),
                                  tflags_d + 4 * t, stop, tensor ? &views[t] : nullptr))
                    return 1;
                LAUNCH(launch_accumulate(feat_t(t), n_local, assign_t(t), K, shift[t], acc_t(t), tflags_d + 4 * t, stop, ctx->stream));
            }
            if (ctx->comm && ctx->nccl.all_reduce_sum_i64(ctx->comm, ctx->bt_acc.as<long long>(), T * acc_words, ctx->stream))
                return ctx->fail("ncclAllReduce(sum) failed");
            LAUNCH(launch_lloyd_control_batch(ctx->bt_acc.as<long long>(), acc_words, (uint32_t)T, K, iter, 0u, tflags_d, 4u, ctl_d,
                                              ctx->comm ? sticky_d : nullptr, ctx->stream));
            for (size_t t = 0; t < T; ++t) {
                const uint32_t* stop = ctl_d + 3 * t;
                if (tensor && !ctx->comm)
                    LAUNCH(launch_exact_errors(feat_t(t), n_local, cent_t(t), assign_t(t), errors_t(t), stop, ctl_d + 3 * t + 2, ctx->stream));
                LAUNCH(launch_finalize_centroids(acc_t(t), K, shift[t], cent_t(t), tflags_d + 4 * t, stop, ctx->stream));
                if (!ctx->comm)
                    LAUNCH(launch_reseed_empty(feat_t(t), n_local, errors_t(t), acc_t(t), K, cent_t(t), tflags_d + 4 * t, stop, ctx->stream));
            }
            if ((iter == 3 || iter == 7 || iter == 15) && iter + 1 < cfg->max_iterations) {  // all converged / redo needed?
                CK(cudaMemcpyAsync(h_words, ctl_d, (7 * T + 1) * 4, cudaMemcpyDeviceToHost, ctx->stream));
                CK(cudaStreamSynchronize(ctx->stream));
                bool all_stopped = true;
                for (size_t t = 0; t < T; ++t) all_stopped = all_stopped && h_words[3 * t] != 0;
                if (all_stopped || h_words[7 * T]) break;
            }
        }
        CK(cudaMemcpyAsync(h_words, ctl_d, (7 * T + 1) * 4, cudaMemcpyDeviceToHost, ctx->stream));
        CK(cudaStreamSynchronize(ctx->stream));
        if (h_words[7 * T]) {
            *redo = true;
            return 0;
        }
        for (size_t t = 0; t < T; ++t) {
            passes[t] = h_words[3 * t + 1];
            reseeds_async += h_words[3 * T + 4 * t + 2];
        }
        return 0;
    };
    auto run_sync = [&]() -> int {
        if (reset_state()) return 1;
        if (ctx->comm) CK(cudaMemcpyAsync(ctx->bt_cent.p, ctx->bt_cent0.p, T * (size_t)K * 256, cudaMemcpyDeviceToDevice, ctx->stream));
        CK(cudaMemsetAsync(ctx->bt_acc.p, 0, T * acc_words * 8, ctx->stream));
        std::vector<char> active(T, 1);
        std::fill(passes.begin(), passes.end(), 0u);
        size_t n_active = T;
        for (uint32_t iter = 0; iter < cfg->max_iterations && n_active; ++iter) {
            for (size_t t = 0; t < T; ++t) {
                if (!active[t]) continue;
                CK(cudaMemsetAsync(flags, 0, 8, ctx->stream));  // changed, empty
                if (assign_blocks(ctx, feat_t(t), n_local, cent_t(t), K, cfg->assign_mode, assign_t(t), tensor ? nullptr : errors_t(t),
                                  flags, nullptr, tensor ? &views[t] : nullptr))
                    return 1;
                ++passes[t];
                CK(cudaMemsetAsync(acc_t(t), 0, acc_words * 8, ctx->stream));
                LAUNCH(launch_accumulate(feat_t(t), n_local, assign_t(t), K, shift[t], acc_t(t), flags, nullptr, ctx->stream));
            }
            // one exchange step per pass for the whole batch (converged textures ride along: their slices are stale, unused)
            if (ctx->comm && ctx->nccl.all_reduce_sum_i64(ctx->comm, ctx->bt_acc.as<long long>(), T * acc_words, ctx->stream))
                return ctx->fail("ncclAllReduce(sum) failed");
            for (size_t t = 0; t < T; ++t)
                if (active[t]) LAUNCH(launch_count_empty(acc_t(t), K, empty_d + t, ctx->stream));
            CK(cudaMemcpy2DAsync(h_words, 8, ctx->bt_acc.as<long long>() + (size_t)K * 65, acc_words * 8, 8, T, cudaMemcpyDeviceToHost,
                                 ctx->stream));
            CK(cudaMemcpyAsync(h_words + 2 * T, empty_d, T * 4, cudaMemcpyDeviceToHost, ctx->stream));
            CK(cudaStreamSynchronize(ctx->stream));
            for (size_t t = 0; t < T; ++t) {
                if (!active[t]) continue;
                long long changed;
                std::memcpy(&changed, h_words + 2 * t, 8);
                const uint32_t n_empty = h_words[2 * T + t];
                if (!changed && iter > 0) {  // src/vq_encoder.cpp:341
                    active[t] = 0;
                    --n_active;
                    continue;
                }
                if (tensor && n_empty)  // errors[i] = distance to the assigned centroid, before the update (:336)
                    LAUNCH(launch_exact_errors(feat_t(t), n_local, cent_t(t), assign_t(t), errors_t(t), nullptr, nullptr, ctx->stream));
                CK(cudaMemsetAsync(flags + F_EMPTY, 0, 4, ctx->stream));
                LAUNCH(launch_finalize_centroids(acc_t(t), K, shift[t], cent_t(t), flags, nullptr, ctx->stream));
                if (!n_empty) continue;
                if (!ctx->comm) {
                    LAUNCH(launch_reseed_empty(feat_t(t), n_local, errors_t(t), acc_t(t), K, cent_t(t), flags, nullptr, ctx->stream));
                } else {
                    Blocks mine;
                    mine.n = n_local;
                    mine.ch = channels;
                    mine.feat = feat_t(t);
                    if (reseed_sharded(ctx, mine, first_block, K, cent_t(t), errors_t(t), acc_t(t), flags)) return 1;
                }
            }
        }
        return 0;
    };
    // Whole-batch launches: when the shards are tile-aligned the rows of all textures are one matrix, so every kernel
    // of a pass is launched ONCE for the batch (the front-end reloads its resident centroid operand when a CTA crosses
    // into the next texture) -- a pass costs the same handful of launches whatever the number of textures.
    const bool no_fuse = debug_opts().batch_per_texture;  // A/B timing and tests
    const uint32_t k_pad = tc_padded_k(K);
    const bool fused = !no_fuse && tensor && n_local > 0 && n_local % 128 == 0 && k_pad <= 512 && accumulate_batch_supported(n_local, K) &&
                       T * n_local < 0xffffffffull;
    TcBatch tb{};
    AccBatch ab{};
    float* scales_d = reinterpret_cast<float*>(sticky_d + 1);
    double* inv_scales_d = reinterpret_cast<double*>(ctx->bt_misc.as<uint8_t>() + ((T * 48 + 256 + 7) & ~(size_t)7));
    if (fused) {
        CK(ctx->tc_b.ensure(T * (size_t)k_pad * 256));
        CK(ctx->tc_bx.ensure(T * (size_t)k_pad * 32));
        CK(ctx->tc_cand.ensure(T * nl * 16));
        CK(ctx->tc_amb.ensure(T * nl * 4));
        std::vector<float> sc(T);
        std::vector<double> isc(T);
        for (size_t t = 0; t < T; ++t) {
            sc[t] = std::ldexp(1.0f, shift[t]);
            isc[t] = std::ldexp(1.0, -shift[t]);
        }
        CK(cudaMemcpyAsync(scales_d, sc.data(), T * 4, cudaMemcpyHostToDevice, ctx->stream));
        CK(cudaMemcpyAsync(inv_scales_d, isc.data(), T * 8, cudaMemcpyHostToDevice, ctx->stream));
        CK(cudaStreamSynchronize(ctx->stream));  // sc / isc are stack-owned
        tb.tiles_per_tex = (uint32_t)(n_local / 128);
        tb.k_pad = k_pad;
        tb.cmax_stride = TC_SMALL_BYTES / 4;
        tb.flags_stride = 4;
        tb.stop_stride = 3;
        tb.rows_per_tex = (uint32_t)n_local;
        tb.n_textures = (uint32_t)T;
        ab.n_textures = (uint32_t)T;
        ab.acc_stride = acc_words;
        ab.scales = scales_d;
        ab.flags_stride = 4;
        ab.stop_stride = 3;
    }
    // one assignment pass over the whole batch; out == bt_assign inside the Lloyd loop, the index map for the quantise
    auto assign_fused = [&](const float* cent_all, uint32_t* out, uint32_t* fl, const uint32_t* stop) -> int {
        uint8_t* sm = ctx->bt_tcsmall.as<uint8_t>();
        TcBatch b = tb;
        if (!fl) {  // the quantise pass: no changed flags, no moved-row list
            b.flags_stride = 0;
            b.chg_rows = b.chg_old = nullptr;
            b.chg_count = nullptr;
        }
        if (!stop) b.stop_stride = 0;
        b.rgb = (uint32_t)ctx->tc_rgb;
        LAUNCH(launch_tc_prep_centroids_batch(cent_all, K, (uint32_t)T, sm, TC_SMALL_BYTES, TC_MU, TC_CMAX, ctx->tc_b.as<float>(),
                                              ctx->tc_bx.as<float>(), ctx->stream, ctx->tc_rgb));
        const bool prof = ctx->profile && ctx->pev_used + 2 <= ctx->pev.size();
        if (prof) cudaEventRecord(ctx->pev[ctx->pev_used++], ctx->stream);
        LAUNCH(launch_assign_tc(ctx->bt_tca.as<float>(), ctx->bt_xnorm.as<float>(), T * n_local, ctx->tc_b.as<float>(), ctx->tc_bx.as<float>(),
                                K, reinterpret_cast<const float*>(sm + TC_CMAX), ctx->tc_cand.as<uint4>(), out, fl,
                                ctx->tc_amb.as<uint32_t>(), flags + F_AMB, stop, ctx->stream, &b));
        if (prof) cudaEventRecord(ctx->pev[ctx->pev_used++], ctx->stream);
        LAUNCH(launch_rerank_sparse(ctx->bt_feat.as<float>(), cent_all, K, ctx->tc_cand.as<uint4>(), ctx->tc_amb.as<uint32_t>(), flags + F_AMB,
                                    flags + F_AMB + 1, out, fl, flags + F_PAIRS, ctx->stream, &b));
        return 0;
    };
    auto run_async_fused = [&](bool* redo) -> int {
        *redo = false;
        if (reset_state()) return 1;
        // Delta update: this rank's partial sums persist in bt_acc_local and are adjusted by the rows that
        // moved; a copy goes through the all-reduce.  One list and one mode for the whole batch.
        const bool no_delta = debug_opts().no_delta;
        const bool delta = !no_delta && delta_batch_supported(K, (uint32_t)T);
        long long* acc_loc = ctx->bt_acc.as<long long>();
        uint32_t* dctl = nullptr;
        tb.chg_rows = tb.chg_old = nullptr;
        tb.chg_count = nullptr;
        if (delta) {
            CK(ctx->tc_chg.ensure(T * nl * 8));
            CK(ctx->delta_ctl.ensure(256));
            dctl = ctx->delta_ctl.as<uint32_t>();
            CK(cudaMemsetAsync(dctl, 0, 256, ctx->stream));
            tb.chg_rows = ctx->tc_chg.as<uint32_t>();
            tb.chg_old = tb.chg_rows + T * nl;
            tb.chg_count = dctl;
            if (ctx->comm) {
                CK(ctx->bt_acc_local.ensure(T * acc_words * 8));
                acc_loc = ctx->bt_acc_local.as<long long>();
            }
        }
        // Fused pass for the block-sharded batch (no device reseed there): front-end, re-rank (+ update prologue), one update
        // kernel, [copy + all-reduce], one tail kernel for all textures (control, finalize, next operands).
        const bool fuse_tail = ctx->comm && delta && !debug_opts().no_tail_fuse && accumulate_auto_supported(n_local, K, (uint32_t)T);
        // the exchange step inside the tail kernel, over peer memory (no copy, no NCCL launch); else copy + ncclAllReduce
        bool peer_xchg = false;
        if (fuse_tail && xchg_ensure(ctx, T * acc_words, &peer_xchg)) return 1;
        if (fuse_tail) {
            uint8_t* sm = ctx->bt_tcsmall.as<uint8_t>();
            CK(cudaMemset2DAsync(sm + TC_CMAX, TC_SMALL_BYTES, 0, 8, T, ctx->stream));
            LAUNCH(launch_tc_prep_centroids_batch(ctx->bt_cent.as<float>(), K, (uint32_t)T, sm, TC_SMALL_BYTES, TC_MU, TC_CMAX, ctx->tc_b.as<float>(),
                                                  ctx->tc_bx.as<float>(), ctx->stream, ctx->tc_rgb));
            TcBatch b = tb;
            b.rgb = (uint32_t)ctx->tc_rgb;
            AccPrep prep{acc_loc, dctl, K, (uint32_t)(T * n_local / 4), 0u, (uint32_t)T, acc_words};
            TailArgs ta{};
            ta.acc = ctx->bt_acc.as<long long>();
            ta.acc_stride = acc_words;
            ta.n_textures = (uint32_t)T;
            ta.K = K;
            ta.k_pad = k_pad;
            ta.break_after = 0u;
            ta.flags = tflags_d;
            ta.flags_stride = 4;
            ta.ctl = ctl_d;
            ta.sticky = sticky_d;
            ta.inv_scales = inv_scales_d;
            ta.cent = ctx->bt_cent.as<float>();
            ta.small = sm;
            ta.small_stride = TC_SMALL_BYTES;
            ta.mu_off = TC_MU;
            ta.cmax_off = TC_CMAX;
            ta.b_main = ctx->tc_b.as<float>();
            ta.b_ext = ctx->tc_bx.as<float>();
            ta.rgb = ctx->tc_rgb;
            for (uint32_t iter = 0; iter < cfg->max_iterations; ++iter) {
                const float* cmax = reinterpret_cast<const float*>(sm + TC_CMAX) + (iter & 1u);
                const bool prof = ctx->profile && ctx->pev_used + 2 <= ctx->pev.size();
                if (prof) cudaEventRecord(ctx->pev[ctx->pev_used++], ctx->stream);
                ctx->stage_mark(255);
                LAUNCH(launch_assign_tc(ctx->bt_tca.as<float>(), ctx->bt_xnorm.as<float>(), T * n_local, ctx->tc_b.as<float>(), ctx->tc_bx.as<float>(), K,
                                        cmax, ctx->tc_cand.as<uint4>(), ctx->bt_assign.as<uint32_t>(), tflags_d, ctx->tc_amb.as<uint32_t>(),
                                        flags + F_AMB, ctl_d, ctx->stream, &b));
                if (prof) cudaEventRecord(ctx->pev[ctx->pev_used++], ctx->stream);
                ctx->stage_mark(0);
                prep.force_full = iter == 0 ? 1u : 0u;
                LAUNCH(launch_rerank_sparse(ctx->bt_feat.as<float>(), ctx->bt_cent.as<float>(), K, ctx->tc_cand.as<uint4>(), ctx->tc_amb.as<uint32_t>(),
                                            flags + F_AMB, flags + F_AMB + 1, ctx->bt_assign.as<uint32_t>(), tflags_d, flags + F_PAIRS, ctx->stream,
                                            &b, &prep));
                ctx->stage_mark(1);
                LAUNCH(launch_accumulate_auto(ctx->bt_feat.as<float>(), n_local, ctx->bt_assign.as<uint32_t>(), K, 0, acc_loc, tb.chg_rows, tb.chg_old,
                                              dctl, tflags_d, ctl_d, ctx->stream, &ab));
                ctx->stage_mark(2);
                if (peer_xchg) {
                    if (iter == 2 && debug_opts().xchg_desync_rank == ctx->rank) ++ctx->xchg_epoch;   // test hook
                    xchg_args(ctx, acc_loc, T * acc_words, &ta.x, &ta.acc);
                    ctx->xchg_path = ta.x.one_shot ? YUPVQ_XCHG_PEER_ONE_SHOT : YUPVQ_XCHG_PEER_TWO_SHOT;
                } else {
                    CK(cudaMemcpyAsync(ctx->bt_acc.p, acc_loc, T * acc_words * 8, cudaMemcpyDeviceToDevice, ctx->stream));
                    if (ctx->nccl.all_reduce_sum_i64(ctx->comm, ctx->bt_acc.as<long long>(), T * acc_words, ctx->stream))
                        return ctx->fail("ncclAllReduce(sum) failed");
                }
                ta.iter = iter;
                LAUNCH(launch_lloyd_tail(ta, ctx->stream));
                ctx->stage_mark(3);
                // all converged / redo needed?  Lagged poll (see run_lloyd): read back behind the next pass, no drain
                if ((iter == 3 || iter == 7 || iter == 15) && iter + 1 < cfg->max_iterations) {
                    CK(cudaMemcpyAsync(h_words, ctl_d, (7 * T + 1) * 4, cudaMemcpyDeviceToHost, ctx->stream));
                    CK(cudaEventRecord(ctx->ev_poll, ctx->stream));
                }
                if (iter == 4 || iter == 8 || iter == 16) {
                    CK(cudaEventSynchronize(ctx->ev_poll));
                    bool all_stopped = true;
                    for (size_t t = 0; t < T; ++t) all_stopped = all_stopped && h_words[3 * t] != 0;
                    if (all_stopped || h_words[7 * T]) break;
                }
            }
        }
        for (uint32_t iter = 0; iter < cfg->max_iterations && !fuse_tail; ++iter) {
            if (assign_fused(ctx->bt_cent.as<float>(), ctx->bt_assign.as<uint32_t>(), tflags_d, ctl_d)) return 1;
            if (delta) {
                LAUNCH(launch_acc_prepare_batch(acc_loc, K, T * n_local, iter == 0 ? 1u : 0u, dctl, ab, ctx->stream));
                LAUNCH(launch_accumulate_batch(ctx->bt_feat.as<float>(), n_local, ctx->bt_assign.as<uint32_t>(), K, acc_loc, tflags_d, ctl_d,
                                               ab, ctx->stream, dctl));
                LAUNCH(launch_accumulate_delta(ctx->bt_feat.as<float>(), ctx->bt_assign.as<uint32_t>(), K, 0, acc_loc, tb.chg_rows, tb.chg_old,
                                               dctl, tflags_d, ctl_d, ctx->stream, &ab, (uint32_t)n_local));
                if (ctx->comm)
                    CK(cudaMemcpyAsync(ctx->bt_acc.p, acc_loc, T * acc_words * 8, cudaMemcpyDeviceToDevice, ctx->stream));
            } else {
                CK(cudaMemsetAsync(ctx->bt_acc.p, 0, T * acc_words * 8, ctx->stream));
                LAUNCH(launch_accumulate_batch(ctx->bt_feat.as<float>(), n_local, ctx->bt_assign.as<uint32_t>(), K, ctx->bt_acc.as<long long>(),
                                               tflags_d, ctl_d, ab, ctx->stream));
            }
            if (ctx->comm && ctx->nccl.all_reduce_sum_i64(ctx->comm, ctx->bt_acc.as<long long>(), T * acc_words, ctx->stream))
                return ctx->fail("ncclAllReduce(sum) failed");
            LAUNCH(launch_lloyd_control_batch(ctx->bt_acc.as<long long>(), acc_words, (uint32_t)T, K, iter, 0u, tflags_d, 4u, ctl_d,
                                              ctx->comm ? sticky_d : nullptr, ctx->stream));
            if (!ctx->comm)
                for (size_t t = 0; t < T; ++t)
                    LAUNCH(launch_exact_errors(feat_t(t), n_local, cent_t(t), assign_t(t), errors_t(t), ctl_d + 3 * t, ctl_d + 3 * t + 2, ctx->stream));
            LAUNCH(launch_finalize_centroids_batch(ctx->bt_acc.as<long long>(), acc_words, (uint32_t)T, K, inv_scales_d, ctx->bt_cent.as<float>(),
                                                   tflags_d, 4u, ctl_d, 3u, ctx->stream));
            if (!ctx->comm)
                for (size_t t = 0; t < T; ++t)
                    LAUNCH(launch_reseed_empty(feat_t(t), n_local, errors_t(t), acc_t(t), K, cent_t(t), tflags_d + 4 * t, ctl_d + 3 * t, ctx->stream));
            if ((iter == 3 || iter == 7 || iter == 15) && iter + 1 < cfg->max_iterations) {  // all converged / redo needed?
                CK(cudaMemcpyAsync(h_words, ctl_d, (7 * T + 1) * 4, cudaMemcpyDeviceToHost, ctx->stream));
                CK(cudaStreamSynchronize(ctx->stream));
                bool all_stopped = true;
                for (size_t t = 0; t < T; ++t) all_stopped = all_stopped && h_words[3 * t] != 0;
                if (all_stopped || h_words[7 * T]) break;
            }
        }
        CK(cudaMemcpyAsync(h_words, ctl_d, (7 * T + 1) * 4, cudaMemcpyDeviceToHost, ctx->stream));
        CK(cudaStreamSynchronize(ctx->stream));
        if (peer_xchg && xchg_check(ctx)) return 1;   // before anything is believed: a timed-out exchange left garbage behind
        if (h_words[7 * T]) {
            *redo = true;
            return 0;
        }
        for (size_t t = 0; t < T; ++t) {
            passes[t] = h_words[3 * t + 1];
            reseeds_async += h_words[3 * T + 4 * t + 2];
        }
        return 0;
    };
    bool redo = force_sync;
    if (!force_sync && (fused ? run_async_fused(&redo) : run_async(&redo))) return 1;
    if (redo && run_sync()) return 1;
    const int t3 = tm.mark();

    // ---- centroids -> pixels, final index map of this rank's shard (QuantizeBlocks, :447-471) -------------
    if (fused) {
        CK(ctx->cent_feat.ensure(T * (size_t)K * 256));
        CK(ctx->out_idx.ensure(T * nl * 4));
        LAUNCH(launch_oklab_to_pixels(ctx->bt_cent.as<float>(), T * (size_t)K, channels, ctx->bt_centpx.as<uint8_t>(), ctx->stream));
        LAUNCH(launch_pixels_to_features(ctx->bt_centpx.as<uint8_t>(), T * (size_t)K, channels, nullptr, 0, ctx->cent_feat.as<float>(), ctx->stream));
        CK(cudaMemsetAsync(flags + F_AMB, 0, 8, ctx->stream));
        if (assign_fused(ctx->cent_feat.as<float>(), ctx->out_idx.as<uint32_t>(), nullptr, nullptr)) return 1;
        for (size_t t = 0; t < T; ++t) {
            if (out_pixel_centroids && from_device(ctx, out_pixel_centroids[t], centpx_t(t), (size_t)K * 16 * channels)) return 1;
            if (out_indices && from_device(ctx, out_indices[t], ctx->out_idx.as<uint32_t>() + t * nl, n_local * 4)) return 1;
        }
    }
    CK(ctx->cent_feat.ensure((size_t)K * 256));
    CK(ctx->out_idx.ensure(nl * 4));
    for (size_t t = 0; t < T && !fused; ++t) {
        LAUNCH(launch_oklab_to_pixels(cent_t(t), K, channels, centpx_t(t), ctx->stream));
        LAUNCH(launch_pixels_to_features(centpx_t(t), K, channels, nullptr, 0, ctx->cent_feat.as<float>(), ctx->stream));
        if (assign_blocks(ctx, feat_t(t), n_local, ctx->cent_feat.as<float>(), K, cfg->assign_mode, ctx->out_idx.as<uint32_t>(),
                          nullptr, nullptr, nullptr, tensor ? &views[t] : nullptr))
            return 1;
        if (out_pixel_centroids && from_device(ctx, out_pixel_centroids[t], centpx_t(t), (size_t)K * 16 * channels)) return 1;
        if (out_indices && from_device(ctx, out_indices[t], ctx->out_idx.p, n_local * 4)) return 1;
    }
    const int t4 = tm.mark();
    CK(cudaMemcpyAsync(ctx->h_small, flags + F_RESEEDS, 4, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaMemcpyAsync(ctx->h_small + 1, flags + F_PAIRS, 8, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    if (stats) {
        stats->n_blocks = (uint32_t)(n_local * T);
        stats->codebook_size = K;
        stats->seed_rounds = seed_rounds;
        for (size_t t = 0; t < T; ++t) stats->lloyd_passes += passes[t];
        stats->reseeds = ctx->h_small[0] + ctx->reseeds_sharded + reseeds_async;
        stats->rerank_blocks = ctx->h_small[1];
        stats->rerank_overflow = ctx->h_small[2];
        stats->ms_seed = tm.ms(t0, t1);      // phase A (whole-texture decode + seeding of the owned textures)
        stats->ms_decode = tm.ms(t1, t2);    // phase B (shard decode, scale, tensor operands)
        stats->ms_lloyd = tm.ms(t2, t3);
        stats->ms_quantize = tm.ms(t3, t4);  // incl. centroid conversion and the D2H of the results
        stats->ms_total = tm.ms(t0, t4);
        stats->gpu_launches = ctx->launches;
    }
    return 0;
}

// HDR images (BASELINE config 4) over the ranks of the communicator: the general lock-step batch above.
int yupvq_encode_hdr_batch_sharded(yupvq_ctx* ctx, const float* const* images, size_t n_images, uint32_t width, uint32_t height,
                                   int channels, size_t first_block, size_t n_local, const yupvq_config* cfg,
                                   float* const* out_float_centroids, uint32_t* const* out_indices, yupvq_stats* stats) {
    if (!ctx) return 1;
    Scope scope(ctx->device);
    init_stats(stats);
    ctx->launches = 0;
    ctx->tc_feat = nullptr;
    ctx->maxabs_feat = nullptr;
    ctx->tc_rgb = rgb_layout(channels);
    const size_t n_total = (size_t)((width + 3) / 4) * ((height + 3) / 4);
    ctx->xchg_path = ctx->comm ? YUPVQ_XCHG_NCCL : YUPVQ_XCHG_NONE;
    if (check_cfg(ctx, cfg, n_total, channels)) return 1;
    if (!images || n_images == 0) return ctx->fail("no images");
    if (n_images > 4096) return ctx->fail("too many images in one batch (%zu > 4096)", n_images);
    if (first_block + n_local > n_total) return ctx->fail("shard [%zu, %zu) exceeds %zu blocks", first_block, first_block + n_local, n_total);
    BatchSrc src;
    src.hdr = true;
    src.data = reinterpret_cast<const void* const*>(images);
    src.width = width;
    src.height = height;
    return batch_sharded_general(ctx, src, n_images, n_total, first_block, n_local, channels, cfg,
                                 reinterpret_cast<void* const*>(out_float_centroids), out_indices, stats);
}

int yupvq_encode_hdr_sharded(yupvq_ctx* ctx, const float* image, uint32_t width, uint32_t height, int channels, size_t first_block,
                             size_t n_local, const yupvq_config* cfg, float* out_float_centroids, uint32_t* out_indices,
                             yupvq_stats* stats) {
    const float* img[1] = {image};
    float* cents[1] = {out_float_centroids};
    uint32_t* idx[1] = {out_indices};
    return yupvq_encode_hdr_batch_sharded(ctx, img, 1, width, height, channels, first_block, n_local, cfg, cents, idx, stats);
}

}  // extern "C"
