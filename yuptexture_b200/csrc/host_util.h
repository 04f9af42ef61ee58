// Host-side helpers shared by the kernel translation units: the parsed-once debug switches and a
// per-device "do this once" guard for cudaFuncSetAttribute (function attributes are per device).
#pragma once
#include <atomic>
#include <cuda_runtime.h>
#include <mutex>
#include <stdint.h>

namespace yupvq {

// A/B timing and test switches.  Read ONCE from the environment (the first call of debug_opts(), which
// yupvq_create makes); nothing in the per-pass code touches getenv.  None of them changes a result:
// every switch selects between two implementations that the tests prove bit-identical.
struct DebugOpts {
    bool seed_per_round = false;     // YUPVQ_SEED_PER_ROUND      one launch per K-Means++ round instead of the persistent kernel
    bool seed_tri = false;           // YUPVQ_SEED_TRI            LDR triangle-inequality skip (slower on noisy textures)
    bool seed_no_tri = false;        // YUPVQ_SEED_NO_TRI         HDR triangle-inequality skip off
    bool seed_no_pal = false;        // YUPVQ_SEED_NO_PAL         BC1 palette-form seeding off (pixel-plane kernel)
    bool seed_no_smem = false;       // YUPVQ_SEED_NO_SMEM        no shared-memory resident segments in the pixel-plane kernel
    bool lloyd_sync = false;         // YUPVQ_LLOYD_SYNC          host-synchronous Lloyd loop
    bool no_delta = false;           // YUPVQ_NO_DELTA            full accumulate on every pass
    bool no_rgb48 = false;           // YUPVQ_NO_RGB48            64-column tensor operands for textures without alpha
    bool no_split_decode = false;    // YUPVQ_NO_SPLIT_DECODE     one decode before the seeding
    bool no_decode_fuse = false;     // YUPVQ_NO_DECODE_FUSE      separate mean / max |x| / operand passes after the feature decode
    bool batch_per_texture = false;  // YUPVQ_BATCH_PER_TEXTURE   lock-step batch with per-texture launches
    bool no_tail_fuse = false;       // YUPVQ_NO_TAIL_FUSE        separate control / errors / finalize / reseed / prep launches
    bool tc_v1 = false;              // YUPVQ_TC_V1               tensor front-end with the one-barrier-per-tile epilogue (no merge warps)
    bool tc_force_batch = false;     // YUPVQ_TC_FORCE_BATCH      single texture through the batch instantiation of the front-end (codegen A/B)
    bool no_peer_xchg = false;       // YUPVQ_NO_PEER_XCHG        sharded batch: copy + ncclAllReduce instead of the in-kernel peer exchange
    int xchg_mode = 0;               // YUPVQ_XCHG_MODE           peer exchange: 1 = one shot, 2 = two shots (default: by size)
    int xchg_timeout_ms = 30000;     // YUPVQ_XCHG_TIMEOUT_MS     peer exchange: how long a CTA waits for a peer's counter before the call fails
    int xchg_desync_rank = -1;       // YUPVQ_XCHG_DESYNC_RANK    test hook: that rank skips one exchange number at pass 2 (its wait must time out)
    int async_dbg = 0;               // YUPVQ_ASYNC_DBG           bit 0: sync after every pass, bit 3: per-pass delta statistics on stderr
};
const DebugOpts& debug_opts();       // defined in api.cu

// Runs f() once per CUDA device (the current one); later calls on that device are a relaxed load.
// Safe against concurrent callers: the attribute is set before the bit becomes visible.
template <typename F>
inline cudaError_t once_per_device(std::atomic<uint64_t>& done, F&& f) {
    int dev = 0;
    cudaGetDevice(&dev);
    const uint64_t bit = 1ull << (dev & 63);
    if (done.load(std::memory_order_acquire) & bit) return cudaSuccess;
    static std::mutex mu;
    std::lock_guard<std::mutex> lock(mu);
    if (done.load(std::memory_order_relaxed) & bit) return cudaSuccess;
    const cudaError_t e = f();
    if (e == cudaSuccess) done.fetch_or(bit, std::memory_order_release);
    return e;
}

}  // namespace yupvq
