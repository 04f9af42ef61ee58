// Internal launch interface between the C-ABI/orchestration (api.cu) and the kernel translation units.
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

namespace yupvq {

// ---- convert.cu : K1 / K1h / K6 -------------------------------------------------
// Pixel blocks live in HBM as `ch` planes of uint4: plane p holds bytes [16p, 16p+16) of every
// block's 16*ch-byte pixel block, planes[p * plane_stride + block].  Seeding reads them coalesced.
cudaError_t colour_tables_init();
// DecodeFuse (optional, with feat != null): the same pass also writes the tensor front-end's per-block operands for the
// given mean row -- a_prime (n x 128 bf16), xnorm (n) -- and max-reduces |x| into *max_abs_bits (float bits; zero it
// first): the work of launch_tc_mean / launch_tc_prep_blocks / launch_max_abs without re-reading the features.
struct DecodeFuse {
    const float* mu = nullptr;        // 64 floats
    float* a_prime = nullptr;
    float* xnorm = nullptr;
    unsigned int* max_abs_bits = nullptr;
    int rgb = 0;
};
cudaError_t launch_decode_bcn(const uint8_t* bc, size_t n, int fmt, int ch, uint4* planes, size_t plane_stride,
                              float* feat, uint8_t* px_blocks, cudaStream_t s, const DecodeFuse* fuse = nullptr);
cudaError_t launch_pixels_to_features(const uint8_t* px_blocks, size_t n, int ch, uint4* planes, size_t plane_stride,
                                      float* feat, cudaStream_t s);
cudaError_t launch_floats_to_features(const float* float_blocks, size_t n, int ch, float* feat, cudaStream_t s);
cudaError_t launch_image_to_features(const float* image, uint32_t w, uint32_t h, int ch, float* feat, float* tiles,
                                     cudaStream_t s);
cudaError_t launch_oklab_to_pixels(const float* oklab, size_t n, int ch, uint8_t* out, cudaStream_t s);
cudaError_t launch_oklab_to_floats(const float* oklab, size_t n, int ch, float* out, cudaStream_t s);
cudaError_t launch_gather_rows_f32(const float* src, const uint32_t* rows, size_t n_rows, int row_floats, float* dst,
                                   cudaStream_t s);
cudaError_t launch_gather_blocks(const uint4* planes, size_t plane_stride, const float* feat, const uint32_t* rows,
                                 size_t n_rows, int ch, uint4* out_planes, size_t out_stride, float* out_feat,
                                 cudaStream_t s);
cudaError_t launch_planes_to_blocks(const uint4* planes, size_t plane_stride, const uint32_t* rows, size_t n_rows,
                                    int ch, uint8_t* out_px_blocks, cudaStream_t s);
cudaError_t launch_gather_codebook(const uint8_t* codebook, uint32_t n_entries, uint32_t block_size,
                                   const uint32_t* indices, size_t n, uint8_t* out, cudaStream_t s);

// ---- seed.cu : K2 / K2h ----------------------------------------------------------------------------
struct SeedState {
    float* min_dist;              // n
    unsigned long long* partial;  // grid partial sums (uint64 for LDR, double bits for HDR)
    unsigned int* counter;        // last-block-done ticket
    uint32_t* seed_idx;           // K chosen block indices (device)
    const double* canon;          // K-1 canonical uniforms in [0,1) (device)
    uint32_t* status;             // [0] degenerate_from (0 = none), [1] rounds drawn
    int grid;                     // CTAs per round (fixed for the whole seeding)
};
size_t seed_grid_for(size_t n);
size_t seed_partial_bytes(int grid);
cudaError_t launch_seed_init(const SeedState& st, size_t n, uint32_t first_idx, cudaStream_t s);
cudaError_t launch_seed_round_ldr(const SeedState& st, const uint4* planes, size_t plane_stride, size_t n, int ch,
                                  uint32_t round, uint32_t K, cudaStream_t s);
// all K-1 rounds in one cooperative launch; cudaErrorNotSupported -> use the per-round launches
// pal5 (optional, ch == 3 only): the blocks in BC1 palette form (launch_bc1_palette_planes), 5 planes of pal_stride words
cudaError_t launch_seed_persistent(const SeedState& st, const uint4* planes, size_t plane_stride, const float* feat, size_t n,
                                   int ch, bool hdr, uint32_t K, cudaStream_t s, const uint32_t* pal5 = nullptr,
                                   size_t pal_stride = 0);
// convert.cu: BC1 blocks -> 5 x uint32 palette planes (R, G, B palette bytes; selector nibbles of texels 0-7, 8-15)
cudaError_t launch_bc1_palette_planes(const uint8_t* bc, size_t n, uint32_t* pal5, size_t stride, cudaStream_t s);
cudaError_t launch_seed_round_hdr(const SeedState& st, const float* feat, size_t n, uint32_t round, uint32_t K,
                                  cudaStream_t s);

// ---- assign.cu : K3 --------------------------------------------------------------------------------
// Exact fp32 assignment in the reference's operation order (also the re-rank kernel).
// flags[0] |= 1 when any assignment changed.
cudaError_t launch_assign_exact(const float* feat, size_t n, const float* cent, uint32_t K, uint32_t* assign,
                                float* errors, uint32_t* flags, const uint32_t* stop, cudaStream_t s);
// errors[i] = exact distance to the assigned centroid (used before an empty-cluster reseed in tensor mode)
// runs only when *stop == 0 and *n_empty != 0 (either pointer may be null = unconditional)
cudaError_t launch_exact_errors(const float* feat, size_t n, const float* cent, const uint32_t* assign, float* errors,
                                const uint32_t* stop, const uint32_t* n_empty, cudaStream_t s);

// ---- assign_tc.cu : K3, tensor engine (tcgen05 front-end + exact re-rank) ------------------------------
// Per block set (once):  mu = column means, A' = rna_tf32(w (x - mu)), xnorm = |w (x - mu)|.
// Per pass: B' / folded norms / cmax from the current centroids, the MMA front-end writes one uint4
// candidate list per block (7 x uint16 + count; count 0xffff = full exact scan), the re-rank turns
// the lists into assignments with the reference's exact arithmetic.
// Lock-step batch descriptor: the rows of all textures form one matrix (texture t = rows [t * rows_per_tex, ...),
// rows_per_tex a multiple of 128), the per-texture centroid operands are stacked in the B buffers (k_pad rows
// each) and the per-texture scalar words sit at fixed strides.  Single texture: tc_single().
struct TcBatch {
    uint32_t tiles_per_tex;   // 128-row tiles per texture
    uint32_t k_pad;           // rows of b_main / b_ext per texture
    uint32_t cmax_stride;     // floats between the textures' cmax words
    uint32_t flags_stride;    // uint32 words between the textures' changed flags
    uint32_t stop_stride;     // uint32 words between the textures' stop words
    uint32_t rows_per_tex;    // rows per texture (the re-rank maps a row to its texture's centroids)
    uint32_t n_textures;
    // delta update (null = off): every row whose assignment changes is appended as
    // (row, previous centroid) so the update touches only those rows (launch_accumulate_delta)
    uint32_t* chg_rows;
    uint32_t* chg_old;
    unsigned int* chg_count;
    // 1 = the operands were prepared in the 48-column layout (see launch_tc_prep_blocks): 10 MMAs per tile-chunk
    uint32_t rgb;
};
inline TcBatch tc_single() { return TcBatch{0xffffffffu, 0u, 0u, 0u, 0u, 0xffffffffu, 1u, nullptr, nullptr, nullptr, 0u}; }
bool tc_supported(uint32_t K);
uint32_t tc_padded_k(uint32_t K);  // rows of b_main / b_ext
cudaError_t launch_tc_mean(const float* feat, size_t n, double* sums64, float* mu, cudaStream_t s);
// rgb != 0 (textures without an alpha channel: the alpha feature is the constant 1.0, so its 16 centred columns are
// exactly zero in every block): the 48 colour columns are packed first and the folded norm terms take three of the 16
// freed columns of the hi half, so a tile-chunk needs 3 + 3 + 4 = 10 K-steps instead of 13.  Both operands and the
// front-end launch must agree on the flag.
cudaError_t launch_tc_prep_blocks(const float* feat, size_t n, const float* mu, float* a_prime, float* xnorm,
                                  cudaStream_t s, int rgb = 0);
cudaError_t launch_tc_prep_centroids(const float* cent, uint32_t K, const float* mu, float* b_main, float* b_ext,
                                     float* cmax, cudaStream_t s, int rgb = 0);
// batch: texture t reads cent + t*K*64, mu at small + t*small_stride + mu_off, writes rows [t*k_pad, ..) of b_main /
// b_ext and max-reduces cmax at small + t*small_stride + cmax_off (zeroed here)
cudaError_t launch_tc_prep_centroids_batch(const float* cent, uint32_t K, uint32_t n_textures, uint8_t* small, size_t small_stride,
                                           size_t mu_off, size_t cmax_off, float* b_main, float* b_ext, cudaStream_t s, int rgb = 0);
// `stop` (may be null): device word; when non-zero the kernel returns at once (Lloyd passes are enqueued
// ahead of the convergence test, see launch_lloyd_control).
// assign != null: rows with exactly one candidate get their assignment (and the changed flag) written by the
// front-end, the other rows are appended to amb_list (counter amb_count) for launch_rerank_sparse.
// assign == null: every row's candidate list goes to `cand` for the dense launch_rerank (errors[] wanted).
cudaError_t launch_assign_tc(const float* a_prime, const float* xnorm, size_t n, const float* b_main,
                             const float* b_ext, uint32_t K, const float* cmax, uint4* cand, uint32_t* assign,
                             uint32_t* flags, uint32_t* amb_list, unsigned int* amb_count, const uint32_t* stop,
                             cudaStream_t s, const TcBatch* batch = nullptr);
// Optional epilogue of launch_rerank_sparse (fused Lloyd pass): the last CTA to finish does what
// launch_acc_prepare does -- decides full / delta for this pass' update, zeroes the accumulator(s) for a full pass and
// the changed word(s), hands the moved-row count over -- so the pass needs no separate launch for it.  dctl == null: off.
struct AccPrep {
    long long* acc;
    uint32_t* dctl;
    uint32_t K;
    uint32_t full_above;   // more moved rows than this -> full recompute
    uint32_t force_full;
    uint32_t n_textures;
    size_t acc_stride;
};
// batch != null: n is the row count of ALL textures, K the per-texture codebook size
cudaError_t launch_rerank_sparse(const float* feat, const float* cent, uint32_t K, const uint4* cand,
                                 const uint32_t* amb_list, unsigned int* amb_count, unsigned int* ticket,
                                 uint32_t* assign, uint32_t* flags, unsigned int* stat_pairs, cudaStream_t s,
                                 const TcBatch* batch = nullptr, const AccPrep* prep = nullptr);
cudaError_t launch_rerank(const float* feat, size_t n, const float* cent, uint32_t K, const uint4* cand,
                          uint32_t* assign, float* errors, uint32_t* flags, unsigned int* stat_pairs,
                          const uint32_t* stop, cudaStream_t s);

// ---- update.cu : K4 / K5 -----------------------------------------------------------------------------
// Accumulator layout (int64 words): [K*64 sums | K counts | 1 changed flag].  Sums are 64-bit fixed
// point (value * 2^shift): associative, therefore order / CTA / GPU-count independent; the whole
// buffer is what gets all-reduced when blocks are sharded across GPUs.
// flags: [0] changed (written by assign), [1] #empty clusters, [2] reseeds (running total).
cudaError_t launch_accumulate(const float* feat, size_t n, const uint32_t* assign, uint32_t K, int shift,
                              long long* acc, const uint32_t* flags, const uint32_t* stop, cudaStream_t s,
                              const uint32_t* dctl = nullptr);
// lock-step batch: texture t = rows [t * rows_per_tex, ..) of feat / assign, accumulator acc + t * acc_stride (int64 words),
// fixed-point scale scales[t] (device), changed flag flags[t * flags_stride], stop word stop[t * stop_stride]
struct AccBatch {
    uint32_t n_textures;
    size_t acc_stride;
    const float* scales;
    uint32_t flags_stride;
    uint32_t stop_stride;
};
// Delta update.  dctl (uint32 words, own cache lines): [0] changed-row counter the assignment kernels append to,
// [32] rows to process this pass, [33] mode (1 = full recompute, 0 = delta).
//   launch_acc_prepare     decides the mode for this pass (full on pass 0 / when more than n/4 rows changed / when the
//                          list is off), zeroes the accumulator for a full pass and always the changed word, moves the
//                          counter to [32] and clears it;
//   launch_accumulate      (full) runs only when dctl == null or dctl[33] == 1;
//   launch_accumulate_delta runs only when dctl[33] == 0: sums[new] += x, sums[old] -= x, counts likewise, for the
//                          listed rows -- exact, because the sums are integers.
cudaError_t launch_acc_prepare(long long* acc, uint32_t K, size_t n, uint32_t force_full, uint32_t* dctl, const uint32_t* stop,
                               cudaStream_t s);
cudaError_t launch_accumulate_delta(const float* feat, const uint32_t* assign, uint32_t K, int shift, long long* acc,
                                    const uint32_t* chg_rows, const uint32_t* chg_old, const uint32_t* dctl,
                                    const uint32_t* flags, const uint32_t* stop, cudaStream_t s,
                                    const AccBatch* batch = nullptr, uint32_t rows_per_tex = 0xffffffffu);
// full or delta update in ONE launch, the mode read from dctl[33] on the device (fused Lloyd pass); batch != null: the
// lock-step batch (scales from batch->scales, shift ignored)
bool accumulate_auto_supported(size_t rows_per_tex, uint32_t K, uint32_t n_textures);
cudaError_t launch_accumulate_auto(const float* feat, size_t rows_per_tex, const uint32_t* assign, uint32_t K, int shift, long long* acc,
                                   const uint32_t* chg_rows, const uint32_t* chg_old, const uint32_t* dctl, const uint32_t* flags,
                                   const uint32_t* stop, cudaStream_t s, const AccBatch* batch = nullptr);
// lock-step batch: one list / one mode for all textures (needs K * n_textures <= 4096, else every pass is full)
bool delta_batch_supported(uint32_t K, uint32_t n_textures);
cudaError_t launch_acc_prepare_batch(long long* acc, uint32_t K, size_t total_rows, uint32_t force_full, uint32_t* dctl,
                                     const AccBatch& bt, cudaStream_t s);
bool accumulate_batch_supported(size_t rows_per_tex, uint32_t K);
cudaError_t launch_accumulate_batch(const float* feat, size_t rows_per_tex, const uint32_t* assign, uint32_t K, long long* acc,
                                    const uint32_t* flags, const uint32_t* stop, const AccBatch& bt, cudaStream_t s,
                                    const uint32_t* dctl = nullptr);
cudaError_t launch_finalize_centroids_batch(const long long* acc, size_t acc_stride, uint32_t n_textures, uint32_t K,
                                            const double* inv_scales, float* cent, uint32_t* flags, uint32_t flags_stride,
                                            const uint32_t* stop, uint32_t stop_stride, cudaStream_t s);
// centroid = float(sum) * (1.0f / count) (reference src/vq_encoder.cpp:366-368); flags[1] = #empty clusters
cudaError_t launch_finalize_centroids(const long long* acc, uint32_t K, int shift, float* cent, uint32_t* flags,
                                      const uint32_t* stop, cudaStream_t s);
cudaError_t launch_count_empty(const long long* acc, uint32_t K, uint32_t* out, cudaStream_t s);
// The between-assign-and-update test of the reference's Lloyd loop (src/vq_encoder.cpp:341 / :572), on the
// device so that passes can be enqueued without a host round trip: reads the (all-reduced) changed word
// acc[K*65]; if nothing changed and iter > break_after sets ctl[0] (stop) else ++ctl[1] (passes run);
// writes the number of empty clusters to ctl[2] and clears flags[0] (changed) and flags[1] (#empty).
cudaError_t launch_lloyd_control(const long long* acc, uint32_t K, uint32_t iter, uint32_t break_after, uint32_t* flags,
                                 uint32_t* ctl, cudaStream_t s);
// the same for texture blockIdx.x of a lock-step batch: acc + t * acc_stride, flags + t * flags_stride, ctl + 3 t;
// *sticky = 1 (if non-null) when a running texture has an empty cluster
cudaError_t launch_lloyd_control_batch(const long long* acc, size_t acc_stride, uint32_t n_textures, uint32_t K, uint32_t iter,
                                       uint32_t break_after, uint32_t* flags, uint32_t flags_stride, uint32_t* ctl,
                                       uint32_t* sticky, cudaStream_t s);
// serial-semantics reseed of empty clusters (ascending c, first maximum of errors, then zero it)
cudaError_t launch_reseed_empty(const float* feat, size_t n, float* errors, const long long* acc, uint32_t K,
                                float* cent, uint32_t* flags, const uint32_t* stop, cudaStream_t s);
// out[0] = max error (float bits), out[1] = index of its first occurrence
cudaError_t launch_argmax_error(const float* errors, size_t n, uint32_t* out2, cudaStream_t s);
cudaError_t launch_max_abs(const float* x, size_t n, uint32_t* out_bits, cudaStream_t s);

// ---- tail.cu : the fused tail of a Lloyd pass (control + errors-if-needed + finalize + reseed + next pass' operands) ------
// Strides address texture t of a lock-step batch (n_textures == 1: all strides unused).  cmax is double-buffered by pass
// parity: the front-end of pass p reads small[cmax_off + 4 * (p & 1)], this kernel max-reduces the next pass' value into slot
// (iter + 1) & 1 and zeroes slot iter & 1 for the tail after it.  errors == null (or n_textures > 1): no device reseed;
// `sticky` (may be null) is raised when a running texture has an empty cluster.
// Exchange step of a block-sharded pass as part of the tail kernel: a two-shot all-reduce of the int64
// partial sums over NVLink peer memory instead of copy + ncclAllReduce.  Every rank pushes slice j of its partial sums into
// rank j's staging area, rank j adds the R contributions (integer adds: any order gives the same bits) and pushes the
// totals of its slice into every rank's result buffer; monotonic counters (one per phase and source rank, bumped with
// red.release.sys by every CTA, polled with ld.acquire.sys) order the two phases -- no grid-wide barrier, no host round trip.
// The data buffers alternate with the parity of a per-context epoch.  n_ranks == 0: off.
constexpr int kXchgMaxRanks = 16;
struct XchgArgs {
    int n_ranks = 0, rank = 0;
    int one_shot = 0;                         // 1: every rank receives every peer's whole vector and adds on read (small vectors)
    uint32_t epoch = 0;                       // number of this exchange (1, 2, ...): counters reach epoch * grid size; parity = epoch & 1
    size_t words = 0;                         // int64 words exchanged
    size_t slice = 0;                         // words per owner slice (ceil(words / n_ranks))
    size_t slot_stride = 0;                   // words between the per-source-rank slots of a staging area
    const long long* src = nullptr;           // this rank's partial sums
    long long* peer_stage[kXchgMaxRanks];     // rank j's staging area of this parity (n_ranks slots), mapped into this process
    long long* peer_result[kXchgMaxRanks];    // rank j's result buffer of this parity
    uint32_t* peer_flags[kXchgMaxRanks];      // rank j's counters: [2 phases][kXchgMaxRanks sources]
    // Bounded wait: a counter that has not reached its target timeout_ns after a CTA began to spin on it (a peer died, failed
    // on the host or launched a different sequence) raises *err (this rank's own memory, sticky) and the kernel goes on
    // with whatever the buffers hold -- every access stays in bounds -- so that it ends and the host can fail the call.  Once
    // *err is set the exchange is skipped altogether.  err == null: wait for ever.
    uint32_t* err = nullptr;
    unsigned long long timeout_ns = 0;
};
// word offset of the error word inside a rank's counter area (behind the [2][kXchgMaxRanks] counters)
constexpr int kXchgErrWord = 2 * kXchgMaxRanks;
struct TailArgs {
    XchgArgs x;                // n_ranks > 1: `acc` is this rank's result buffer, filled by the exchange at the head of the kernel
    const long long* acc;      // (all-reduced) accumulators
    size_t acc_stride;         // int64 words between textures
    uint32_t n_textures, K, k_pad, iter, break_after;
    uint32_t* flags;           // [changed, empty, reseeds, ...] per texture
    uint32_t flags_stride;
    uint32_t* ctl;             // [stop, passes, empty clusters] per texture (stride 3)
    uint32_t* sticky;
    double inv_scale;          // 2^-shift (single texture)
    const double* inv_scales;  // per texture (batch), else null
    float* cent;               // n_textures x K x 64
    uint8_t* small;            // per-texture scalar block: mu (64 floats) at mu_off, cmax[2] at cmax_off
    size_t small_stride, mu_off, cmax_off;
    float* b_main;             // n_textures x k_pad operand rows
    float* b_ext;
    int rgb;
    // single-texture reseed path
    const float* feat;
    size_t n_rows;
    const uint32_t* assign;
    float* errors;
    float* scratch_val;        // one word per CTA
    unsigned long long* scratch_idx;
};
cudaError_t launch_lloyd_tail(const TailArgs& args, cudaStream_t s);

// ---- bcn_encode.cu : N4, whole-texture BCn encoders (BC1, BC3, BC4, BC5) ---------------------------------------
// image: height x width x channels bytes, row-major; out: ceil(w/4) * ceil(h/4) blocks in row-major block order.
// swap_rb: encode a red/blue-swapped view (3- and 4-channel sources); alpha_threshold: BC1 punch-through (4 channels).
bool bcn_encode_supported(int fmt);
cudaError_t launch_bcn_encode(const uint8_t* image, uint32_t width, uint32_t height, uint32_t channels, int fmt, int swap_rb,
                              uint32_t alpha_threshold, uint8_t* out, cudaStream_t s);

// ---- sad.cu : A21, SAD-metric Lloyd (all integer, bit-exact) ---------------------------------------------
// argmin over K pixel-block centroids (K x 16*ch bytes) of sum |a - b|; errors[i] = the minimum (as float)
cudaError_t launch_sad_assign(const uint4* planes, size_t stride, size_t n, int ch, const uint8_t* cent_px, uint32_t K,
                              uint32_t* assign, float* errors, uint32_t* flags, const uint32_t* stop, cudaStream_t s);
// pixel planes -> N x 64 fp32 rows (integer-valued, zero padded): the rows the shared accumulate / reseed kernels read
cudaError_t launch_sad_expand(const uint4* planes, size_t stride, size_t n, int ch, float* rows, cudaStream_t s);
cudaError_t launch_sad_pack(const float* cent_f, uint32_t K, int ch, uint8_t* cent_px, cudaStream_t s);
// centroid byte = sum / count (truncating); flags[1] = #empty clusters
cudaError_t launch_sad_finalize(const long long* acc, uint32_t K, float* cent_f, uint32_t* flags, const uint32_t* stop,
                                cudaStream_t s);

}  // namespace yupvq
