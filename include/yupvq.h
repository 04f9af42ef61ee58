/*
 * yupvq.h -- C ABI of the B200-native YupTexture vector-quantisation hot path.
 *
 * This is the drop-in boundary: plain pointers and sizes, no C++ or torch types.
 * Every entry point names the reference interface it replaces (paths relative to the reference
 * repository, YupCore/YupTexture @ 7c8d4366).  The C++ classes in include/vq_encoder.h and
 * include/vq_bcn_compressor.h (same names/signatures as the reference's) are thin wrappers over
 * these calls; INTEGRATION.md shows the binding a reference maintainer would add.
 *
 * Conventions
 *   - All functions return 0 on success, non-zero on failure; yupvq_last_error(ctx) gives the text
 *     (the C++ wrapper rethrows it as std::runtime_error, the reference's convention,
 *     src/vq_bcn_compressor.cpp:99).
 *   - Data pointers may be host OR device pointers (detected with cudaPointerGetAttributes);
 *     host buffers are staged through pinned memory owned by the context.  Outputs are caller-owned.
 *   - A context is bound to one CUDA device and is not safe for concurrent calls (same contract as
 *     one VQBCnCompressor instance, include/vq_bcn_compressor.h:16-25).
 *   - There is no CPU fallback: without a CUDA device yupvq_create fails.
 *   - Block order is row-major over 4x4 tiles, pixel order row-major inside a tile, channels
 *     interleaved -- the layout of src/vq_bcn_compressor.cpp:142-146 / :235-252.
 */
#ifndef YUPVQ_H
#define YUPVQ_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define YUPVQ_API __attribute__((visibility("default")))
#else
#define YUPVQ_API
#endif

typedef struct yupvq_ctx yupvq_ctx;

/* BCFormat values of include/vq_bcn_types.h:27-36. */
enum { YUPVQ_BC1 = 1, YUPVQ_BC2 = 2, YUPVQ_BC3 = 3, YUPVQ_BC4 = 4, YUPVQ_BC5 = 5, YUPVQ_BC6H = 6, YUPVQ_BC7 = 7 };
/* DistanceMetric values of include/vq_bcn_types.h:105-108. */
enum { YUPVQ_METRIC_SAD = 0, YUPVQ_METRIC_OKLAB = 1 };
/* Assignment engine: AUTO picks the tcgen05 front-end + exact re-rank when the shape allows it. */
enum { YUPVQ_ASSIGN_AUTO = 0, YUPVQ_ASSIGN_EXACT = 1, YUPVQ_ASSIGN_TENSOR = 2 };

/* Mirrors CompressionConfig (include/vq_encoder.h:22-41) + the seed the reference lacks
 * (src/vq_encoder.cpp:24-26 seeds std::mt19937 from std::random_device). */
typedef struct yupvq_config {
    uint32_t codebook_size;   /* K; use yupvq_codebook_size(quality) for the reference mapping      */
    uint32_t max_iterations;  /* CompressionConfig::maxIterations (32)                              */
    int32_t  metric;          /* YUPVQ_METRIC_*  (HDR ignores it: always Oklab)                     */
    float    sample_ratio;    /* CompressionConfig::fastModeSampleRatio; >=1 or <=0 -> all blocks   */
    uint64_t seed;            /* std::mt19937 seed; 0 -> std::random_device (reference behaviour)   */
    int32_t  assign_mode;     /* YUPVQ_ASSIGN_*                                                     */
    int32_t  reserved;
} yupvq_config;

typedef struct yupvq_stats {
    uint32_t n_blocks;        /* blocks clustered (after sampling)                                  */
    uint32_t codebook_size;
    uint32_t lloyd_passes;    /* assignment passes run inside BuildCodebook                         */
    uint32_t seed_rounds;     /* K-Means++ rounds that drew a sample                                */
    uint32_t reseeds;         /* empty clusters re-seeded over all passes                           */
    uint32_t rerank_blocks;   /* exact distance evaluations of the re-rank over all passes (tensor)  */
    float ms_h2d, ms_decode, ms_seed, ms_lloyd, ms_finalize, ms_quantize, ms_d2h, ms_total;
    uint32_t gpu_launches;    /* kernels launched by this call                                      */
    uint32_t rerank_overflow; /* blocks that needed a full exact scan (empty/overfull candidate list) */
} yupvq_stats;

/* ---- lifetime ------------------------------------------------------------------------------ */
YUPVQ_API int yupvq_create(int device, yupvq_ctx** out);
YUPVQ_API void yupvq_destroy(yupvq_ctx* ctx);
YUPVQ_API const char* yupvq_last_error(const yupvq_ctx* ctx); /* ctx may be NULL: creation error */
YUPVQ_API const char* yupvq_version(void);

/* Device-side stopwatch on the context's stream (CUDA events): brackets any sequence of calls.
 * The reference wraps its calls in std::chrono (example/example.cpp:293-304); this is the
 * device-clock equivalent the bench uses. */
YUPVQ_API int yupvq_timer_start(yupvq_ctx* ctx);
YUPVQ_API int yupvq_timer_stop(yupvq_ctx* ctx, float* out_ms);

/* Per-launch timing of the dominant kernel (the tcgen05 assignment front-end) inside real calls:
 * while enabled, every launch of it in the Lloyd / quantise passes is bracketed by a CUDA event pair
 * on the context's stream (up to 128 launches); yupvq_profile_read synchronises, returns the summed
 * time and the number of launches, and rearms.  Used by bench.py for the roofline line only (the
 * timed region of the bench runs with profiling off). */
YUPVQ_API int yupvq_profile(yupvq_ctx* ctx, int enable);
YUPVQ_API int yupvq_profile_read(yupvq_ctx* ctx, float* out_ms_front_end, uint32_t* out_launches);
/* Stage times of the Lloyd passes run while profiling was on (CUDA events between the launches of a pass), summed over the
 * passes: out_ms = { tensor front-end, exact re-rank, update (full or delta), tail (with the exchange step when the
 * blocks are sharded) }.  Used by bench.py to explain where a multi-GPU step spends its time. */
YUPVQ_API int yupvq_profile_stages(yupvq_ctx* ctx, float out_ms[4], uint32_t* out_passes);

/* CompressionConfig::SetQuality, src/vq_encoder.cpp:12-18 (powers hard-wired 4..10). */
YUPVQ_API uint32_t yupvq_codebook_size(float quality);
/* Fill cfg the way VQBCnCompressor::Compress does from CompressionParams,
 * src/vq_bcn_compressor.cpp:128-134. */
YUPVQ_API void yupvq_config_init(yupvq_config* cfg, float quality, uint32_t max_iterations, int metric,
                                 float sample_ratio, uint64_t seed);

/* ---- whole-path entry points ----------------------------------------------------------------- */

/* LDR VQ from BCn bytes: the three starred lines of VQBCnCompressor::Compress
 * (src/vq_bcn_compressor.cpp:142-150): per-block BCn decode, BuildCodebook (up to, not including,
 * the Compressonator centroid encode of src/vq_encoder.cpp:434-439) and QuantizeBlocks.
 *   bc_blocks           n_blocks * blockSize bytes (8 for BC1/BC4, else 16)
 *   out_pixel_centroids K * 16*channels bytes   (outPixelCentroids of BuildCodebook)
 *   out_indices         n_blocks uint32         (QuantizeBlocks result)                          */
YUPVQ_API int yupvq_encode_bcn(yupvq_ctx* ctx, const uint8_t* bc_blocks, size_t n_blocks, int bc_format,
                               int channels, const yupvq_config* cfg, uint8_t* out_pixel_centroids,
                               uint32_t* out_indices, yupvq_stats* stats);

/* HDR VQ from the raw float image: VQBCnCompressor::CompressHDR, src/vq_bcn_compressor.cpp:230-256
 * (edge-clamped 4x4 tile gather, BuildCodebookHDR up to the BC6H encode, QuantizeBlocksHDR).
 *   image               height*width*channels floats, row-major
 *   out_float_centroids K * 16*channels floats                                                    */
YUPVQ_API int yupvq_encode_hdr(yupvq_ctx* ctx, const float* image, uint32_t width, uint32_t height,
                               int channels, const yupvq_config* cfg, float* out_float_centroids,
                               uint32_t* out_indices, yupvq_stats* stats);

/* Batch of independent same-shape BC textures run in lock-step (BASELINE config 5).  Each texture
 * gets its own codebook; seeds are cfg->seed + t.  Pointers are arrays of n_textures pointers.  */
YUPVQ_API int yupvq_encode_bcn_batch(yupvq_ctx* ctx, const uint8_t* const* bc_blocks, size_t n_textures,
                                     size_t n_blocks, int bc_format, int channels, const yupvq_config* cfg,
                                     uint8_t* const* out_pixel_centroids, uint32_t* const* out_indices,
                                     yupvq_stats* stats);

/* ---- the reference's four public VQEncoder calls, on decoded blocks --------------------------- */

/* VQEncoder::BuildCodebook, src/vq_encoder.cpp:260-433 (pixel blocks already decoded). */
YUPVQ_API int yupvq_build_codebook(yupvq_ctx* ctx, const uint8_t* pixel_blocks, size_t n_blocks, int channels,
                                   const yupvq_config* cfg, uint8_t* out_pixel_centroids, yupvq_stats* stats);
/* VQEncoder::QuantizeBlocks, src/vq_encoder.cpp:447-486. */
YUPVQ_API int yupvq_quantize_blocks(yupvq_ctx* ctx, const uint8_t* pixel_blocks, size_t n_blocks, int channels,
                                    const uint8_t* pixel_centroids, uint32_t codebook_size, int metric,
                                    int assign_mode, uint32_t* out_indices, yupvq_stats* stats);
/* VQEncoder::BuildCodebookHDR, src/vq_encoder.cpp:493-613. */
YUPVQ_API int yupvq_build_codebook_hdr(yupvq_ctx* ctx, const float* float_blocks, size_t n_blocks, int channels,
                                       const yupvq_config* cfg, float* out_float_centroids, yupvq_stats* stats);
/* VQEncoder::QuantizeBlocksHDR, src/vq_encoder.cpp:629-658. */
YUPVQ_API int yupvq_quantize_blocks_hdr(yupvq_ctx* ctx, const float* float_blocks, size_t n_blocks, int channels,
                                        const float* float_centroids, uint32_t codebook_size, int assign_mode,
                                        uint32_t* out_indices, yupvq_stats* stats);

/* ---- kernel-level entry points (unit parity; each is one stage of the path) ------------------- */

/* BCnCompressor::Decompress of one 4x4 block, src/bcn_compressor.cpp:191-227, for n blocks. */
YUPVQ_API int yupvq_decode_bcn(yupvq_ctx* ctx, const uint8_t* bc_blocks, size_t n_blocks, int bc_format,
                               int channels, uint8_t* out_pixel_blocks);
/* VQEncoder::PixelBlockToOklabBlock, src/vq_encoder.cpp:96-102 (out: n * 64 floats). */
YUPVQ_API int yupvq_pixels_to_oklab(yupvq_ctx* ctx, const uint8_t* pixel_blocks, size_t n_blocks, int channels,
                                    float* out_oklab);
/* VQEncoder::OklabBlockToPixelBlock, src/vq_encoder.cpp:105-111. */
YUPVQ_API int yupvq_oklab_to_pixels(yupvq_ctx* ctx, const float* oklab, size_t n_blocks, int channels,
                                    uint8_t* out_pixel_blocks);
/* VQEncoder::RgbaFloatBlockToOklabBlock, src/vq_encoder.cpp:179-185. */
YUPVQ_API int yupvq_floats_to_oklab(yupvq_ctx* ctx, const float* float_blocks, size_t n_blocks, int channels,
                                    float* out_oklab);
/* VQEncoder::OklabBlockToRgbaFloatBlock, src/vq_encoder.cpp:188-194. */
YUPVQ_API int yupvq_oklab_to_floats(yupvq_ctx* ctx, const float* oklab, size_t n_blocks, int channels,
                                    float* out_float_blocks);
/* The 4x4 tile gather of src/vq_bcn_compressor.cpp:235-252 (out: n_blocks * 16*channels floats). */
YUPVQ_API int yupvq_gather_tiles(yupvq_ctx* ctx, const float* image, uint32_t width, uint32_t height,
                                 int channels, float* out_float_blocks);
/* The assignment loop of src/vq_encoder.cpp:330-340 on Oklab features: argmin with strict '<'
 * (lowest index wins) and errors[i] = min distance (OklabBlockDistanceSq_SIMD, :212-237).
 * out_errors may be NULL. */
YUPVQ_API int yupvq_assign(yupvq_ctx* ctx, const float* features, size_t n_blocks, const float* centroids,
                           uint32_t codebook_size, int assign_mode, uint32_t* out_indices, float* out_errors,
                           yupvq_stats* stats);
/* K-Means++ seeding of src/vq_encoder.cpp:287-313 (LDR, SAD^2 on pixels): out_seed_blocks[k] is the
 * index of the block chosen as seed k.  Returns through *out_rounds the rounds that drew a sample. */
YUPVQ_API int yupvq_seed_ldr(yupvq_ctx* ctx, const uint8_t* pixel_blocks, size_t n_blocks, int channels,
                             uint32_t codebook_size, uint64_t seed, uint32_t* out_seed_blocks, uint32_t* out_rounds);
/* K-Means++ seeding of src/vq_encoder.cpp:526-554 (HDR, weighted L2 on Oklab features). */
YUPVQ_API int yupvq_seed_hdr(yupvq_ctx* ctx, const float* features, size_t n_blocks, uint32_t codebook_size,
                             uint64_t seed, uint32_t* out_seed_blocks, uint32_t* out_rounds);
/* One Lloyd update of src/vq_encoder.cpp:343-375: sums/counts, c = sum * (1.0f / count), empty
 * clusters re-seeded with serial semantics from `errors` (modified in place on the device copy). */
YUPVQ_API int yupvq_update_centroids(yupvq_ctx* ctx, const float* features, size_t n_blocks,
                                     const uint32_t* assignments, const float* errors, uint32_t codebook_size,
                                     float* inout_centroids, uint32_t* out_counts);
/* VQBCnCompressor::DecompressToBCn gather, src/vq_bcn_compressor.cpp:350-358:
 * out[i] = codebook[indices[i]] (blocks with an out-of-range index are left zero, as the reference
 * leaves them default-initialised). */
YUPVQ_API int yupvq_gather_codebook(yupvq_ctx* ctx, const uint8_t* codebook, uint32_t n_entries, uint32_t block_size,
                                    const uint32_t* indices, size_t n_blocks, uint8_t* out_bc_blocks);

/* The same update for callers that cannot allocate GPU memory.  No counterpart in the reference (its update works in
 * std::vector on the host); this exists because every other entry point needs a yupvq_ctx, and creating or growing a
 * context calls cudaMalloc / cudaMallocHost.  That rules the library out for a pipeline embedded in an engine or tool
 * whose own allocator already owns the whole device, for code in a frame loop that cannot afford the process-wide stall
 * of a driver allocation, and for plugins whose host application hands them buffers and forbids them device memory of
 * their own.  This call takes no context and allocates nothing: its working memory (fixed-point sums, counts and a few
 * flag words) is an array built into the library, 520 KiB, which every GPU that loads libyupvq.so carries until the
 * process exits.
 *   - every buffer is DEVICE (or managed) memory owned by the caller, valid on `device`; host pointers are refused
 *     (staging them would need an allocation).  cuda_stream is a cudaStream_t of that device (NULL = default stream).
 *   - codebook_size <= YUPVQ_STATIC_MAX_CODEBOOK (what the array was sized for); larger values are an error.
 *     assignments[i] must be < codebook_size, as for yupvq_update_centroids.
 *   - inout_errors is modified IN PLACE by the empty-cluster reseed (the error of each block taken as a new centroid
 *     becomes 0, src/vq_encoder.cpp:373); yupvq_update_centroids works on a private copy instead.
 *   - out_counts (may be NULL): codebook_size uint32 in device memory.
 *   - centroids and counts are bit-identical to yupvq_update_centroids on the same inputs.
 *   - all calls on a device share the one array.  It keeps the previous call's contents, so each call first clears the
 *     words it uses; two calls whose GPU work overlaps (e.g. two host threads) overwrite each other's sums -- serialise
 *     them.  The call returns after its work on cuda_stream has finished.
 *   - the first call on each device also resolves the array's device address and sets a kernel attribute; make that
 *     first call outside any stream capture.
 *   - returns 0, or non-zero with a message in err_text (caller's buffer of err_len bytes; may be NULL). */
#define YUPVQ_STATIC_MAX_CODEBOOK 1024
YUPVQ_API int yupvq_update_centroids_static(int device, void* cuda_stream, const float* features, size_t n_blocks,
                                            const uint32_t* assignments, float* inout_errors, uint32_t codebook_size,
                                            float* inout_centroids, uint32_t* out_counts, char* err_text, size_t err_len);

/* ---- N4: the whole-texture BCn pre-pass on the GPU ---------------------------------------------- */

/* BCnCompressor::Compress, src/bcn_compressor.cpp:45-116 (called at src/vq_bcn_compressor.cpp:94), for BC1, BC3, BC4
 * and BC5: image (height x width x channels bytes, row-major, host or device) -> ceil(w/4) * ceil(h/4) blocks in
 * row-major block order (host or device).  The reference hands this stage to the closed-source Compressonator
 * encoders; the blocks written here are not bit-identical to its output but decode to at least its PSNR
 * (tests/test_gpu_bcn_encode.py).  swap_rb != 0 encodes a red/blue-swapped view of a 3- or 4-channel source:
 * Compressonator reads CMP_FORMAT_RGB_888 sources that way, so the reference's VQ path (flipRGB == false) corresponds
 * to swap_rb = (channels == 3) and its bypass path (flipRGB == true, :61-76) to swap_rb = 0.  alpha_threshold: BC1 with
 * a 4-channel source puts texels with alpha below it on the transparent entry (CMP_CompressOptions::nAlphaThreshold,
 * :103-107).  out_ms_kernel (may be NULL): CUDA-event time of the encode kernel.  A device out_bc_blocks aligned to the
 * block size (8 or 16 bytes) is written in place; any other pointer (host memory, or a misaligned slice of a device
 * buffer) is filled through a staging copy. */
YUPVQ_API int yupvq_bcn_encode_supported(int bc_format);
YUPVQ_API int yupvq_bcn_encode_image(yupvq_ctx* ctx, const uint8_t* image, uint32_t width, uint32_t height, int channels,
                                     int bc_format, int swap_rb, int alpha_threshold, uint8_t* out_bc_blocks,
                                     float* out_ms_kernel);
/* Pre-pass + VQ in one call (VQBCnCompressor::Compress, src/vq_bcn_compressor.cpp:94-150, without the host-side
 * centroid encodes and container): the BCn blocks stay in HBM between the encoder and yupvq_encode_bcn's decode.
 * out_bc_blocks may be NULL (the VQ path does not need them on the host). */
YUPVQ_API int yupvq_compress_image(yupvq_ctx* ctx, const uint8_t* image, uint32_t width, uint32_t height, int channels,
                                   int bc_format, int swap_rb, int alpha_threshold, const yupvq_config* cfg,
                                   uint8_t* out_bc_blocks, uint8_t* out_pixel_centroids, uint32_t* out_indices,
                                   yupvq_stats* stats, float* out_ms_bcn_encode);

/* ---- multi-GPU (one process per GPU) -------------------------------------------- */

/* NCCL bootstrap: rank 0 calls yupvq_comm_unique_id, the host runtime (torch.distributed, MPI, a
 * file) hands the 128 bytes to every rank, then every rank calls yupvq_comm_init. */
YUPVQ_API int yupvq_comm_unique_id(yupvq_ctx* ctx, uint8_t out_id[128]);
YUPVQ_API int yupvq_comm_init(yupvq_ctx* ctx, const uint8_t id[128], int n_ranks, int rank);
/* Only the *_sharded entry points below use the communicator.  Every other call on a context that has one
 * (yupvq_encode_bcn, yupvq_build_codebook[_hdr], yupvq_encode_bcn_batch, ...) stays rank-local: ranks may encode
 * unrelated textures with them at the same time. */
/* How the last *_sharded call on this context exchanged the per-pass partial sums (no reference equivalent: the reference
 * has one encoder per Compress call, src/vq_bcn_compressor.cpp:136).  PEER_*: inside the pass-tail kernel over NVLink peer
 * memory (one shot = every rank reads all vectors, two shots = reduce-scatter then all-gather); NCCL: copy +
 * ncclAllReduce (peers not mappable, grids differ, batch beyond the fused path, or after a timed-out peer exchange);
 * NONE: no communicator or no sharded call yet. */
enum { YUPVQ_XCHG_NONE = 0, YUPVQ_XCHG_PEER_ONE_SHOT = 1, YUPVQ_XCHG_PEER_TWO_SHOT = 2, YUPVQ_XCHG_NCCL = 3 };
YUPVQ_API int yupvq_exchange_path(const yupvq_ctx* ctx);
/* Block-sharded LDR VQ of ONE texture: this rank holds blocks [first_block, first_block + n_local) of a texture
 * of n_total blocks; the whole texture's BCn bytes are passed on every rank.  It is the lock-step batch below with a
 * single texture: rank 0 runs the K-Means++ seeding (with cfg->seed == 0 it alone draws from std::random_device) and
 * the seed rows go out with one all-reduce; every rank then runs the Lloyd passes on its shard without a host round
 * trip per pass, and per pass the K x 64 fixed-point sums, K counts and the changed flag are all-reduced over NCCL.
 * out_indices receives this rank's n_local indices. */
YUPVQ_API int yupvq_encode_bcn_sharded(yupvq_ctx* ctx, const uint8_t* bc_blocks, size_t n_total, size_t first_block,
                                       size_t n_local, int bc_format, int channels, const yupvq_config* cfg,
                                       uint8_t* out_pixel_centroids, uint32_t* out_indices, yupvq_stats* stats);

/* Lock-step batch over block shards (BASELINE config 5): n_textures textures of n_total
 * blocks each, this rank owns blocks [first_block, first_block + n_local) of EVERY texture.
 * Texture t is seeded on rank t % n_ranks (the K seed rows are combined
 * by one all-reduce), all textures run their Lloyd passes together and ONE all-reduce per pass carries
 * the K x 65 fixed-point sums/counts of the whole batch.  Texture t uses seed cfg->seed + t, as
 * yupvq_encode_bcn_batch does; results are identical to the single-GPU call.  Without a
 * communicator (yupvq_comm_init not called) it is the single-GPU lock-step batch (n_local = n_total).
 * out_pixel_centroids[t]: K x 16*channels bytes; out_indices[t]: n_local indices.
 * shard_only (may be NULL = every pointer is a whole texture): shard_only[t] != 0 says bc_blocks[t] points at this
 * rank's n_local blocks only.  A rank needs the whole texture just for the textures it seeds (t % n_ranks == rank),
 * so with the flags set its input (and H2D traffic) is n_textures / n_ranks whole textures plus one shard of each of
 * the others, instead of growing with n_textures. */
YUPVQ_API int yupvq_encode_bcn_batch_sharded(yupvq_ctx* ctx, const uint8_t* const* bc_blocks, size_t n_textures,
                                             size_t n_total, size_t first_block, size_t n_local, int bc_format,
                                             int channels, const yupvq_config* cfg,
                                             uint8_t* const* out_pixel_centroids, uint32_t* const* out_indices,
                                             yupvq_stats* stats, const uint8_t* shard_only);
/* Variants on the two entry points above (src/vq_encoder.cpp:262-284, :380-430): with cfg->sample_ratio in (0, 1)
 * (fast mode) the ranks split each texture's SAMPLE list instead of the texture -- the list is drawn on the seeding
 * rank and broadcast -- and with cfg->metric == YUPVQ_METRIC_SAD the integer SAD kernels run; both need the whole
 * texture on every rank (shard_only must be NULL or all zero).  Results equal yupvq_encode_bcn with seed + t. */

/* HDR images over the ranks (BASELINE config 4; src/vq_encoder.cpp:493-658): n_images float images of width x height
 * x channels, every rank passes every image whole and receives the index map of its block range
 * [first_block, first_block + n_local) of each (blocks in row-major 4x4-tile order, as yupvq_encode_hdr).
 * Image t is seeded on rank t % n_ranks with seed cfg->seed + t; one all-reduce per Lloyd pass; the reference's
 * `iter > 1` break rule.  out_float_centroids[t]: K x 16*channels floats. */
YUPVQ_API int yupvq_encode_hdr_batch_sharded(yupvq_ctx* ctx, const float* const* images, size_t n_images, uint32_t width,
                                             uint32_t height, int channels, size_t first_block, size_t n_local,
                                             const yupvq_config* cfg, float* const* out_float_centroids,
                                             uint32_t* const* out_indices, yupvq_stats* stats);
/* One image: yupvq_encode_hdr_batch_sharded with n_images == 1. */
YUPVQ_API int yupvq_encode_hdr_sharded(yupvq_ctx* ctx, const float* image, uint32_t width, uint32_t height, int channels,
                                       size_t first_block, size_t n_local, const yupvq_config* cfg,
                                       float* out_float_centroids, uint32_t* out_indices, yupvq_stats* stats);

#ifdef __cplusplus
}
#endif
#endif /* YUPVQ_H */
