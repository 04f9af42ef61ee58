// yupvq_update_centroids_static: the Lloyd centroid update (reference src/vq_encoder.cpp:343-375) for callers that
// must not allocate GPU memory.
//
// Who that is: an asset pipeline running inside an engine or tool whose own allocator has already claimed the whole
// device (any cudaMalloc fails or evicts something the host application needs), code running in a frame loop where an
// allocation would stall every stream of the process while the driver maps memory, and plugins that are handed
// buffers by their host application and are forbidden to create device memory of their own.  Every other entry point of
// this library takes a yupvq_ctx, and both yupvq_create and the growth of a context's buffers call cudaMalloc /
// cudaMallocHost -- so those callers cannot use them.
//
// What this entry point does instead: the caller passes every input and output as device memory, and the one piece of
// working memory the update needs -- the 64-bit fixed-point sums and counts of all K clusters plus a few flag words --
// lives in a fixed array that is part of the library image (g_update_scratch below).  The array is sized for
// K <= 1024, the largest codebook CompressionConfig::SetQuality can ask for, which makes it 520 KiB; every GPU that
// loads the library holds those 520 KiB until the process exits, whether or not this entry point is ever called.
// Nothing here calls cudaMalloc, cudaFree, cudaMallocHost or a library routine that allocates.
//
// Consequences of the shared array, which the header repeats for users:
//   - it still holds the previous call's sums: the words a call uses are cleared at its start;
//   - two calls whose GPU work overlaps (two host threads, same device) use the same words and corrupt each other.
//     The call returns only after its own work has finished, so one thread calling repeatedly is always safe.
//   - the first call on a device looks up the array's device address (cudaGetSymbolAddress) and opts the accumulation
//     kernel in to its shared-memory size; make that first call outside any stream capture.
//
// The arithmetic is the existing update's, kernel for kernel (launch_max_abs, launch_accumulate,
// launch_finalize_centroids, launch_reseed_empty): results are bit-identical to yupvq_update_centroids.
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <mutex>

#include "../../include/yupvq.h"
#include "common.cuh"
#include "kernels.h"

using namespace yupvq;

namespace {

constexpr uint32_t kStaticMaxK = YUPVQ_STATIC_MAX_CODEBOOK;
constexpr size_t kFlagBytes = 256;                                   // flag words in front, on their own lines
constexpr size_t kAccWordsMax = (size_t)kStaticMaxK * 65 + 1;        // [K*64 sums | K counts | changed], int64 words
constexpr size_t kScratchWords = kFlagBytes / 8 + kAccWordsMax;
// flag words (uint32) at the head of the scratch: the first three are the layout the update kernels expect
enum { SF_CHANGED = 0, SF_EMPTY = 1, SF_RESEEDS = 2, SF_MAXABS = 3 };

// The working memory.  In HOST code this name is only a handle for cudaGetSymbolAddress -- it is not a device address
// and must never be passed to a kernel; scratch_address() below turns it into one, once per device.
static __device__ __align__(256) unsigned long long g_update_scratch[kScratchWords];

// Device code may name the array directly: the counts sit K*64 words into the accumulator.
__global__ void static_counts_kernel(uint32_t K, uint32_t* __restrict__ out_counts) {
    const uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c < K) out_counts[c] = (uint32_t)g_update_scratch[kFlagBytes / 8 + (size_t)K * 64 + c];
}

constexpr int kMaxDevices = 64;
std::mutex g_setup_mu;
void* g_scratch_addr[kMaxDevices] = {};   // device address of g_update_scratch on each device, filled on first use

// `device` must be the current device.
cudaError_t scratch_address(int device, void** out) {
    std::lock_guard<std::mutex> lock(g_setup_mu);
    if (!g_scratch_addr[device]) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_update_scratch);
        if (e != cudaSuccess) return e;
        g_scratch_addr[device] = p;
    }
    *out = g_scratch_addr[device];
    return cudaSuccess;
}

// the fixed-point scale of the sums: n values of magnitude <= max_abs must add up below 2^62 (same rule as the
// allocating update in api.cu, so both produce the same bits)
int sum_shift(size_t n_total, float max_abs) {
    int bits_n = 1;
    while (((size_t)1 << bits_n) < n_total && bits_n < 40) ++bits_n;
    int e = 0;
    if (max_abs > 0.f && std::isfinite(max_abs)) {
        (void)std::frexp(max_abs, &e);
    } else if (!std::isfinite(max_abs)) {
        e = 128;
    }
    const int shift = 62 - bits_n - (e > 0 ? e : 0);
    return shift < -60 ? -60 : (shift > 40 ? 40 : shift);
}

int fail(char* err, size_t err_len, const char* fmt, ...) {
    if (err && err_len) {
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(err, err_len, fmt, ap);
        va_end(ap);
    }
    return 1;
}

bool on_device(const void* p) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}

// makes `device` current for the call and puts the caller's device back afterwards
struct DeviceGuard {
    int prev = -1, target = -1;
    cudaError_t enter(int device) {
        target = device;
        if (cudaError_t e = cudaGetDevice(&prev); e != cudaSuccess) return e;
        return prev == device ? cudaSuccess : cudaSetDevice(device);
    }
    ~DeviceGuard() {
        if (prev >= 0 && prev != target) cudaSetDevice(prev);
    }
};

}  // namespace

#define SCK(call)                                                                                                 \
    do {                                                                                                          \
        const cudaError_t e_ = (call);                                                                            \
        if (e_ != cudaSuccess) return fail(err_text, err_len, "%s: %s", #call, cudaGetErrorString(e_));           \
    } while (0)

extern "C" YUPVQ_API int yupvq_update_centroids_static(int device, void* cuda_stream, const float* features, size_t n_blocks,
                                                       const uint32_t* assignments, float* inout_errors,
                                                       uint32_t codebook_size, float* inout_centroids, uint32_t* out_counts,
                                                       char* err_text, size_t err_len) {
    if (err_text && err_len) err_text[0] = 0;
    const uint32_t K = codebook_size;
    if (n_blocks == 0 || K == 0) return fail(err_text, err_len, "empty input");
    if (K > kStaticMaxK)
        return fail(err_text, err_len,
                    "codebook_size %u is beyond the %u clusters the built-in scratch array holds; "
                    "yupvq_update_centroids has no such limit (it allocates)",
                    K, kStaticMaxK);
    if (device < 0 || device >= kMaxDevices) return fail(err_text, err_len, "device %d out of range", device);
    if (!features || !assignments || !inout_errors || !inout_centroids) return fail(err_text, err_len, "null pointer argument");
    DeviceGuard guard;
    SCK(guard.enter(device));
    // host buffers would have to be staged, and staging needs an allocation: refuse them instead of faulting in a kernel
    if (!on_device(features) || !on_device(assignments) || !on_device(inout_errors) || !on_device(inout_centroids) ||
        (out_counts && !on_device(out_counts)))
        return fail(err_text, err_len, "every buffer must be device (or managed) memory: this entry point cannot stage host buffers");

    cudaStream_t s = reinterpret_cast<cudaStream_t>(cuda_stream);
    void* base = nullptr;
    SCK(scratch_address(device, &base));
    uint32_t* flags = reinterpret_cast<uint32_t*>(base);
    long long* acc = reinterpret_cast<long long*>(reinterpret_cast<uint8_t*>(base) + kFlagBytes);
    const size_t acc_words = (size_t)K * 65 + 1;

    // the array keeps whatever the previous call left: clear the flag words and this call's accumulator
    SCK(cudaMemsetAsync(base, 0, kFlagBytes + acc_words * 8, s));
    SCK(launch_max_abs(features, n_blocks * 64, flags + SF_MAXABS, s));
    float max_abs = 0.f;   // a 4-byte read-back into a local: the driver stages it through its own buffer
    SCK(cudaMemcpyAsync(&max_abs, flags + SF_MAXABS, 4, cudaMemcpyDeviceToHost, s));
    SCK(cudaStreamSynchronize(s));
    const int shift = sum_shift(n_blocks, max_abs);
    SCK(launch_accumulate(features, n_blocks, assignments, K, shift, acc, flags, nullptr, s));
    SCK(launch_finalize_centroids(acc, K, shift, inout_centroids, flags, nullptr, s));
    SCK(launch_reseed_empty(features, n_blocks, inout_errors, acc, K, inout_centroids, flags, nullptr, s));
    if (out_counts) {
        static_counts_kernel<<<(K + 255) / 256, 256, 0, s>>>(K, out_counts);
        SCK(cudaGetLastError());
    }
    SCK(cudaStreamSynchronize(s));
    return 0;
}
