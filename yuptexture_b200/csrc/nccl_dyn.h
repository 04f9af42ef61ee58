// NCCL bound at run time (dlopen) so that single-GPU use has no NCCL dependency and the library
// shares whatever libnccl.so.2 the host process already loaded (e.g. torch's bundled one).
// Only the handful of calls the VQ path needs: the per-pass K x 65 all-reduce.
#pragma once
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <stddef.h>
#include <stdint.h>

#include <cstring>
#include <string>

namespace yupvq {

struct NcclUniqueId { char internal[128]; };  // ncclUniqueId, nccl.h:38

struct NcclApi {
    void* handle = nullptr;
    std::string error;
    int (*p_get_unique_id)(NcclUniqueId*) = nullptr;
    int (*p_comm_init_rank)(void**, int, NcclUniqueId, int) = nullptr;
    int (*p_comm_destroy)(void*) = nullptr;
    int (*p_all_reduce)(const void*, void*, size_t, int, int, void*, cudaStream_t) = nullptr;
    int (*p_all_gather)(const void*, void*, size_t, int, void*, cudaStream_t) = nullptr;
    int (*p_broadcast)(const void*, void*, size_t, int, int, void*, cudaStream_t) = nullptr;
    const char* (*p_error_string)(int) = nullptr;

    enum { kUint8 = 1, kUint32 = 3, kInt64 = 4 };  // ncclDataType_t
    enum { kSum = 0, kMax = 2 };                   // ncclRedOp_t

    int load() {
        if (handle) return 0;
        const char* names[] = {"libnccl.so.2", "libnccl.so"};
        for (const char* n : names) {
            handle = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
            if (handle) break;
        }
        if (!handle) {
            error = dlerror() ? dlerror() : "libnccl.so.2 not found";
            return 1;
        }
        auto sym = [&](const char* n) { return dlsym(handle, n); };
        p_get_unique_id = reinterpret_cast<decltype(p_get_unique_id)>(sym("ncclGetUniqueId"));
        p_comm_init_rank = reinterpret_cast<decltype(p_comm_init_rank)>(sym("ncclCommInitRank"));
        p_comm_destroy = reinterpret_cast<decltype(p_comm_destroy)>(sym("ncclCommDestroy"));
        p_all_reduce = reinterpret_cast<decltype(p_all_reduce)>(sym("ncclAllReduce"));
        p_all_gather = reinterpret_cast<decltype(p_all_gather)>(sym("ncclAllGather"));
        p_broadcast = reinterpret_cast<decltype(p_broadcast)>(sym("ncclBroadcast"));
        p_error_string = reinterpret_cast<decltype(p_error_string)>(sym("ncclGetErrorString"));
        if (!p_get_unique_id || !p_comm_init_rank || !p_comm_destroy || !p_all_reduce || !p_all_gather || !p_broadcast) {
            error = "libnccl is missing required symbols";
            return 1;
        }
        return 0;
    }
    int unique_id(uint8_t out[128]) {
        NcclUniqueId id;
        if (p_get_unique_id(&id)) return 1;
        std::memcpy(out, id.internal, 128);
        return 0;
    }
    int init_rank(void** comm, int n, const uint8_t id_bytes[128], int rank) {
        NcclUniqueId id;
        std::memcpy(id.internal, id_bytes, 128);
        return p_comm_init_rank(comm, n, id, rank);
    }
    void destroy(void* comm) {
        if (p_comm_destroy && comm) p_comm_destroy(comm);
    }
    int all_reduce_sum_i64(void* comm, long long* buf, size_t count, cudaStream_t s) {
        return p_all_reduce(buf, buf, count, kInt64, kSum, comm, s);
    }
    int all_reduce_max_u32(void* comm, uint32_t* buf, size_t count, cudaStream_t s) {
        return p_all_reduce(buf, buf, count, kUint32, kMax, comm, s);
    }
    int all_gather_bytes(void* comm, const void* send, void* recv, size_t bytes_per_rank, cudaStream_t s) {
        return p_all_gather(send, recv, bytes_per_rank, kUint8, comm, s);
    }
    int broadcast_bytes(void* comm, void* buf, size_t bytes, int root, cudaStream_t s) {
        return p_broadcast(buf, buf, bytes, kUint8, root, comm, s);
    }
};

}  // namespace yupvq
