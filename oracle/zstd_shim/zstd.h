/* TEST INFRASTRUCTURE -- prototype-only stand-in for <zstd.h> so that the UNMODIFIED reference
 * src/vq_bcn_compressor.cpp (its `#include <zstd.h>`, :3) compiles here: the reference vendors zstd as a git
 * submodule that is absent from /root/reference, and the image has libzstd.so.1 (1.5.5) but no header.
 * Only the 17 functions and 3 parameter ids that translation unit uses are declared, with the signatures and
 * enumerator values of zstd's stable public API (v1.4.0+); the reference objects link against the system
 * libzstd.so.1.  Used by oracle/build_ref.sh only. */
#ifndef YUP_ZSTD_SHIM_H
#define YUP_ZSTD_SHIM_H
#include <stddef.h>
#ifdef __cplusplus
extern "C" {
#endif
typedef struct ZSTD_CCtx_s ZSTD_CCtx;
typedef struct ZSTD_DCtx_s ZSTD_DCtx;
typedef struct ZSTD_CDict_s ZSTD_CDict;
typedef struct ZSTD_DDict_s ZSTD_DDict;
typedef enum { ZSTD_c_compressionLevel = 100, ZSTD_c_enableLongDistanceMatching = 160, ZSTD_c_nbWorkers = 400 } ZSTD_cParameter;
#define ZSTD_CONTENTSIZE_UNKNOWN (0ULL - 1)
#define ZSTD_CONTENTSIZE_ERROR (0ULL - 2)
ZSTD_CCtx* ZSTD_createCCtx(void);
size_t ZSTD_freeCCtx(ZSTD_CCtx* cctx);
ZSTD_DCtx* ZSTD_createDCtx(void);
size_t ZSTD_freeDCtx(ZSTD_DCtx* dctx);
size_t ZSTD_compressBound(size_t srcSize);
size_t ZSTD_CCtx_setParameter(ZSTD_CCtx* cctx, ZSTD_cParameter param, int value);
size_t ZSTD_compress2(ZSTD_CCtx* cctx, void* dst, size_t dstCapacity, const void* src, size_t srcSize);
size_t ZSTD_compress_usingCDict(ZSTD_CCtx* cctx, void* dst, size_t dstCapacity, const void* src, size_t srcSize, const ZSTD_CDict* cdict);
unsigned ZSTD_isError(size_t code);
const char* ZSTD_getErrorName(size_t code);
ZSTD_CDict* ZSTD_createCDict(const void* dictBuffer, size_t dictSize, int compressionLevel);
size_t ZSTD_freeCDict(ZSTD_CDict* cdict);
ZSTD_DDict* ZSTD_createDDict(const void* dictBuffer, size_t dictSize);
size_t ZSTD_freeDDict(ZSTD_DDict* ddict);
unsigned long long ZSTD_getFrameContentSize(const void* src, size_t srcSize);
size_t ZSTD_decompressDCtx(ZSTD_DCtx* dctx, void* dst, size_t dstCapacity, const void* src, size_t srcSize);
size_t ZSTD_decompress_usingDDict(ZSTD_DCtx* dctx, void* dst, size_t dstCapacity, const void* src, size_t srcSize, const ZSTD_DDict* ddict);
#ifdef __cplusplus
}
#endif
#endif
