/* internal.h — shared by the translation units of libmsbfs.so (not installed). */
#ifndef MSBFS_INTERNAL_H
#define MSBFS_INTERNAL_H

#include "../../include/msbfs.h"

#include <cstdarg>
#include <cstdio>

/* records the message for msbfs_last_error() and returns `code` */
int msbfs_fail(int code, const char* fmt, ...);
void msbfs_clear_error();

#ifdef __CUDACC__
#include <cuda_runtime.h>
#include <nvtx3/nvToolsExt.h>

/* NVTX range for the life of a scope: graph load, CSR build, a run, every lane batch.  Costs a
 * few nanoseconds when no tool is attached; shows up as named spans in Nsight Systems / ncu.
 * (The reference has only its two wall-clock windows, main.cu:235-297, 301-399.) */
struct MsbfsNvtxRange {
    explicit MsbfsNvtxRange(const char* name) { nvtxRangePushA(name); }
    ~MsbfsNvtxRange() { nvtxRangePop(); }
    MsbfsNvtxRange(const MsbfsNvtxRange&) = delete;
    MsbfsNvtxRange& operator=(const MsbfsNvtxRange&) = delete;
};
#define MSBFS_CUDA(call)                                                                  \
    do {                                                                                  \
        cudaError_t _e = (call);                                                          \
        if (_e != cudaSuccess)                                                            \
            return msbfs_fail(MSBFS_ERR_CUDA, "%s failed: %s (%s:%d)", #call,             \
                              cudaGetErrorString(_e), __FILE__, __LINE__);                \
    } while (0)
#endif

#endif
