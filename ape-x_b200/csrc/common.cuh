// common.cuh — shared host/device helpers for libper_b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdarg.h>
#include <stdio.h>
#include <string>

#include "../../include/per_b200.h"

namespace perb {

typedef unsigned long long u64;

// ---- error plumbing -------------------------------------------------------------------
inline std::string &err_slot() {
    static thread_local std::string s;
    return s;
}
inline int fail(int code, const char *fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    err_slot() = buf;
    return code;
}
#define PER_CUDA(call)                                                                       \
    do {                                                                                     \
        cudaError_t e__ = (call);                                                            \
        if (e__ != cudaSuccess)                                                              \
            return perb::fail(PER_ERR_CUDA, "%s failed: %s (%s:%d)", #call,                  \
                              cudaGetErrorString(e__), __FILE__, __LINE__);                  \
    } while (0)

// Makes the handle's device current for the duration of one ABI call and puts the
// caller's device back afterwards (the caller is usually torch, which tracks its own).
struct DeviceGuard {
    int prev = -1;
    bool ok = true;
    explicit DeviceGuard(int dev) {
        if (cudaGetDevice(&prev) != cudaSuccess) { ok = false; return; }
        if (prev != dev && cudaSetDevice(dev) != cudaSuccess) ok = false;
        want = dev;
    }
    ~DeviceGuard() {
        if (prev >= 0 && prev != want) cudaSetDevice(prev);
    }
    int want = -1;
};

inline int ilog2_exact(int64_t x) {
    int l = 0;
    while ((int64_t(1) << l) < x) ++l;
    return l;
}
inline int64_t round_up(int64_t x, int64_t m) { return (x + m - 1) / m * m; }

// Launch with (or without) programmatic stream serialization: a kernel launched this way may become
// resident while its predecessor in the stream is still running and must execute pdl_wait() before it
// touches anything the predecessor reads or writes.  Only kernels that do so are launched through here
// with pdl = true.
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, bool pdl, Args... args) {
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof cfg);
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}

// ---- device helpers -------------------------------------------------------------------
// Timeline (diagnostics, per_tune "timeline"): every sample / gather / update launch owns TL_SLOTS consecutive
// entries of the list (its launch sequence number is a kernel argument, so launches captured in a CUDA graph
// keep their entries across replays); thread 0 of CTA 0 stores (tag << 56) | globaltimer ns with plain stores —
// no atomics, nothing on the kernel's critical path waits for the stamp.  tl == null: off.
constexpr int TL_SLOTS = 6;
__device__ __forceinline__ void tl_stamp(unsigned long long *tl, unsigned slot, unsigned tag) {
    if (!tl) return;
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    tl[slot] = ((unsigned long long)tag << 56) | (t & ((1ull << 56) - 1));
}

// Programmatic dependent launch: let the next kernel of the stream become resident / wait until the
// previous one has completed and its writes are visible.
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

// 64-bit L2-only loads: tree nodes are rewritten by other CTAs / earlier phases of the
// same kernel, so the update path never reads them through L1.
__device__ __forceinline__ double ldcg_f64(const double *p) { return __ldcg(p); }
__device__ __forceinline__ double2 ldcg_f64x2(const double *p) {
    return __ldcg(reinterpret_cast<const double2 *>(p));
}

__device__ __forceinline__ double shfl_f64(double x, int src) {
    return __shfl_sync(0xffffffffu, x, src);
}

// max over non-negative doubles / floats through their integer bit patterns
__device__ __forceinline__ void atomic_max_pos_f64(double *addr, double x) {
    atomicMax(reinterpret_cast<u64 *>(addr), (u64)__double_as_longlong(x));
}
__device__ __forceinline__ void atomic_max_pos_f32(float *addr, float x) {
    atomicMax(reinterpret_cast<unsigned int *>(addr), __float_as_uint(x));
}

// priority ** alpha (memory.py:317).  alpha == 1 and alpha == 0 are exact shortcuts.
__device__ __forceinline__ double pow_alpha(double p, double alpha) {
    if (alpha == 1.0) return p;
    if (alpha == 0.0) return 1.0;
    return pow(p, alpha);
}

}  // namespace perb
