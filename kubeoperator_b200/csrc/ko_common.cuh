// Shared host/device helpers of libkorollup.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <mutex>
#include <string>

#include "../../include/korollup.h"

#define KO_VERSION_INT 100  // 0.1.0

void ko_set_error(const char* fmt, ...);

#define KO_CUDA(expr)                                                                         \
    do {                                                                                      \
        cudaError_t _e = (expr);                                                              \
        if (_e != cudaSuccess) {                                                              \
            ko_set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__,    \
                         __LINE__);                                                           \
            return (_e == cudaErrorMemoryAllocation) ? KO_ENOMEM : KO_ECUDA;                  \
        }                                                                                     \
    } while (0)

#define KO_REQUIRE(cond, msg)            \
    do {                                 \
        if (!(cond)) {                   \
            ko_set_error("%s", msg);     \
            return KO_EINVAL;            \
        }                                \
    } while (0)

// A device buffer that only grows; owned by the ctx.
struct KoBuf {
    void* p = nullptr;
    size_t cap = 0;
};

struct ko_ctx {
    int device = 0;
    int num_sms = 0;
    std::mutex mu;
    cudaStream_t stream = nullptr;       // compute stream of host-pointer entry points
    cudaStream_t copy_stream = nullptr;  // H2D stream of the chunked series pipeline
    cudaEvent_t ev_copied[2] = {nullptr, nullptr};
    cudaEvent_t ev_done[2] = {nullptr, nullptr};
    cudaEvent_t ev_k0 = nullptr, ev_k1 = nullptr;  // around the long-series kernels of the last series rollup
    cudaEvent_t ev_p0 = nullptr, ev_p1 = nullptr;  // around its first series_flow launch alone (roofline timing)
    bool k_timed = false;
    cudaStream_t last_series_stream = nullptr;  // stream of the last ko_series_rollup_dev
    int* last_series_scratch = nullptr;         // its counters (device)
    KoBuf scratch;       // work lists / counters / sort buffers (kernel scratch)
    KoBuf stage[2];      // double-buffered device staging of series values
    KoBuf in_small;      // small inputs (offsets, node columns, ...)
    KoBuf out_small;     // device outputs before D2H
    void* pinned_out = nullptr;  // pinned bounce buffer for small D2H results
    size_t pinned_out_cap = 0;
    ko_counters counters = {0, 0, 0, 0};
    bool alive = false;
};

int ko_buf_reserve(KoBuf* b, size_t bytes);
int ko_pinned_reserve(ko_ctx* ctx, size_t bytes);

static inline size_t ko_align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

#ifdef __CUDACC__
// Order-preserving u64 image of an f64 (NaN handled by the callers).
__device__ __forceinline__ uint64_t ko_okey(double x) {
    uint64_t u = (uint64_t)__double_as_longlong(x);
    return (u >> 63) ? ~u : (u | 0x8000000000000000ull);
}
__device__ __forceinline__ double ko_key2d(uint64_t k) {
    uint64_t u = (k >> 63) ? (k & 0x7FFFFFFFFFFFFFFFull) : ~k;
    return __longlong_as_double((long long)u);
}
#define KO_KEY_NAN 0xFFFFFFFFFFFFFFFFull
#define KO_KEY_PINF 0xFFF0000000000000ull  // ko_okey(+inf)
#define KO_KEY_NINF 0x000FFFFFFFFFFFFFull  // ko_okey(-inf)

__device__ __forceinline__ double ko_ld_stream(const double* p) {
    double v;
    asm volatile("ld.global.nc.L1::no_allocate.f64 %0, [%1];" : "=d"(v) : "l"(p));
    return v;
}
#endif
