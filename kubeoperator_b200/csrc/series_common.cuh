// Shared by the series kernels (series_rollup.cu, series_flow.cu): launch arguments and inline-PTX helpers.
#pragma once
#include <math.h>

#include "ko_common.cuh"

namespace {

constexpr int NI_MAX = KO_MAX_QUANTILES;

struct SeriesArgs {
    const double* values;
    const int64_t* offsets;
    int64_t n_series;
    double* out_sum;
    double* out_mean;
    double* out_min;
    double* out_max;
    double* out_last;
    int64_t* out_count;
    double* out_quant;
    // counters (16 ints, cleared at the start of every call):
    //   [0] long series (classify)          [1] work cursor of series_flow, first attempt
    //   [2] device error flag               [3] series that took more than one pass in series_long
    //   [4] series deferred by the first attempt          [5] work cursor of the wide second attempt
    //   [6] series deferred again (-> series_long)         [7] work cursor of series_long
    //   [8] longest series of the call (classify)
    int* counters;
    int* long_list;
    int* defer_list;  // series the first attempt of series_flow could not finish (counters[4] entries)
    // The long-series kernels are launched three times over different lists (see launch_long_path):
    const int* work_list;   // series to process ...
    int work_cnt;           // ... their number is counters[work_cnt]
    int work_cur;           // index of the kernel's work cursor in counters
    int* out_list;          // series_flow: where the series it cannot finish go ...
    int out_cnt;            // ... counted in counters[out_cnt]
    int wide;               // series_flow: second attempt, brackets 5 sigma wide, one undivided candidate pool
    int pack;               // n_series < 2^27: the first attempt's list carries one entry per unanswered quantile,
                            // (quantile << 27) | series; otherwise one entry per series
    // series_flow: candidates that do not fit a CTA's shared-memory pool spill to spill[blockIdx.x * spill_cap ...]
    unsigned long long* spill;
    int spill_cap;
    int nq;
    double q[NI_MAX];   // ascending
    int qcol[NI_MAX];   // output column of q[j]
};

// ------------------------------------------------------------------------------------------------ PTX
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred P1;\n"
        "KO_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
        "@P1 bra KO_DONE;\n"
        "bra KO_WAIT;\n"
        "KO_DONE:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
// Wait for roles that idle for long: one look at the barrier, then a plain nanosleep (measured on B200:
// nanosleep(400) suspends a warp for ~1 000 cycles, whatever the number of resident warps).  Whether the long waits
// poll every 400 ns or every 4 us, or sit in a try_wait loop, makes no measurable difference to the kernel.
template <int NS>
__device__ __forceinline__ void mbar_wait_sleep(uint64_t* bar, uint32_t parity) {
    const uint32_t addr = smem_u32(bar);
    for (;;) {
        uint32_t done;
        asm volatile(
            "{\n"
            ".reg .pred P1;\n"
            "mbarrier.test_wait.parity.shared::cta.b64 P1, [%1], %2;\n"
            "selp.u32 %0, 1, 0, P1;\n"
            "}\n"
            : "=r"(done)
            : "r"(addr), "r"(parity)
            : "memory");
        if (done) break;
        __nanosleep(NS);
    }
}
// TMA bulk copy global -> shared, completion counted in bytes on an mbarrier (SASS: UBLKCP).
__device__ __forceinline__ void tma_bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
            smem_u32(dst)),
        "l"(src), "r"(bytes), "r"(smem_u32(bar))
        : "memory");
}
// Bulk prefetch into L2: no shared memory, no completion to wait for (SASS: UBLKPF).
__device__ __forceinline__ void bulk_prefetch_l2(const void* src, uint32_t bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(src), "r"(bytes) : "memory");
}
// One look at a barrier, no waiting.
__device__ __forceinline__ bool mbar_test(uint64_t* bar, uint32_t parity) {
    uint32_t done;
    asm volatile(
        "{\n"
        ".reg .pred P1;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2;\n"
        "selp.u32 %0, 1, 0, P1;\n"
        "}\n"
        : "=r"(done)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return done != 0;
}
__device__ __forceinline__ void mbar_wait_addr(uint32_t bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred P1;\n"
        "KO_WAITA:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
        "@P1 bra KO_DONEA;\n"
        "bra KO_WAITA;\n"
        "KO_DONEA:\n"
        "}\n" ::"r"(bar),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ void mbar_arrive_addr(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
// Shared-memory atomics by explicit state space: through a generic pointer the compiler emits the slow generic
// ATOM path instead of ATOMS/REDS.
__device__ __forceinline__ void red_shared_add(const void* p, unsigned v) {
    asm volatile("red.shared.add.u32 [%0], %1;" ::"r"(smem_u32(p)), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned atom_shared_add(const void* p, unsigned v) {
    unsigned r;
    asm volatile("atom.shared.add.u32 %0, [%1], %2;" : "=r"(r) : "r"(smem_u32(p)), "r"(v) : "memory");
    return r;
}
__device__ __forceinline__ void red_shared_min64(const void* p, unsigned long long v) {
    asm volatile("red.shared.min.u64 [%0], %1;" ::"r"(smem_u32(p)), "l"(v) : "memory");
}
__device__ __forceinline__ void red_shared_max64(const void* p, unsigned long long v) {
    asm volatile("red.shared.max.u64 [%0], %1;" ::"r"(smem_u32(p)), "l"(v) : "memory");
}
__device__ __forceinline__ void st_generic_u64(unsigned long long addr, unsigned long long v) {
    asm volatile("st.u64 [%0], %1;" ::"l"(addr), "l"(v) : "memory");
}
__device__ __forceinline__ double lds_f64(uint32_t addr) {
    double x;
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(x) : "r"(addr));
    return x;
}
__device__ __forceinline__ void lds_v2f64(uint32_t addr, double& x, double& y) {
    asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(x), "=d"(y) : "r"(addr));
}

__device__ __forceinline__ unsigned nearest_rank_dev(unsigned n, double q) {
    double r = ceil(q * (double)n);
    if (!(r >= 1.0)) r = 1.0;
    if (r > (double)n) r = (double)n;
    return (unsigned)r;
}

}  // namespace
