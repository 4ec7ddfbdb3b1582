// Window statistics over columnar f64 series: sum / count / mean / min / max / last + exact nearest-rank
// quantiles.  BUILD-DEFINED semantics (oracle/rollup_oracle.py series_stats); the reference only ships
// PromQL strings for this (core/apps/kubeops_api/prometheus_client.py:90-111), see include/korollup.h.
//
// HBM-bound by design: 8 algorithmic bytes per sample, read once.  Kernels of one call:
//   classify_series   — builds the work list of long series (scratch) and validates offsets
//   series_small      — one warp per series of <= 1024 samples: coalesced streaming loads, shuffle reductions;
//                       quantiles from a register bitonic sort (<= 256 samples) or a warp-private shared-memory
//                       radix select
//   series_plan       — one warp per long series: a strided 1024-point pre-sample, sorted in registers with
//                       shuffles, gives a value bracket per quantile (or a list of distinct levels for
//                       integer-valued series); the plan goes to a window in scratch
//   series_pipe       — persistent, one CTA of 23 warps per SM, warp-specialised: plan loader, TMA producer
//                       (cp.async.bulk + mbarrier ring, tiles asked into L2 ahead of it), 12 stream warps that
//                       fuse the statistics with "count below the bracket, keep what is inside it" in ONE pass,
//                       a prologue warp, 8 finish warps that select the exact order statistic among the kept
//                       candidates of the previous series meanwhile
//   series_pipe again — second attempt with wide brackets over the few quantiles the first could not answer
//   series_long       — exact multi-pass slow path (histogram refinement in key space) over what is left
// Every bracket is verified by exact counts, so the result is always the oracle's value.
#include <math.h>

#include "ko_common.cuh"

namespace {

constexpr int SMALL_MAX = 1024;      // series up to this length take the warp path
constexpr int NCONS = 256;           // consumer threads of series_long
constexpr int NCWARPS = NCONS / 32;
constexpr int NTHREADS = NCONS + 32; // + one producer warp
constexpr int TILE = 1024;           // doubles per ring stage (8 KB)
constexpr int STAGES = 4;
constexpr int SAMPLE_BIG = 1024;     // pre-sample size for L >= 8192
constexpr int SAMPLE_SMALL = 256;    // pre-sample size for shorter long series
constexpr int SLOTS = 32;            // thread-private candidate slots per consumer thread
constexpr int SLOTS_H = 24;          // ... when a histogram interval is active (its per-bin min/max keys use the rest)
constexpr int OVQ = 1024;            // shared overflow queue entries
constexpr int NI_MAX = KO_MAX_QUANTILES;
constexpr int HBINS = 256;
constexpr double CSIG = 3.0;         // bracket half width in binomial sigmas of the pre-sample

enum { M_COUNT = 0, M_CAPTURE = 1, M_HIST = 2 };
enum { A_NONE = 0, A_SELECT = 1 };

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
    int* counters;  // [0] n_long  [2] error flag  [3] series needing extra passes  [4] deferred by the fast path
                    // [6] deferred again by its wide second attempt  [7] work cursor of series_long
    int* long_list;
    int* defer_list;  // series the pipelined fast path could not finish (counters[4] entries)
    // The long-series kernels are launched three times over different lists (see launch_long_path):
    const int* work_list;   // series to process ...
    int work_cnt;           // ... their number is counters[work_cnt]
    int work_cur;           // series_long: index of its work cursor in counters
    int* out_list;          // series_pipe: where the series it cannot finish go ...
    int out_cnt;            // ... counted in counters[out_cnt]
    int wide;               // series_pipe: second attempt, one quantile per work item, twice the bracket width
    int pack;               // n_series < 2^27: the first attempt's list carries the unanswered quantiles in bits 27..30
    // series_plan writes one plan per work item w in [w_begin, w_begin + plan_cap) to plans[w - w_begin];
    // series_pipe, launched after it over the same window, fetches them
    void* plans;
    long long w_begin;
    int plan_cap;
    int nq;
    double q[NI_MAX];   // ascending
    int qcol[NI_MAX];   // output column of q[j]
};

struct Plan {
    int ni;
    int mode[NI_MAX];
    int hshift[NI_MAX];
    int slot0[NI_MAX], nslots[NI_MAX];
    int ov0[NI_MAX], ovcap[NI_MAX];
    unsigned long long kA[NI_MAX], kB[NI_MAX];
    double tA[NI_MAX], tB[NI_MAX];
    // per quantile
    int q_iv[NI_MAX];
    int q_done[NI_MAX];
    int q_act[NI_MAX];
    unsigned q_rho[NI_MAX];
    double q_val[NI_MAX];
    // pass results
    unsigned cA[NI_MAX];
    unsigned nin[NI_MAX];
    unsigned ovcnt[NI_MAX];
    int more;
    int passes;
    int fast;       // first pass can use the specialised loop (capture / count-only brackets + upper tail)
    int tail_iv;    // index of the upper-tail interval of the fast pass, or -1
    unsigned cmp_cnt;
    double trash;   // sink of captures that find both the private slots and the overflow queue full
    unsigned nvalid;
    int sample_valid;
    double sum, mn, mx, last;
    // radix select scratch
    int sel_digit;
    unsigned sel_cum, sel_cnt;
    unsigned long long sel_key;
    int work;
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
// Wait for roles that may idle for microseconds: suspend inside try_wait (time hint) and back off with
// nanosleep, so that waiting warps do not steal issue slots from the warps that do the work.
__device__ __forceinline__ void mbar_wait_sleep(uint64_t* bar, uint32_t parity) {
    const uint32_t addr = smem_u32(bar);
    for (;;) {
        uint32_t done;
        asm volatile(
            "{\n"
            ".reg .pred P1;\n"
            "mbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2, %3;\n"
            "selp.u32 %0, 1, 0, P1;\n"
            "}\n"
            : "=r"(done)
            : "r"(addr), "r"(parity), "r"(2000u)
            : "memory");
        if (done) break;
        __nanosleep(400);
    }
}
// Wait that suspends inside try_wait and is woken by the barrier: no polling sleep, no busy loop.
__device__ __forceinline__ void mbar_wait_hint(uint64_t* bar, uint32_t parity) {
    const uint32_t addr = smem_u32(bar);
    for (;;) {
        uint32_t done;
        asm volatile(
            "{\n"
            ".reg .pred P1;\n"
            "mbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2, %3;\n"
            "selp.u32 %0, 1, 0, P1;\n"
            "}\n"
            : "=r"(done)
            : "r"(addr), "r"(parity), "r"(20000u)
            : "memory");
        if (done) break;
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
__device__ __forceinline__ double lds_f64(uint32_t addr) {
    double x;
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(x) : "r"(addr));
    return x;
}
__device__ __forceinline__ void lds_v2f64(uint32_t addr, double& x, double& y) {
    asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(x), "=d"(y) : "r"(addr));
}
__device__ __forceinline__ void cons_sync() { asm volatile("bar.sync 1, %0;" ::"n"(NCONS) : "memory"); }

__device__ __forceinline__ unsigned nearest_rank_dev(unsigned n, double q) {
    double r = ceil(q * (double)n);
    if (!(r >= 1.0)) r = 1.0;
    if (r > (double)n) r = (double)n;
    return (unsigned)r;
}

// ------------------------------------------------------------------------------------ classify kernel
__global__ void classify_series(SeriesArgs a) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= a.n_series) return;
    int64_t L = a.offsets[i + 1] - a.offsets[i];
    if (L < 0 || L >= 2147483647LL) {
        atomicExch(&a.counters[2], 1);
        return;
    }
    if (L > SMALL_MAX) {
        int pos = atomicAdd(&a.counters[0], 1);
        a.long_list[pos] = (int)i;
    }
}

// ------------------------------------------------------------------------------- warp path (<= 1024)
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_min(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmin(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ double warp_max(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ unsigned warp_sum_u(unsigned v) { return __reduce_add_sync(0xffffffffu, v); }

// Locate the digit whose cumulative count first reaches rho.  Called by one full warp on a 256-bin
// histogram; returns (digit, count before the digit, count of the digit) in every lane.
__device__ __forceinline__ void warp_find_digit(const unsigned* hist, unsigned rho, int lane, int& digit,
                                                unsigned& before, unsigned& cnt) {
    unsigned c[8];
    unsigned s = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        c[i] = hist[lane * 8 + i];
        s += c[i];
    }
    unsigned incl = s;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        unsigned t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    unsigned ballot = __ballot_sync(0xffffffffu, incl >= rho);
    int src = ballot ? (__ffs(ballot) - 1) : 31;
    int d = 0;
    unsigned b = 0, n = 0;
    if (lane == src) {
        unsigned run = incl - s;
        d = 7;
        n = c[7];
        b = run + (s - c[7]);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (run + c[i] >= rho) {
                d = i;
                b = run;
                n = c[i];
                break;
            }
            run += c[i];
        }
        d += lane * 8;
    }
    digit = __shfl_sync(0xffffffffu, d, src);
    before = __shfl_sync(0xffffffffu, b, src);
    cnt = __shfl_sync(0xffffffffu, n, src);
}

// MSB-first radix select over keys[0..n) held in shared memory by ONE warp; NaN keys are skipped.
// kmin/kmax bound the valid keys, so the digits above their first difference are skipped.
__device__ unsigned long long warp_radix_select(const unsigned long long* keys, int n, unsigned rho,
                                                unsigned long long kmin, unsigned long long kmax,
                                                unsigned* hist, int lane) {
    unsigned long long diff = kmin ^ kmax;
    if (diff == 0) return kmin;
    int top = 63 - __clzll((long long)diff);
    int shift = (top >> 3) << 3;
    unsigned long long mask = (shift + 8 >= 64) ? 0ull : (~0ull << (shift + 8));
    unsigned long long prefix = kmin & mask;
    for (;;) {
        for (int i = lane; i < HBINS; i += 32) hist[i] = 0;
        __syncwarp();
        for (int i = lane; i < n; i += 32) {
            unsigned long long k = keys[i];
            if (k != KO_KEY_NAN && (k & mask) == prefix) red_shared_add(&hist[(unsigned)(k >> shift) & 255u], 1u);
        }
        __syncwarp();
        int digit;
        unsigned before, cnt;
        warp_find_digit(hist, rho, lane, digit, before, cnt);
        __syncwarp();
        rho -= before;
        prefix |= (unsigned long long)digit << shift;
        mask |= 0xFFull << shift;
        if (shift == 0) return prefix;
        shift -= 8;
    }
}

// Short windows (<= 256 samples, e.g. 5 min .. 1 h at 15 s): the warp sorts the series' keys in registers, K per
// lane (sorted position p lives in lane p / K, register p % K), with a bitonic network of shuffles; every
// quantile is then a single register read.  Stale samples carry KO_KEY_NAN and sort to the end.
template <int K>
__device__ __forceinline__ void warp_sort_keys(unsigned long long (&r)[K], int lane) {
    constexpr int N = 32 * K;
#pragma unroll
    for (int k = 2; k <= N; k <<= 1) {
#pragma unroll
        for (int j = k >> 1; j > 0; j >>= 1) {
            if (j >= K) {   // partner in another lane
                const int m = j / K;
                const bool lower = (lane & m) == 0;
#pragma unroll
                for (int e = 0; e < K; ++e) {
                    const bool up = (((lane * K) | e) & k) == 0;
                    const unsigned long long o = __shfl_xor_sync(0xffffffffu, r[e], m);
                    const unsigned long long lo = r[e] < o ? r[e] : o, hi = r[e] < o ? o : r[e];
                    r[e] = (lower == up) ? lo : hi;
                }
            } else {        // partner in another register of the same lane
#pragma unroll
                for (int e = 0; e < K; ++e) {
                    if ((e & j) == 0) {
                        const bool up = (((lane * K) | e) & k) == 0;
                        const unsigned long long x0 = r[e], x1 = r[e | j];
                        const unsigned long long lo = x0 < x1 ? x0 : x1, hi = x0 < x1 ? x1 : x0;
                        r[e] = up ? lo : hi;
                        r[e | j] = up ? hi : lo;
                    }
                }
            }
        }
    }
}
template <int K>
__device__ __forceinline__ void small_sorted_quantiles(const SeriesArgs& a, const double* v, int L, unsigned nv,
                                                       int64_t sid, int lane) {
    unsigned long long r[K];
#pragma unroll
    for (int e = 0; e < K; ++e) {
        const int i = lane + 32 * e;   // coalesced; which register holds which sample does not matter before the sort
        double x = __longlong_as_double(0x7FF8000000000000LL);
        if (i < L) x = ko_ld_stream(v + i);
        r[e] = (x == x) ? ko_okey(x + 0.0) : KO_KEY_NAN;
    }
    warp_sort_keys<K>(r, lane);
#pragma unroll
    for (int j = 0; j < NI_MAX; ++j) {
        if (j < a.nq) {
            double res = __longlong_as_double(0x7FF8000000000000LL);
            if (nv > 0) {
                const int p = (int)nearest_rank_dev(nv, a.q[j]) - 1;
                unsigned long long mine = 0;
#pragma unroll
                for (int e = 0; e < K; ++e)
                    if (e == (p % K)) mine = r[e];
                res = ko_key2d(__shfl_sync(0xffffffffu, mine, p / K));
            }
            if (lane == 0) a.out_quant[sid * a.nq + a.qcol[j]] = res;
        }
    }
}

__global__ void __launch_bounds__(256) series_small(SeriesArgs a) {
    extern __shared__ __align__(16) unsigned char smem_small[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t sid = (int64_t)blockIdx.x * 8 + warp;
    if (sid >= a.n_series) return;
    const int64_t off0 = a.offsets[sid];
    const int64_t Ll = a.offsets[sid + 1] - off0;
    if (Ll > SMALL_MAX || Ll < 0) return;
    const int L = (int)Ll;
    const double* v = a.values + off0;
    unsigned long long* keys = nullptr;
    unsigned* hist = nullptr;
    if (a.nq > 0) {
        keys = reinterpret_cast<unsigned long long*>(smem_small) + (size_t)warp * SMALL_MAX;
        hist = reinterpret_cast<unsigned*>(smem_small + (size_t)8 * SMALL_MAX * 8) + warp * HBINS;
    }
    double sum = 0.0, mn = INFINITY, mx = -INFINITY;
    unsigned nv = 0;
    int last_idx = -1;
    for (int i = lane; i < L; i += 32) {
        double x = ko_ld_stream(v + i);
        bool ok = (x == x);
        if (ok) {
            sum += x;
            nv++;
            last_idx = i;
        }
        mn = fmin(mn, x);
        mx = fmax(mx, x);
        if (keys && L > 256) keys[i] = ok ? ko_okey(x + 0.0) : KO_KEY_NAN;
    }
    sum = warp_sum(sum);
    mn = warp_min(mn);
    mx = warp_max(mx);
    nv = warp_sum_u(nv);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) last_idx = max(last_idx, __shfl_xor_sync(0xffffffffu, last_idx, o));
    if (lane == 0) {
        const double nan = __longlong_as_double(0x7FF8000000000000LL);
        if (a.out_sum) a.out_sum[sid] = sum;
        if (a.out_mean) a.out_mean[sid] = nv > 0 ? sum / (double)nv : 0.0;
        if (a.out_min) a.out_min[sid] = nv > 0 ? mn + 0.0 : nan;
        if (a.out_max) a.out_max[sid] = nv > 0 ? mx + 0.0 : nan;
        if (a.out_last) a.out_last[sid] = last_idx >= 0 ? v[last_idx] : nan;
        if (a.out_count) a.out_count[sid] = (int64_t)nv;
    }
    if (a.nq > 0 && a.out_quant && L <= 256) {   // sorted in registers (the samples are re-read: they sit in L1)
        if (L <= 32) small_sorted_quantiles<1>(a, v, L, nv, sid, lane);
        else if (L <= 64) small_sorted_quantiles<2>(a, v, L, nv, sid, lane);
        else if (L <= 128) small_sorted_quantiles<4>(a, v, L, nv, sid, lane);
        else small_sorted_quantiles<8>(a, v, L, nv, sid, lane);
    } else if (a.nq > 0 && a.out_quant) {
        __syncwarp();
        const unsigned long long kmin = ko_okey(mn + 0.0), kmax = ko_okey(mx + 0.0);
        for (int j = 0; j < a.nq; ++j) {
            double r = __longlong_as_double(0x7FF8000000000000LL);
            if (nv > 0) {
                unsigned rho = nearest_rank_dev(nv, a.q[j]);
                r = ko_key2d(warp_radix_select(keys, L, rho, kmin, kmax, hist, lane));
            }
            if (lane == 0) a.out_quant[sid * a.nq + a.qcol[j]] = r;
        }
    }
}

// ------------------------------------------------------------------------------ CTA path (> 1024)
struct LongSmem {
    // Fixed offsets from the dynamic shared-memory base (no pointer table: it would live in local memory and turn
    // every access into a generic load).
    unsigned char* base;
    __device__ __forceinline__ double* ring_() const { return reinterpret_cast<double*>(base); }
    __device__ __forceinline__ unsigned char* pool_() const { return base + (size_t)STAGES * TILE * 8; }
    __device__ __forceinline__ double* ovq_() const {
        return reinterpret_cast<double*>(base + (size_t)STAGES * TILE * 8 + (size_t)SLOTS * NCONS * 8);
    }
    __device__ __forceinline__ unsigned* hist_() const {
        return reinterpret_cast<unsigned*>(base + (size_t)STAGES * TILE * 8 + (size_t)SLOTS * NCONS * 8 + (size_t)OVQ * 8);
    }
    __device__ __forceinline__ unsigned* selhist_() const { return hist_() + NI_MAX * HBINS; }
    __device__ __forceinline__ double* red_() const { return reinterpret_cast<double*>(selhist_() + HBINS); }
    __device__ __forceinline__ Plan* plan_() const { return reinterpret_cast<Plan*>(red_() + NCWARPS * 4); }
    __device__ __forceinline__ uint64_t* full_() const {
        return reinterpret_cast<uint64_t*>(reinterpret_cast<unsigned char*>(plan_()) + 1024);
    }
    __device__ __forceinline__ uint64_t* empty_() const { return full_() + STAGES; }
};

constexpr int CMP_MAX = NI_MAX * HBINS * 4 / 8;  // compacted keys of the radix select alias the histograms
constexpr size_t LONG_SMEM_BYTES = (size_t)STAGES * TILE * 8 + (size_t)SLOTS * NCONS * 8 + (size_t)OVQ * 8 +
                                   (size_t)NI_MAX * HBINS * 4 + (size_t)HBINS * 4 +
                                   (size_t)NCWARPS * 4 * 8 + 1024 /* Plan */ + (size_t)2 * STAGES * 8;
static_assert(sizeof(Plan) <= 1024, "Plan must fit its shared-memory slot");
static_assert(LONG_SMEM_BYTES <= 115712, "two CTAs per SM need <= 113 KB each");

struct Acc {
    double sum, mn, mx;
    unsigned nvalid;
    unsigned cA[NI_MAX];
    unsigned cin[NI_MAX];
    uint32_t p[NI_MAX];     // next free private slot (byte offset into pool)
    uint32_t pend[NI_MAX];
};

// Rare: the thread's private slots are full — claim an entry of the interval's shared overflow queue and return
// its shared-memory address (a scratch word when the queue is full too; the attempt is still counted).
__device__ __noinline__ uint32_t overflow_addr(Plan* pl, double* ovq, int i) {
    const unsigned pos = atom_shared_add(&pl->ovcnt[i], 1u);
    if (pos < (unsigned)pl->ovcap[i]) return smem_u32(ovq + pl->ov0[i] + pos);
    return smem_u32(&pl->trash);
}
__device__ __forceinline__ void sts_f64(uint32_t addr, double x) {
    asm volatile("st.shared.f64 [%0], %1;" ::"r"(addr), "d"(x) : "memory");
}
// Store x into the thread's next private slot of interval i (or the overflow queue when they are full).
__device__ __forceinline__ void capture(uint32_t pool32, uint32_t& p, uint32_t pend, const LongSmem& sm, int i,
                                        double x) {
    uint32_t addr = pool32 + p;
    if (p == pend) addr = overflow_addr(sm.plan_(), sm.ovq_(), i);
    else p += NCONS * 8;
    sts_f64(addr, x);
}

// Histogram mode (series with heavy ties, brackets too wide for shared memory, refinement passes): 256 bins
// linear in key space over the bracket, plus the smallest and largest key seen per bin — a bin that holds a
// single distinct value (the usual case for integer-valued series) settles its quantile without another pass.
__device__ __forceinline__ unsigned long long* hist_bmin(unsigned char* pool) {
    return reinterpret_cast<unsigned long long*>(pool + (size_t)SLOTS_H * NCONS * 8);
}
__device__ __forceinline__ unsigned long long* hist_bmax(unsigned char* pool) {
    return hist_bmin(pool) + NI_MAX * HBINS;
}
__device__ __noinline__ void hist_add(const Plan* pl, unsigned* hist, unsigned char* pool, int i, double x) {
    const unsigned long long k = ko_okey(x + 0.0);
    const unsigned b = (unsigned)((k - pl->kA[i]) >> pl->hshift[i]);
    // Lanes that arrive together and hit the same bin are combined (series with few distinct values would
    // otherwise serialise 32-way on one shared-memory word): one add of the group size, one min, one max.
    const unsigned act = __activemask();
    const unsigned peers = __match_any_sync(act, b);
    const int lane = threadIdx.x & 31;
    unsigned long long kmn = k, kmx = k;
    for (unsigned m = peers & (peers - 1); m; m &= m - 1) {   // fold the peers' keys into the group extremes
        const unsigned long long o = __shfl_sync(peers, k, __ffs(m) - 1);
        kmn = o < kmn ? o : kmn;
        kmx = o > kmx ? o : kmx;
    }
    {
        const unsigned long long o = __shfl_sync(peers, k, __ffs(peers) - 1);
        kmn = o < kmn ? o : kmn;
        kmx = o > kmx ? o : kmx;
    }
    if (lane == __ffs(peers) - 1) {
        red_shared_add(&hist[i * HBINS + b], (unsigned)__popc(peers));
        red_shared_min64(&hist_bmin(pool)[i * HBINS + b], kmn);
        red_shared_max64(&hist_bmax(pool)[i * HBINS + b], kmx);
    }
}

// sum += x, nvalid += 1 unless x is NaN (predicated, no selects)
__device__ __forceinline__ void acc_valid(double x, double& sum, unsigned& nv) {
    asm("{\n.reg .pred p;\nsetp.eq.f64 p, %2, %2;\n@p add.f64 %0, %0, %2;\n@p add.u32 %1, %1, 1;\n}"
        : "+d"(sum), "+r"(nv)
        : "d"(x));
}

// ---- fast first pass: NB value brackets [tA, tB) (capture or count-only) + an optional upper tail x >= tT.
template <int NB, bool TAIL>
struct FastRegs {
    double tA[NB > 0 ? NB : 1], tB[NB > 0 ? NB : 1];
    bool cap[NB > 0 ? NB : 1];
    double tT;
};

template <int NB, bool TAIL>
__device__ __forceinline__ void fast_elems4(const double (&x)[4], Acc& acc, const FastRegs<NB, TAIL>& fr,
                                            const LongSmem& sm, uint32_t pool32) {
#pragma unroll
    for (int e = 0; e < 4; ++e) {
        acc_valid(x[e], acc.sum, acc.nvalid);
        if (x[e] < acc.mn) acc.mn = x[e];
        if (x[e] > acc.mx) acc.mx = x[e];
    }
#pragma unroll
    for (int i = 0; i < NB; ++i) {
        bool in[4];
        bool any = false;
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            const bool lt = x[e] < fr.tA[i];
            acc.cA[i] += lt ? 1u : 0u;
            in[e] = !lt && (x[e] < fr.tB[i]);
            any = any || in[e];
        }
        if (any) {
            if (fr.cap[i]) {
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    if (in[e]) capture(pool32, acc.p[i], acc.pend[i], sm, i, x[e]);
                }
            } else {
#pragma unroll
                for (int e = 0; e < 4; ++e) acc.cin[i] += in[e] ? 1u : 0u;
            }
        }
    }
    if (TAIL) {
        constexpr int i = NB;
        bool in[4];
        bool any = false;
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            in[e] = x[e] >= fr.tT;
            any = any || in[e];
        }
        if (any) {
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                if (in[e]) capture(pool32, acc.p[i], acc.pend[i], sm, i, x[e]);
            }
        }
    }
}

template <int NB, bool TAIL>
__device__ __forceinline__ void fast_elem1(double x, Acc& acc, const FastRegs<NB, TAIL>& fr, const LongSmem& sm) {
    const double nan = __longlong_as_double(0x7FF8000000000000LL);
    const double v[4] = {x, nan, nan, nan};  // NaN is neither valid nor inside any bracket
    fast_elems4<NB, TAIL>(v, acc, fr, sm, smem_u32(sm.pool_()));
}

template <int NB, bool TAIL>
__device__ __forceinline__ void consume_fast(const double* __restrict__ v, int L, int head, int body, int ntiles,
                                             uint32_t g, Acc& acc, const LongSmem& sm, int ctid, int lane) {
    FastRegs<NB, TAIL> fr;
#pragma unroll
    for (int i = 0; i < NB; ++i) {
        fr.tA[i] = sm.plan_()->tA[i];
        fr.tB[i] = sm.plan_()->tB[i];
        fr.cap[i] = sm.plan_()->mode[i] == M_CAPTURE;
    }
    fr.tT = TAIL ? sm.plan_()->tA[NB] : 0.0;
    if (ctid == 0 && head) fast_elem1<NB, TAIL>(v[0], acc, fr, sm);
    if (ctid == 1 && head + body < L) fast_elem1<NB, TAIL>(v[L - 1], acc, fr, sm);
    const uint32_t full0 = smem_u32(sm.full_()), empty0 = smem_u32(sm.empty_());
    const uint32_t ring0 = smem_u32(sm.ring_()) + (uint32_t)ctid * 16u;
    const uint32_t pool32 = smem_u32(sm.pool_());
    uint32_t s = g % STAGES, par = (g / STAGES) & 1u;
    int rem2 = body >> 1;  // double2 elements left
    for (int t = 0; t < ntiles; ++t) {
        mbar_wait_addr(full0 + s * 8u, par);
        const uint32_t base = ring0 + s * (uint32_t)(TILE * 8);
        double x[4];
        if (rem2 >= TILE / 2) {
            lds_v2f64(base, x[0], x[1]);
            lds_v2f64(base + NCONS * 16, x[2], x[3]);
        } else {
            const double nan = __longlong_as_double(0x7FF8000000000000LL);
            x[0] = x[1] = x[2] = x[3] = nan;
            if (ctid < rem2) lds_v2f64(base, x[0], x[1]);
            if (ctid + NCONS < rem2) lds_v2f64(base + NCONS * 16, x[2], x[3]);
        }
        fast_elems4<NB, TAIL>(x, acc, fr, sm, pool32);
        __syncwarp();
        if (lane == 0) mbar_arrive_addr(empty0 + s * 8u);
        rem2 -= TILE / 2;
        if (++s == STAGES) {
            s = 0;
            par ^= 1u;
        }
    }
}

// ---- generic pass (refinement passes, histogram first passes): NI intervals of any mode.
template <int NI, bool STATS>
__device__ __forceinline__ void generic_elem(double x, Acc& acc, const double (&tA)[NI_MAX],
                                             const double (&tB)[NI_MAX], const int (&mode)[NI_MAX],
                                             const LongSmem& sm, uint32_t pool32) {
    if (STATS) {
        acc_valid(x, acc.sum, acc.nvalid);
        if (x < acc.mn) acc.mn = x;
        if (x > acc.mx) acc.mx = x;
    }
#pragma unroll
    for (int i = 0; i < NI; ++i) {
        const bool lt = x < tA[i];
        acc.cA[i] += lt ? 1u : 0u;
        const bool in = !lt && (x < tB[i]);
        if (in) {
            if (mode[i] == M_CAPTURE) capture(pool32, acc.p[i], acc.pend[i], sm, i, x);
            else if (mode[i] == M_HIST) hist_add(sm.plan_(), sm.hist_(), sm.pool_(), i, x);
            else acc.cin[i]++;
        }
    }
}

template <int NI, bool STATS>
__device__ __noinline__ void consume_generic(const double* __restrict__ v, int L, int head, int body, int ntiles,
                                             uint32_t g, Acc* accp, const LongSmem* smp, int ctid, int lane) {
    const LongSmem& sm = *smp;
    Acc acc = *accp;
    double tA[NI_MAX], tB[NI_MAX];
    int mode[NI_MAX];
#pragma unroll
    for (int i = 0; i < NI_MAX; ++i) {
        tA[i] = (i < NI) ? sm.plan_()->tA[i] : 0.0;
        tB[i] = (i < NI) ? sm.plan_()->tB[i] : 0.0;
        mode[i] = (i < NI) ? sm.plan_()->mode[i] : M_COUNT;
    }
    const uint32_t pool32 = smem_u32(sm.pool_());
    if (ctid == 0 && head) generic_elem<NI, STATS>(v[0], acc, tA, tB, mode, sm, pool32);
    if (ctid == 1 && head + body < L) generic_elem<NI, STATS>(v[L - 1], acc, tA, tB, mode, sm, pool32);
    const uint32_t full0 = smem_u32(sm.full_()), empty0 = smem_u32(sm.empty_());
    const uint32_t ring0 = smem_u32(sm.ring_()) + (uint32_t)ctid * 16u;
    uint32_t s = g % STAGES, par = (g / STAGES) & 1u;
    int rem2 = body >> 1;
    for (int t = 0; t < ntiles; ++t) {
        mbar_wait_addr(full0 + s * 8u, par);
        const uint32_t base = ring0 + s * (uint32_t)(TILE * 8);
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            if (ctid + h * NCONS < rem2) {
                double x0, x1;
                lds_v2f64(base + h * NCONS * 16, x0, x1);
                generic_elem<NI, STATS>(x0, acc, tA, tB, mode, sm, pool32);
                generic_elem<NI, STATS>(x1, acc, tA, tB, mode, sm, pool32);
            }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive_addr(empty0 + s * 8u);
        rem2 -= TILE / 2;
        if (++s == STAGES) {
            s = 0;
            par ^= 1u;
        }
    }
    *accp = acc;
}

template <bool STATS>
__device__ __forceinline__ void consume_generic_dispatch(int ni, const double* v, int L, int head, int body,
                                                         int ntiles, uint32_t g, Acc* acc, const LongSmem* sm,
                                                         int ctid, int lane) {
    switch (ni) {
        case 0: consume_generic<0, STATS>(v, L, head, body, ntiles, g, acc, sm, ctid, lane); break;
        case 1: consume_generic<1, STATS>(v, L, head, body, ntiles, g, acc, sm, ctid, lane); break;
        case 2: consume_generic<2, STATS>(v, L, head, body, ntiles, g, acc, sm, ctid, lane); break;
        case 3: consume_generic<3, STATS>(v, L, head, body, ntiles, g, acc, sm, ctid, lane); break;
        default: consume_generic<4, STATS>(v, L, head, body, ntiles, g, acc, sm, ctid, lane); break;
    }
}

// ---- pre-sample sort: 64-bit packed keys (top 54 key bits | 10-bit sample index), KPT keys per thread held in
// registers.  Compare-exchange partners inside a thread are swapped in registers, partners inside the warp by
// shuffles, the few remaining stages through shared memory.  n = KPT * NCONS keys, element index = tid*KPT + e.
template <int KPT>
__device__ __forceinline__ void sort_sample_regs(unsigned long long (&r)[KPT], unsigned long long* buf, int ctid) {
    constexpr int N = KPT * NCONS;
#pragma unroll
    for (int k = 2; k <= N; k <<= 1) {
#pragma unroll
        for (int j = k >> 1; j > 0; j >>= 1) {
            if (j < KPT) {  // partner in the same thread
#pragma unroll
                for (int e = 0; e < KPT; ++e) {
                    if ((e & j) == 0) {
                        const int i = ctid * KPT + e;
                        const bool up = (i & k) == 0;
                        const unsigned long long a = r[e], b = r[e | j];
                        const bool sw = (a > b) == up;
                        r[e] = sw ? b : a;
                        r[e | j] = sw ? a : b;
                    }
                }
            } else if (j / KPT < 32) {  // partner in the same warp
                const int m = j / KPT;
                const bool lower = (ctid & m) == 0;
#pragma unroll
                for (int e = 0; e < KPT; ++e) {
                    const int i = ctid * KPT + e;
                    const bool up = (i & k) == 0;
                    const unsigned long long o = __shfl_xor_sync(0xffffffffu, r[e], m);
                    const bool take_min = (lower == up);
                    const bool o_less = o < r[e];
                    r[e] = (o_less == take_min) ? o : r[e];
                }
            } else {  // partner in another warp: exchange through shared memory
                const int m = j / KPT;
                cons_sync();
#pragma unroll
                for (int e = 0; e < KPT; ++e) buf[ctid * KPT + e] = r[e];
                cons_sync();
                const bool lower = (ctid & m) == 0;
#pragma unroll
                for (int e = 0; e < KPT; ++e) {
                    const int i = ctid * KPT + e;
                    const bool up = (i & k) == 0;
                    const unsigned long long o = buf[(ctid ^ m) * KPT + e];
                    const bool take_min = (lower == up);
                    const bool o_less = o < r[e];
                    r[e] = (o_less == take_min) ? o : r[e];
                }
            }
        }
    }
    cons_sync();
#pragma unroll
    for (int e = 0; e < KPT; ++e) buf[ctid * KPT + e] = r[e];
    cons_sync();
}

// ---- CTA-wide MSB-first radix select over the captured candidates of interval iv (thread-private slots plus
// overflow queue).  All candidate keys lie in [kA, kB).  Returns the key of local rank rho (1-based).  After the
// first pass the few candidates that carry the chosen digit are compacted into a list, so later passes are short.
template <typename F>
__device__ __forceinline__ void for_each_candidate(const LongSmem& sm, const double* ovq, int ovn, uint32_t p0,
                                                   uint32_t pcur, int ctid, F f) {
    for (uint32_t p = p0; p != pcur; p += NCONS * 8) f(ko_okey(*reinterpret_cast<const double*>(sm.pool_() + p) + 0.0));
    for (int i = ctid; i < ovn; i += NCONS) f(ko_okey(ovq[i] + 0.0));
}

__device__ unsigned long long cta_radix_select(const LongSmem& sm, int iv, unsigned rho, uint32_t p0,
                                               uint32_t pcur, int ctid, int lane) {
    Plan* pl = sm.plan_();
    const unsigned long long kA = pl->kA[iv], kB = pl->kB[iv];
    const unsigned long long diff = kA ^ (kB - 1);
    if (diff == 0) return kA;
    const int top = 63 - __clzll((long long)diff);
    int shift = (top >> 3) << 3;
    unsigned long long mask = (shift + 8 >= 64) ? 0ull : (~0ull << (shift + 8));
    unsigned long long prefix = kA & mask;
    const int ovn = (int)min(pl->ovcnt[iv], (unsigned)pl->ovcap[iv]);
    const double* ovq = sm.ovq_() + pl->ov0[iv];
    unsigned long long* cmp = reinterpret_cast<unsigned long long*>(sm.hist_());  // free once the pass was decided
    bool use_list = false;
    int nlist = 0;
    for (;;) {
        sm.selhist_()[ctid] = 0;
        cons_sync();
        if (use_list) {
            for (int i = ctid; i < nlist; i += NCONS) {
                const unsigned long long k = cmp[i];
                if ((k & mask) == prefix) red_shared_add(&sm.selhist_()[(unsigned)(k >> shift) & 255u], 1u);
            }
        } else {
            for_each_candidate(sm, ovq, ovn, p0, pcur, ctid, [&](unsigned long long k) {
                if ((k & mask) == prefix) red_shared_add(&sm.selhist_()[(unsigned)(k >> shift) & 255u], 1u);
            });
        }
        cons_sync();
        if (ctid < 32) {
            int digit;
            unsigned before, cnt;
            warp_find_digit(sm.selhist_(), rho, lane, digit, before, cnt);
            if (lane == 0) {
                pl->sel_digit = digit;
                pl->sel_cum = before;
                pl->sel_cnt = cnt;
                pl->cmp_cnt = 0;
            }
        }
        cons_sync();
        rho -= pl->sel_cum;
        prefix |= (unsigned long long)pl->sel_digit << shift;
        mask |= 0xFFull << shift;
        const unsigned cnt = pl->sel_cnt;
        if (shift == 0) return prefix;
        if (cnt == 1) {  // one candidate carries this prefix: fetch it
            if (use_list) {
                for (int i = ctid; i < nlist; i += NCONS)
                    if ((cmp[i] & mask) == prefix) pl->sel_key = cmp[i];
            } else {
                for_each_candidate(sm, ovq, ovn, p0, pcur, ctid, [&](unsigned long long k) {
                    if ((k & mask) == prefix) pl->sel_key = k;
                });
            }
            cons_sync();
            return pl->sel_key;
        }
        if (!use_list && cnt <= (unsigned)CMP_MAX) {
            for_each_candidate(sm, ovq, ovn, p0, pcur, ctid, [&](unsigned long long k) {
                if ((k & mask) == prefix) cmp[atom_shared_add(&pl->cmp_cnt, 1u)] = k;
            });
            cons_sync();
            use_list = true;
            nlist = (int)cnt;
        }
        shift -= 8;
    }
}

__device__ __forceinline__ int hist_shift(unsigned long long kA, unsigned long long kB) {
    unsigned long long span = kB - kA - 1;  // kB > kA
    if (span < 256) return 0;
    int top = 63 - __clzll((long long)span);
    return top - 7;
}

// Thread 0: derive the intervals of the first pass from the sorted pre-sample.
// spk: sorted packed keys (54 key bits | 10-bit index into ukeys); ukeys: the full keys in sample order.
__device__ void build_first_plan(Plan* pl, const unsigned long long* spk, const unsigned long long* ukeys,
                                 int valid_s, int L, const SeriesArgs& a) {
    const int nq = a.nq;
    int lo[NI_MAX], hi[NI_MAX];
    int ni = 0;
    for (int j = 0; j < nq; ++j) {
        int aj = -1, bj = valid_s;
        if (valid_s > 0) {
            const double q = a.q[j];
            const double pos = q * (double)valid_s;
            const double sig = sqrt(fmax(q * (1.0 - q), 0.0) * (double)valid_s);
            const int m = (int)ceil(CSIG * sig) + 2;
            aj = (int)floor(pos) - m - 1;
            bj = (int)ceil(pos) + m;
            if (aj < 0) aj = -1;
            if (bj > valid_s) bj = valid_s;
        }
        const int gap = valid_s / 32;
        if (ni > 0 && aj <= hi[ni - 1] + gap) {  // merge with the previous bracket
            if (bj > hi[ni - 1]) hi[ni - 1] = bj;
            if (aj < lo[ni - 1]) lo[ni - 1] = aj;
        } else {
            lo[ni] = aj;
            hi[ni] = bj;
            ni++;
        }
        pl->q_iv[j] = ni - 1;
        pl->q_done[j] = 0;
        pl->q_act[j] = A_NONE;
    }
    pl->ni = ni;
    pl->fast = 1;
    pl->tail_iv = -1;
    if (nq == 0) return;
    int need[NI_MAX];
    bool tail[NI_MAX];
    int total = 0;
    for (int i = 0; i < ni; ++i) {
        // Brackets are clamped to the sample's own extremes (a target beyond them is a rare miss that the
        // refinement passes catch); with no valid sample the bracket is everything.
        const bool any = valid_s > 0;
        const int a0 = lo[i] < 0 ? 0 : lo[i], b0 = hi[i] >= valid_s ? valid_s - 1 : hi[i];
        unsigned long long kA = KO_KEY_NINF, kB = KO_KEY_PINF;
        int mode = M_CAPTURE;
        tail[i] = false;
        if (any) {
            unsigned long long ka = ukeys[spk[a0] & 1023ull], kb = ukeys[spk[b0] & 1023ull];
            if (ka > kb) {  // both in one 54-bit bucket: order inside it is arbitrary
                const unsigned long long t = ka;
                ka = kb;
                kb = t;
            }
            kA = ka;
            if (ka == kb) {  // the sample sees a single value here: count its tie group
                mode = M_COUNT;
                kB = ka + 1;
            } else if (hi[i] >= valid_s) {  // reaches beyond the sample maximum: everything >= kA
                tail[i] = true;
                kB = KO_KEY_PINF + 1;
            } else {
                kB = (kb >= KO_KEY_PINF) ? KO_KEY_PINF : kb + 1;
            }
        }
        if (kB <= kA || (mode == M_COUNT && kA >= KO_KEY_PINF)) {  // only +inf in range
            kA = KO_KEY_PINF - 1;
            kB = KO_KEY_PINF;
            mode = M_COUNT;
            tail[i] = false;
        }
        pl->kA[i] = kA;
        pl->kB[i] = kB;
        pl->mode[i] = mode;
        need[i] = 0;
        if (mode == M_CAPTURE) {
            // the closed bracket holds the whole tie groups of its end values: widen the sample range over them
            int ea = a0, eb = b0;
            if (any) {
                const unsigned long long ta = spk[a0] >> 10, tb = spk[b0] >> 10;
                while (ea > 0 && (spk[ea - 1] >> 10) == ta) --ea;
                while (eb < valid_s - 1 && (spk[eb + 1] >> 10) == tb) ++eb;
                if (tail[i]) eb = valid_s - 1;
            }
            const double mass = valid_s > 0 ? (double)(eb - ea + 3) / (double)valid_s : 1.0;
            const double per_thread = fmin(mass, 1.0) * (double)L / (double)NCONS;
            need[i] = (int)ceil(per_thread) + 1;
            total += need[i];
        }
    }
    int nbr = 0;
    for (int i = 0; i < ni; ++i) nbr += tail[i] ? 0 : 1;
    if (nbr > 2) pl->fast = 0;
    // intervals that cannot be captured in shared memory get a key-space histogram instead
    while (total > (pl->fast ? SLOTS : SLOTS_H)) {
        int big = -1;
        for (int i = 0; i < ni; ++i)
            if (pl->mode[i] == M_CAPTURE && (big < 0 || need[i] > need[big])) big = i;
        if (big < 0) break;
        pl->mode[big] = M_HIST;
        pl->fast = 0;
        total -= need[big];
        need[big] = 0;
    }
    int ncap = 0;
    for (int i = 0; i < ni; ++i) ncap += (pl->mode[i] == M_CAPTURE);
    int slot = 0, spare = (pl->fast ? SLOTS : SLOTS_H) - total, k = 0;
    for (int i = 0; i < ni; ++i) {
        if (tail[i]) {
            if (pl->fast) pl->tail_iv = i;
            else pl->kB[i] = KO_KEY_PINF;  // the generic pass compares against tB; +inf is settled by rank
        }
        pl->tA[i] = ko_key2d(pl->kA[i]);
        pl->tB[i] = ko_key2d(pl->kB[i] > KO_KEY_PINF ? KO_KEY_PINF : pl->kB[i]);
        pl->hshift[i] = hist_shift(pl->kA[i], pl->kB[i]);
        pl->slot0[i] = slot;
        pl->nslots[i] = 0;
        pl->ov0[i] = 0;
        pl->ovcap[i] = 0;
        if (pl->mode[i] == M_CAPTURE) {
            const int extra = spare / ncap;
            pl->nslots[i] = need[i] + extra;
            slot += pl->nslots[i];
            pl->ov0[i] = k * (OVQ / ncap);
            pl->ovcap[i] = OVQ / ncap;
            k++;
        }
    }
}

// Thread 0: brackets of a refinement pass — one interval per unresolved quantile.
// nkA/nkB/npop hold the new key range and its known population per quantile.
__device__ void build_refine_plan(Plan* pl, int nq, const unsigned long long* nkA, const unsigned long long* nkB,
                                  const unsigned* npop) {
    int ni = 0, ncap = 0;
    int need[NI_MAX];
    for (int j = 0; j < nq; ++j) {
        if (pl->q_done[j]) continue;
        const int i = ni++;
        pl->q_iv[j] = i;
        pl->kA[i] = nkA[j];
        pl->kB[i] = nkB[j];
        pl->tA[i] = ko_key2d(nkA[j]);
        pl->tB[i] = ko_key2d(nkB[j]);
        pl->hshift[i] = hist_shift(nkA[j], nkB[j]);
        const unsigned long long per = ((unsigned long long)npop[j] + NCONS - 1) / NCONS;
        need[i] = (int)(per > 100000ull ? 100000ull : per);
        need[i] = need[i] + need[i] / 2 + 3;
        pl->mode[i] = M_CAPTURE;
    }
    pl->ni = ni;
    pl->fast = 0;
    pl->tail_iv = -1;
    // share the private slots evenly; what does not fit is histogrammed
    for (int i = 0; i < ni; ++i) {
        if (need[i] > SLOTS_H / ni) pl->mode[i] = M_HIST;
        else ncap++;
    }
    int k = 0;
    for (int i = 0; i < ni; ++i) {
        pl->slot0[i] = i * (SLOTS_H / ni);
        pl->nslots[i] = (pl->mode[i] == M_CAPTURE) ? SLOTS_H / ni : 0;
        pl->ov0[i] = 0;
        pl->ovcap[i] = 0;
        if (pl->mode[i] == M_CAPTURE) {
            pl->ov0[i] = k * (OVQ / ncap);
            pl->ovcap[i] = OVQ / ncap;
            k++;
        }
    }
}

#ifdef KO_PHASE_TIMING
__device__ unsigned long long g_phase[48];
#define KO_COUNT(reason, wide) atomicAdd(&g_phase[32 + (reason) + ((wide) ? 8 : 0)], 1ull)
#define KO_TICK(slot)                                                   \
    do {                                                                \
        if (blockIdx.x == 0 && threadIdx.x == 0) {                      \
            const unsigned long long _t = clock64();                    \
            g_phase[slot] += _t - _tick;                                \
            _tick = _t;                                                 \
        }                                                               \
    } while (0)
#else
#define KO_TICK(slot) do { } while (0)
#define KO_COUNT(reason, wide) do { } while (0)
#endif

template <int KPT>
__device__ __forceinline__ void presample(const double* __restrict__ v, int L, const LongSmem& sm, int ctid) {
    constexpr int S = KPT * NCONS;
    Plan* pl = sm.plan_();
    unsigned long long* ukeys = reinterpret_cast<unsigned long long*>(sm.pool_());        // full keys, sample order
    unsigned long long* spk = ukeys + SAMPLE_BIG;                                      // sorted packed keys
    unsigned long long r[KPT];
    int nv = 0;
#pragma unroll
    for (int e = 0; e < KPT; ++e) {
        const int i = ctid * KPT + e;
        const long long idx = ((2LL * i + 1) * (long long)L) / (2LL * S);
        const double x = v[idx];
        const bool ok = (x == x);
        const unsigned long long k = ok ? ko_okey(x + 0.0) : KO_KEY_NAN;
        ukeys[i] = k;
        r[e] = (k & ~1023ull) | (unsigned long long)i;
        nv += ok;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) nv += __shfl_xor_sync(0xffffffffu, nv, o);
    if ((ctid & 31) == 0 && nv) red_shared_add(&pl->sample_valid, (unsigned)nv);
    sort_sample_regs<KPT>(r, spk, ctid);  // ends with a consumer barrier: ukeys, spk, sample_valid visible
}

__global__ void __launch_bounds__(NTHREADS, 2) series_long(SeriesArgs a) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    LongSmem sm;
    sm.base = smem_raw;
    Plan* pl = sm.plan_();
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const bool producer = (warp == NCWARPS);
    const int ctid = tid;  // consumer thread index (valid when !producer)
    if (tid == 0) {
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&sm.full_()[s], 1);
            mbar_init(&sm.empty_()[s], NCWARPS);
        }
        mbar_fence_init();
    }
    __syncthreads();
    const int n_long = a.counters[a.work_cnt];
    const int* work_list = a.work_list;
    int* cursor = &a.counters[a.work_cur];
    const double nan = __longlong_as_double(0x7FF8000000000000LL);
    uint32_t g = 0;  // running ring tile index, identical in all threads
#ifdef KO_PHASE_TIMING
    unsigned long long _tick = clock64();
#endif

    for (;;) {
        if (tid == 0) pl->work = atomicAdd(cursor, 1);
        __syncthreads();
        const int w = pl->work;
        if (w >= n_long) break;
        KO_TICK(0);
        const int sid = work_list[w];
        const int64_t off0 = a.offsets[sid];
        const int L = (int)(a.offsets[sid + 1] - off0);
        const double* v = a.values + off0;
        const int head = (int)((reinterpret_cast<uintptr_t>(v) >> 3) & 1);
        const int body = (L - head) & ~1;
        const int ntiles = (body + TILE - 1) / TILE;
        const double* vbody = v + head;
        const int nq = a.nq;

        // ---------------- pre-sample, sort, first plan (consumers); last valid sample (thread 255)
        if (!producer) {
            if (ctid == 0) {
                pl->sample_valid = 0;
                pl->passes = 0;
                pl->more = 0;
                pl->ni = 0;
                pl->fast = 1;
                pl->tail_iv = -1;
            }
            if (ctid == NCONS - 1) {
                double lastv = nan;
                for (int i = L - 1; i >= 0; --i) {
                    const double x = v[i];
                    if (x == x) {
                        lastv = x;
                        break;
                    }
                }
                pl->last = lastv;
            }
            if (nq > 0) {
                cons_sync();
                if (L >= 8192) presample<SAMPLE_BIG / NCONS>(v, L, sm, ctid);
                else presample<SAMPLE_SMALL / NCONS>(v, L, sm, ctid);
                KO_TICK(1);
                if (ctid == 0) {
                    const unsigned long long* ukeys = reinterpret_cast<const unsigned long long*>(sm.pool_());
                    build_first_plan(pl, ukeys + SAMPLE_BIG, ukeys, pl->sample_valid, L, a);
                }
                KO_TICK(2);
            }
        }

        // ---------------- streaming passes
        bool first = true;
        for (;;) {
            __syncthreads();  // plan visible to everyone; previous pass fully consumed
            const int ni = pl->ni;
            if (producer) {
                if (lane == 0) {
                    for (int t = 0; t < ntiles; ++t) {
                        const uint32_t gg = g + (uint32_t)t;
                        const int s = (int)(gg % STAGES);
                        mbar_wait(&sm.empty_()[s], ((gg / STAGES) & 1u) ^ 1u);
                        const uint32_t bytes = (uint32_t)min(TILE, body - t * TILE) * 8u;
                        mbar_expect_tx(&sm.full_()[s], bytes);
                        tma_bulk_g2s(sm.ring_() + (size_t)s * TILE, vbody + (size_t)t * TILE, bytes, &sm.full_()[s]);
                    }
                }
            } else {
                Acc acc;
                acc.sum = 0.0;
                acc.mn = INFINITY;
                acc.mx = -INFINITY;
                acc.nvalid = 0;
                uint32_t p0[NI_MAX];
#pragma unroll
                for (int i = 0; i < NI_MAX; ++i) {
                    acc.cA[i] = 0;
                    acc.cin[i] = 0;
                    p0[i] = (i < ni) ? (uint32_t)(((size_t)pl->slot0[i] * NCONS + ctid) * 8) : 0u;
                    acc.p[i] = p0[i];
                    acc.pend[i] = (i < ni) ? p0[i] + (uint32_t)pl->nslots[i] * NCONS * 8 : 0u;
                }
                if (ctid < NI_MAX) {
                    pl->cA[ctid] = 0;
                    pl->nin[ctid] = 0;
                    pl->ovcnt[ctid] = 0;
                }
                if (!pl->fast) {
                    for (int i = ctid; i < NI_MAX * HBINS; i += NCONS) {
                        sm.hist_()[i] = 0;
                        hist_bmin(sm.pool_())[i] = ~0ull;
                        hist_bmax(sm.pool_())[i] = 0ull;
                    }
                }
                cons_sync();
                if (first && pl->fast) {
                    const int tl = pl->tail_iv >= 0 ? 1 : 0;
                    const int nb = ni - tl;
                    if (tl) {
                        if (nb == 0) consume_fast<0, true>(v, L, head, body, ntiles, g, acc, sm, ctid, lane);
                        else if (nb == 1) consume_fast<1, true>(v, L, head, body, ntiles, g, acc, sm, ctid, lane);
                        else consume_fast<2, true>(v, L, head, body, ntiles, g, acc, sm, ctid, lane);
                    } else {
                        if (nb == 0) consume_fast<0, false>(v, L, head, body, ntiles, g, acc, sm, ctid, lane);
                        else if (nb == 1) consume_fast<1, false>(v, L, head, body, ntiles, g, acc, sm, ctid, lane);
                        else consume_fast<2, false>(v, L, head, body, ntiles, g, acc, sm, ctid, lane);
                    }
                } else if (first) {
                    consume_generic_dispatch<true>(ni, v, L, head, body, ntiles, g, &acc, &sm, ctid, lane);
                } else {
                    consume_generic_dispatch<false>(ni, v, L, head, body, ntiles, g, &acc, &sm, ctid, lane);
                }

                KO_TICK(3);
                // ---- reduce
                if (first) {
                    const double ws = warp_sum(acc.sum), wmn = warp_min(acc.mn), wmx = warp_max(acc.mx);
                    const unsigned wn = warp_sum_u(acc.nvalid);
                    if (lane == 0) {
                        sm.red_()[warp * 4 + 0] = ws;
                        sm.red_()[warp * 4 + 1] = wmn;
                        sm.red_()[warp * 4 + 2] = wmx;
                        sm.red_()[warp * 4 + 3] = (double)wn;
                    }
                }
#pragma unroll
                for (int i = 0; i < NI_MAX; ++i) {
                    if (i < ni) {
                        const unsigned ca = warp_sum_u(acc.cA[i]);
                        const unsigned ci = warp_sum_u(acc.cin[i] + (acc.p[i] - p0[i]) / (NCONS * 8));
                        if (lane == 0) {
                            red_shared_add(&pl->cA[i], ca);
                            red_shared_add(&pl->nin[i], ci);
                        }
                    }
                }
                cons_sync();

                // ---- thread 0: decide per quantile
                unsigned long long nkA[NI_MAX], nkB[NI_MAX];
                unsigned npop[NI_MAX];
                if (ctid == 0) {
                    if (first) {
                        double s = 0.0, mn = INFINITY, mx = -INFINITY, n = 0.0;
                        for (int wq = 0; wq < NCWARPS; ++wq) {
                            s += sm.red_()[wq * 4 + 0];
                            mn = fmin(mn, sm.red_()[wq * 4 + 1]);
                            mx = fmax(mx, sm.red_()[wq * 4 + 2]);
                            n += sm.red_()[wq * 4 + 3];
                        }
                        pl->sum = s;
                        pl->mn = mn;
                        pl->mx = mx;
                        pl->nvalid = (unsigned)n;
                    }
                    const unsigned n = pl->nvalid;
                    if (pl->tail_iv >= 0) {  // upper tail: everything not inside it is below it
                        const int ti = pl->tail_iv;
                        pl->cA[ti] = n - (pl->nin[ti] + pl->ovcnt[ti]);
                    }
                    for (int j = 0; j < nq; ++j) {
                        pl->q_act[j] = A_NONE;
                        if (pl->q_done[j]) continue;
                        if (n == 0) {
                            pl->q_val[j] = nan;
                            pl->q_done[j] = 1;
                            continue;
                        }
                        const int i = pl->q_iv[j];
                        const unsigned r = nearest_rank_dev(n, a.q[j]);
                        const unsigned cA = pl->cA[i];
                        unsigned nb = pl->nin[i] + pl->ovcnt[i];
                        if (pl->mode[i] == M_HIST) {
                            nb = 0;
                            for (int b = 0; b < HBINS; ++b) nb += sm.hist_()[i * HBINS + b];
                        }
                        const unsigned long long kA = pl->kA[i], kB = pl->kB[i];
                        if (r <= cA) {  // below the bracket
                            nkA[j] = ko_okey(pl->mn + 0.0);   // series minimum (known since the first pass)
                            nkB[j] = kA;
                            npop[j] = cA;
                            if (nkB[j] <= nkA[j]) {  // cannot happen (cA would be 0); be safe
                                pl->q_val[j] = ko_key2d(KO_KEY_NINF);
                                pl->q_done[j] = 1;
                            }
                        } else if (r > cA + nb) {  // above the bracket
                            if (kB >= KO_KEY_PINF) {
                                pl->q_val[j] = ko_key2d(KO_KEY_PINF);
                                pl->q_done[j] = 1;
                            } else {
                                nkA[j] = kB;
                                nkB[j] = ko_okey(pl->mx + 0.0);    // series maximum; +inf is settled by rank
                                nkB[j] = nkB[j] >= KO_KEY_PINF ? KO_KEY_PINF : nkB[j] + 1;
                                npop[j] = n - cA - nb;
                                if (nkB[j] <= nkA[j]) {  // nothing finite above the bracket: it is +inf
                                    pl->q_val[j] = ko_key2d(KO_KEY_PINF);
                                    pl->q_done[j] = 1;
                                }
                            }
                        } else {
                            const unsigned rho = r - cA;
                            if (kB - kA == 1 || pl->mode[i] == M_COUNT) {
                                pl->q_val[j] = ko_key2d(kA);
                                pl->q_done[j] = 1;
                            } else if (pl->mode[i] == M_CAPTURE) {
                                if (pl->ovcnt[i] <= (unsigned)pl->ovcap[i]) {
                                    pl->q_act[j] = A_SELECT;
                                    pl->q_rho[j] = rho;
                                } else {  // overflowed: histogram the same bracket next
                                    nkA[j] = kA;
                                    nkB[j] = kB > KO_KEY_PINF ? KO_KEY_PINF : kB;
                                    npop[j] = 0xFFFFFFFFu;
                                }
                            } else {  // M_HIST: narrow to the bin holding the rank
                                unsigned cum = 0;
                                int b = 0;
                                for (; b < HBINS - 1; ++b) {
                                    const unsigned h = sm.hist_()[i * HBINS + b];
                                    if (cum + h >= rho) break;
                                    cum += h;
                                }
                                const int sh = pl->hshift[i];
                                const unsigned long long na = kA + ((unsigned long long)b << sh);
                                unsigned long long nb2 = na + (1ull << sh);
                                if (nb2 > kB || nb2 < na) nb2 = kB;
                                const unsigned long long bmn = hist_bmin(sm.pool_())[i * HBINS + b];
                                if (nb2 - na == 1) {
                                    pl->q_val[j] = ko_key2d(na);
                                    pl->q_done[j] = 1;
                                } else if (bmn == hist_bmax(sm.pool_())[i * HBINS + b]) {  // one distinct value
                                    pl->q_val[j] = ko_key2d(bmn);
                                    pl->q_done[j] = 1;
                                } else {
                                    nkA[j] = na;
                                    nkB[j] = nb2;
                                    npop[j] = sm.hist_()[i * HBINS + b];
                                }
                            }
                        }
                    }
                }
                KO_TICK(4);
                cons_sync();
                // ---- selects over captured candidates (all consumers)
                for (int j = 0; j < nq; ++j) {
                    if (pl->q_act[j] == A_SELECT) {
                        const int i = pl->q_iv[j];
                        uint32_t pp0 = 0, pcur = 0;
#pragma unroll
                        for (int ii = 0; ii < NI_MAX; ++ii)
                            if (ii == i) {
                                pp0 = p0[ii];
                                pcur = acc.p[ii];
                            }
                        const unsigned long long k = cta_radix_select(sm, i, pl->q_rho[j], pp0, pcur, ctid, lane);
                        if (ctid == 0) {
                            pl->q_val[j] = ko_key2d(k);
                            pl->q_done[j] = 1;
                        }
                        cons_sync();
                    }
                }
                KO_TICK(5);
                if (ctid == 0) {
                    int more = 0;
                    for (int j = 0; j < nq; ++j) more |= !pl->q_done[j];
                    pl->passes += 1;
                    if (more && pl->passes > 14) {  // cannot happen: every pass shrinks the key range 256-fold
                        for (int j = 0; j < nq; ++j)
                            if (!pl->q_done[j]) {
                                pl->q_val[j] = nan;
                                pl->q_done[j] = 1;
                            }
                        atomicExch(&a.counters[2], 2);
                        more = 0;
                    }
                    if (more) build_refine_plan(pl, nq, nkA, nkB, npop);
                    pl->more = more;
                }
            }
            g += (uint32_t)ntiles;
            first = false;
            __syncthreads();
            if (!pl->more) break;
        }

        // ---------------- outputs
        if (tid == 0) {
            const unsigned n = pl->nvalid;
            if (a.out_sum) a.out_sum[sid] = pl->sum;
            if (a.out_mean) a.out_mean[sid] = n > 0 ? pl->sum / (double)n : 0.0;
            if (a.out_min) a.out_min[sid] = n > 0 ? pl->mn + 0.0 : nan;
            if (a.out_max) a.out_max[sid] = n > 0 ? pl->mx + 0.0 : nan;
            if (a.out_last) a.out_last[sid] = pl->last;
            if (a.out_count) a.out_count[sid] = (int64_t)n;
            if (a.out_quant)
                for (int j = 0; j < nq; ++j) a.out_quant[(int64_t)sid * nq + a.qcol[j]] = pl->q_val[j];
            if (pl->passes > 1) atomicAdd(&a.counters[3], 1);
        }
        __syncthreads();
        KO_TICK(6);
    }
}


// =========================================================================================================
// series_pipe — the fast path for long series: one CTA per SM, warp-specialised, three series in flight.
//
//   PLANNER  (1 warp)   series k+2: strided pre-sample of 1024 values, sorted in registers (32 keys per lane,
//                       compare-exchange in registers and by shuffles, no shared memory, no CTA barrier),
//                       brackets per quantile, capture plan.
//   PRODUCER (1 warp)   series k+1: TMA bulk copies (cp.async.bulk + mbarrier) into a 5-stage ring.
//   STREAM   (12 warps) series k+1: fused sum / count / min / max + per-bracket counting and capture of the
//                       in-bracket samples into thread-private shared-memory slots (pool k+1 mod 2).
//   FINISH   (4 warps)  series k: combines the partials, radix-selects the exact order statistics among the
//                       captured samples (pool k mod 2), writes the outputs.
// Roles hand over through mbarriers only.  Anything unusual — brackets that do not fit shared memory, heavy tie
// groups, a bracket that missed its rank, an overflowing queue — is not handled here: the series is appended to
// the deferred list and recomputed exactly by series_long afterwards.
#ifdef KO_PHASE_TIMING
#define KO_PT0() const unsigned long long _pt0 = clock64()
#define KO_PT1(slot, cond)                                                   \
    do {                                                                     \
        if (blockIdx.x == 0 && (cond)) g_phase[slot] += clock64() - _pt0;    \
    } while (0)
#else
#define KO_PT0() do { } while (0)
#define KO_PT1(slot, cond) do { } while (0)
#endif
constexpr int P_SW = 12;                       // stream warps
constexpr int P_FW = 8;                        // finish warps
constexpr int P_NS = P_SW * 32;                // 384 stream threads
constexpr int P_NF = P_FW * 32;                // 128 finish threads
constexpr int P_PW = 1;                        // plan loader warp
constexpr int P_PLANS = 4;                     // plan slots
constexpr uint32_t P_PLAN_LOG = 2;
constexpr int P_THREADS = (P_PW + 1 + P_SW + P_FW + 1) * 32;   // plan loader, producer, stream, finish, prologue
constexpr int P_TILE = 4 * P_NS;               // 1536 doubles (12 KB) per ring stage
constexpr int P_STAGES = 3;
#ifndef KO_P_PF
#define KO_P_PF 4
#endif
constexpr int P_PF = KO_P_PF;                   // tiles asked into L2 ahead of the ring
constexpr int P_SLOTS = 24;                    // thread-private candidate slots per stream thread
constexpr int P_OVQ = 640;                     // overflow queue entries per pool
constexpr int P_NI = 3;                        // at most 2 brackets + 1 upper tail
// finish groups: all P_NF threads for one capturing interval, halves for two, three groups of 64 threads for three
static_assert(P_NF >= 192 && P_NF % 64 == 0 && P_NS % 64 == 0, "group sizes");
__host__ __device__ constexpr int fin_group_size(int ng) { return ng == 3 ? 64 : P_NF / ng; }
constexpr int P_LIST = 256;                    // candidates of one sub-bracket kept for the final rank counting
// Second attempt: brackets WIDE_MULT times as wide in rank space.  It works from the same pre-sample, so what it
// has to cover is a sample that was off by more than CSIG sigma; 1.5 x CSIG leaves that 1 in 400 while the
// candidates of a median bracket still fit a thread's 24 slots with room (2 x CSIG filled the overflow queue on a
// few series per 10 000, and each of those then cost a 0.2 ms trip through series_long).
constexpr double WIDE_MULT = 1.5;
constexpr int P_SAMPLE = 1024;
constexpr int P_CROW = P_NF + 4;               // bytes per row of the finish warps' counters
constexpr int P_LEVELS = 15;                   // distinct values a "levels" plan can list
constexpr int P_NTHR = 31;                     // thresholds searched per sample (2 per level + one +inf pad)
constexpr int P_MAX_L_LEVELS = 16000000;       // keeps the 16-bit per-thread counters of a levels plan exact

struct alignas(16) PipePlan {
    int sid;                 // series index, -1 = no more work
    int defer;               // not eligible for the fast path
    int L, head, body, ntiles;
    const double* v;
    double last;
    double smin, smax;       // smallest / largest valid pre-sample value (+inf / -inf without a sample)
    int ni, nb, tail;        // intervals: nb brackets first, then the optional upper tail
    int mode[P_NI];
    int slot0[P_NI], nslots[P_NI];
    unsigned long long kA[P_NI], kB[P_NI];
    double tA[P_NI], tB[P_NI];
    int q_iv[NI_MAX];
    // "levels" plan, for series with few distinct values around the wanted ranks (integer-valued gauges): no
    // candidate is captured; thr holds, ascending, each listed level v and the next double above it, padded
    // with +inf, and the stream warps count the samples per gap between consecutive thresholds.
    int vals, nlev;
    int qsel;                // wide attempt: the one quantile this work item answers (-1: all)
    int item;                // wide attempt: index of the series in the work list
    int vidx[P_LEVELS + 1];
    double thr[P_NTHR + 1];
};

// What the finish warps need to agree on before their first scan: valid count, extremes, which bracket holds which
// rank.  It follows from the stream warps' totals alone, so one thread (the prologue warp) works it out once per series.
struct FinPro {
    unsigned n;                              // valid samples
    int ncapi;                               // capturing intervals
    int failmask;                            // quantiles this attempt cannot answer (bracket missed, queue full)
    unsigned pad;
    double sum, mn, mx;
    unsigned long long kA[P_NI];             // effective bracket start per interval
    int ish[P_NI];                           // (key - kA) >> ish is the sub-bracket
    int ord[P_NI];                           // ordinal among the capturing intervals (-1: none)
    double dval[NI_MAX];                     // quantiles answered by counts alone
    int act_iv[NI_MAX];                      // interval in which quantile j is to be selected (-1: not)
    unsigned rho[NI_MAX];                    // its rank inside that bracket
};

struct PipeResult {          // what the stream warps hand to the finish warps, one per pool
    double sum[P_SW], mn[P_SW], mx[P_SW];          // per stream warp (summed in warp order: reproducible)
    // totals over all stream warps, added with shared-memory reductions, cleared by the finish warps
    unsigned t_nvalid, t_cA[P_NI], t_cin[P_NI], t_ncap[P_NI];
    FinPro pro;
    unsigned ovcnt[P_NI];                // captures that found the thread's private slots full, per interval
    unsigned ovn;                        // ... and in total: entries claimed in the shared overflow queue
    unsigned char ovtag[P_OVQ];          // interval of each overflow queue entry
    unsigned short pcount[P_NI][P_NS];   // captured entries per stream thread
    double trash;                        // sink of captures that find slots and queue full
};

struct PipeSel {             // scratch of the finish warps
    // per-thread byte counters of the 64 sub-brackets, zero between uses; the row stride of 49 words spreads the
    // sub-brackets over the shared-memory banks
    unsigned char cnt8[64][P_CROW];
    double list[NI_MAX][P_LIST];               // candidates of the chosen sub-bracket per quantile
    unsigned tot[P_NI][64];                    // sums of the counters per group
    unsigned char dummy[4][P_NF];              // counters of absent samples
    int bin[NI_MAX];                           // chosen sub-bracket per quantile
    unsigned rho[NI_MAX];                      // rank inside it
    unsigned nlist[NI_MAX];                    // candidates inside it
    // What finish thread 0 still reads after the last barrier of a series (and then resets): two copies, used by
    // alternate series, so that the other finish threads can start on the next series without waiting for it.
    struct Out {
        unsigned list_cnt[NI_MAX];   // entries of list[j]
        int fail;                    // bit j: quantile j could not be answered
        double val[NI_MAX];          // the selected order statistics
    } o[2];
};

__host__ __device__ constexpr size_t align16(size_t x) { return (x + 15) / 16 * 16; }
constexpr size_t PO_RING = 0;
constexpr size_t PO_POOL = PO_RING + (size_t)P_STAGES * P_TILE * 8;
constexpr size_t P_POOL_BYTES = (size_t)P_SLOTS * P_NS * 8;
constexpr size_t PO_OVQ = PO_POOL + 2 * P_POOL_BYTES;
constexpr size_t PO_PLAN = PO_OVQ + 2 * (size_t)P_OVQ * 8;
constexpr size_t PO_RES = PO_PLAN + align16(P_PLANS * sizeof(PipePlan));
constexpr size_t PO_SEL = PO_RES + align16(2 * sizeof(PipeResult));
constexpr size_t PO_BARS = PO_SEL + align16(sizeof(PipeSel));
constexpr size_t PIPE_SMEM_BYTES = PO_BARS + 32 * 8;
static_assert(PIPE_SMEM_BYTES <= 232448, "series_pipe must fit the 227 KB of one SM");

// All regions are fixed offsets from the dynamic shared-memory base: nothing here is indexed at run time
// through a table (a table would be placed in slow local memory).
struct PipeSmem {
    unsigned char* base;
    __device__ __forceinline__ double* ring() const { return reinterpret_cast<double*>(base + PO_RING); }
    __device__ __forceinline__ unsigned char* pool(uint32_t i) const { return base + PO_POOL + i * P_POOL_BYTES; }
    __device__ __forceinline__ double* ovq(uint32_t i) const {
        return reinterpret_cast<double*>(base + PO_OVQ + i * (size_t)P_OVQ * 8);
    }
    __device__ __forceinline__ PipePlan* plan(uint32_t i) const {
        return reinterpret_cast<PipePlan*>(base + PO_PLAN) + i;
    }
    __device__ __forceinline__ PipeResult* res(uint32_t i) const {
        return reinterpret_cast<PipeResult*>(base + PO_RES) + i;
    }
    __device__ __forceinline__ PipeSel* sel() const { return reinterpret_cast<PipeSel*>(base + PO_SEL); }
    __device__ __forceinline__ uint64_t* bars() const { return reinterpret_cast<uint64_t*>(base + PO_BARS); }
};

enum { B_RING_FULL = 0, B_RING_EMPTY = 5, B_PLAN_FULL = 10, B_PLAN_FREE = 18, B_POOL_FULL = 26, B_POOL_FREE = 28, B_HANDOFF = 30 };
static_assert(P_PLANS == (1 << P_PLAN_LOG) && B_PLAN_FREE - B_PLAN_FULL >= P_PLANS && B_POOL_FULL - B_PLAN_FREE >= P_PLANS, "barrier slots");

__device__ __forceinline__ void fin_sync() { asm volatile("bar.sync 2, %0;" ::"n"(P_NF) : "memory"); }

// ------------------------------------------------------------------------------------------ planner
// 1024 packed 32-bit keys, 32 per lane (element index = lane*32 + e), sorted ascending across the warp.
// A packed key is (22-bit position of the sample's key inside the sample's own key range) << 10 | sample index:
// fine enough to place the brackets, half the registers and instructions of a 64-bit sort; the exact bracket
// ends are re-read from the series through the index.
// Bitonic sort in the "flip" form: the first step of every merge pairs i with i ^ (k - 1), the later ones i with
// i ^ j, and the smaller key always goes to the lower index.  No step needs a direction, so a pair of registers
// costs one min and one max.
template <int J>
__device__ __forceinline__ void planner_stage_regs(unsigned (&r)[32]) {
#pragma unroll
    for (int e = 0; e < 32; ++e) {
        if ((e & J) == 0) {
            const unsigned lo = min(r[e], r[e | J]), hi = max(r[e], r[e | J]);
            r[e] = lo;
            r[e | J] = hi;
        }
    }
    if constexpr (J > 1) planner_stage_regs<J / 2>(r);
}
template <int K>   // runs of K / 2 keys inside a lane's registers -> runs of K
__device__ __forceinline__ void planner_merge_regs(unsigned (&r)[32]) {
#pragma unroll
    for (int e = 0; e < 32; ++e) {
        const int p = e ^ (K - 1);
        if (e < p) {
            const unsigned lo = min(r[e], r[p]), hi = max(r[e], r[p]);
            r[e] = lo;
            r[p] = hi;
        }
    }
    if constexpr (K >= 4) planner_stage_regs<K / 4>(r);
}
__device__ __forceinline__ void planner_sort(unsigned (&r)[32], int lane) {
    planner_merge_regs<2>(r);
    planner_merge_regs<4>(r);
    planner_merge_regs<8>(r);
    planner_merge_regs<16>(r);
    planner_merge_regs<32>(r);
#pragma unroll 1
    for (int k = 64; k <= P_SAMPLE; k <<= 1) {
        {   // flip step: lane ^ (k / 32 - 1), register 31 - e
            const int mask = (k >> 5) - 1;
            const bool lower = (lane & (k >> 6)) == 0;
#pragma unroll
            for (int e = 0; e < 16; ++e) {
                const unsigned x = r[e], y = r[31 - e];
                const unsigned ox = __shfl_xor_sync(0xffffffffu, y, mask);
                const unsigned oy = __shfl_xor_sync(0xffffffffu, x, mask);
                r[e] = lower ? min(x, ox) : max(x, ox);
                r[31 - e] = lower ? min(y, oy) : max(y, oy);
            }
        }
#pragma unroll 1
        for (int j = k >> 2; j >= 32; j >>= 1) {  // partner in another lane
            const int m = j >> 5;
            const bool lower = (lane & m) == 0;
#pragma unroll
            for (int e = 0; e < 32; ++e) {
                const unsigned o = __shfl_xor_sync(0xffffffffu, r[e], m);
                r[e] = lower ? min(r[e], o) : max(r[e], o);
            }
        }
        planner_stage_regs<16>(r);
    }
}

// sorted key at position p (uniform across the warp)
__device__ __forceinline__ unsigned planner_at(const unsigned (&r)[32], int p, int lane) {
    unsigned mine = 0;
#pragma unroll
    for (int e = 0; e < 32; ++e)
        if (e == (p & 31)) mine = r[e];
    return __shfl_sync(0xffffffffu, mine, p >> 5);
}

// number of sample keys whose 22-bit position is below t / not above t
__device__ __forceinline__ void planner_count(const unsigned (&r)[32], unsigned t, int& below, int& notabove) {
    int b = 0, na = 0;
#pragma unroll
    for (int e = 0; e < 32; ++e) {
        const unsigned k = r[e] >> 10;
        b += (k < t) ? 1 : 0;
        na += (k <= t) ? 1 : 0;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        b += __shfl_xor_sync(0xffffffffu, b, o);
        na += __shfl_xor_sync(0xffffffffu, na, o);
    }
    below = b;
    notabove = na;
}

__device__ __forceinline__ long long sample_pos(int i, int L) {
    return ((2LL * i + 1) * (long long)L) / (2LL * P_SAMPLE);
}

// tiny fixed-size arrays kept in registers: every access is unrolled to static indices
template <typename T, int N>
struct RegArr {
    T v[N];
    __device__ __forceinline__ T get(int i) const {
        T r = v[0];
#pragma unroll
        for (int t = 1; t < N; ++t)
            if (i == t) r = v[t];
        return r;
    }
    __device__ __forceinline__ void set(int i, T x) {
#pragma unroll
        for (int t = 0; t < N; ++t)
            if (i == t) v[t] = x;
    }
};

// Whole warp, every value uniform: build the plan of series `sid` into *pp.
__device__ __forceinline__ void planner_build(PipePlan* pp, const SeriesArgs& a, int sid, int qsel, int item,
                                              int lane) {
    const int64_t off0 = a.offsets[sid];
    const int L = (int)(a.offsets[sid + 1] - off0);
    const double* v = a.values + off0;
    const int head = (int)((reinterpret_cast<uintptr_t>(v) >> 3) & 1);
    const int body = (L - head) & ~1;
    const int nq = a.nq;
    const double nan = __longlong_as_double(0x7FF8000000000000LL);
    // last valid sample: look at the final 32 samples together, walk further back only if all are stale
    double lastv = nan;
    {
        const int i = L - 1 - lane;
        const double x = (i >= 0) ? v[i] : nan;
        const unsigned okm = __ballot_sync(0xffffffffu, x == x);
        if (okm) {
            lastv = __shfl_sync(0xffffffffu, x, __ffs(okm) - 1);
        } else {
            if (lane == 0)
                for (int t = L - 33; t >= 0; --t) {
                    const double y = v[t];
                    if (y == y) {
                        lastv = y;
                        break;
                    }
                }
            lastv = __shfl_sync(0xffffffffu, lastv, 0);
        }
    }
    int defer = 0;
    int soft = 0;   // the capture plan does not fit; a levels plan may
    int vals = 0, nlev = 0;
    int ni = 0, nb = 0, tail = 0;
    double smin = INFINITY, smax = -INFINITY;   // lane 0 only
    RegArr<int, P_NI> lo, hi;
    RegArr<int, NI_MAX> q_iv;
#pragma unroll
    for (int t = 0; t < P_NI; ++t) lo.v[t] = hi.v[t] = 0;
#pragma unroll
    for (int t = 0; t < NI_MAX; ++t) q_iv.v[t] = 0;
    if (nq > 0) {
        unsigned r[32];
        int nv = 0;
        {
            unsigned long long kk[32];
            unsigned long long kmin = KO_KEY_NAN, kmax = 0ull;
#pragma unroll
            for (int e = 0; e < 32; ++e) {
                const double x = v[sample_pos((lane << 5) | e, L)];
                const bool ok = (x == x);
                kk[e] = ok ? ko_okey(x + 0.0) : KO_KEY_NAN;
                nv += ok;
                if (ok) {
                    kmin = kk[e] < kmin ? kk[e] : kmin;
                    kmax = kk[e] > kmax ? kk[e] : kmax;
                }
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                nv += __shfl_xor_sync(0xffffffffu, nv, o);
                const unsigned long long a1 = __shfl_xor_sync(0xffffffffu, kmin, o);
                const unsigned long long a2 = __shfl_xor_sync(0xffffffffu, kmax, o);
                kmin = a1 < kmin ? a1 : kmin;
                kmax = a2 > kmax ? a2 : kmax;
            }
            const unsigned long long span = kmax > kmin ? kmax - kmin : 0ull;
            const int bits = span ? 64 - __clzll((long long)span) : 0;
            const int sh = bits > 22 ? bits - 22 : 0;
#pragma unroll
            for (int e = 0; e < 32; ++e) {
                const unsigned i = (unsigned)((lane << 5) | e);
                r[e] = (kk[e] == KO_KEY_NAN) ? 0xFFFFFFFFu : ((unsigned)((kk[e] - kmin) >> sh) << 10) | i;
            }
        }
        const int valid_s = nv;
        planner_sort(r, lane);
        if (valid_s < 64) {
            defer = 1;  // too few valid samples to bracket: leave it to the exact slow path
            if (lane == 0) KO_COUNT(0, a.wide);
        }
        if (!defer) {   // the sample's extremes seed the running min / max of the stream warps
            const unsigned p0 = planner_at(r, 0, lane), p1 = planner_at(r, valid_s - 1, lane);
            if (lane == 0) {
                smin = v[sample_pos((int)(p0 & 1023u), L)];
                smax = v[sample_pos((int)(p1 & 1023u), L)];
            }
        }
        int cur_lo = 0, cur_hi = 0;   // the interval being grown
#pragma unroll
        for (int j = 0; j < NI_MAX; ++j) {
            if (j < nq && !defer && (qsel < 0 || j == qsel)) {
                const double q = a.q[j];
                const double pos = q * (double)valid_s;
                const double sig = sqrt(fmax(q * (1.0 - q), 0.0) * (double)valid_s);
                const int m = (int)ceil((a.wide ? WIDE_MULT * CSIG : CSIG) * sig) + 2;
                int aj = (int)floor(pos) - m - 1, bj = (int)ceil(pos) + m;
                if (aj < 0) aj = -1;
                if (bj > valid_s) bj = valid_s;
                const int gap = valid_s / 32;
                if (ni > 0 && aj <= cur_hi + gap) {
                    if (bj > cur_hi) cur_hi = bj;
                    if (aj < cur_lo) cur_lo = aj;
                } else if (ni < P_NI) {
                    cur_lo = aj;
                    cur_hi = bj;
                    ni++;
                } else {
                    defer = 1;
                    if (lane == 0) KO_COUNT(1, a.wide);
                }
                lo.set(ni - 1, cur_lo);
                hi.set(ni - 1, cur_hi);
                q_iv.v[j] = ni - 1;
                if (qsel < 0 && j >= 2 && q_iv.v[j - 2] == ni - 1) soft = 1;   // three quantiles in one interval
            }
        }
        int total = 0, ncap = 0, ties = 0;
        int mode[P_NI], need[P_NI];
        unsigned long long kA[P_NI], kB[P_NI];
        bool last_is_tail = false;
#pragma unroll
        for (int i = 0; i < P_NI; ++i) {
            mode[i] = M_COUNT;
            need[i] = 0;
            kA[i] = kB[i] = 0;
            if (i < ni && !defer) {
                const int lo_i = lo.v[i], hi_i = hi.v[i];
                const int a0 = lo_i < 0 ? 0 : lo_i, b0 = hi_i >= valid_s ? valid_s - 1 : hi_i;
                const unsigned pa = planner_at(r, a0, lane), pb = planner_at(r, b0, lane);
                // exact keys of the two bracket ends (the packed key only carries the top 54 bits)
                double xa = 0.0, xb = 0.0;
                if (lane == 0) {
                    xa = v[sample_pos((int)(pa & 1023u), L)];
                    xb = v[sample_pos((int)(pb & 1023u), L)];
                }
                xa = __shfl_sync(0xffffffffu, xa, 0);
                xb = __shfl_sync(0xffffffffu, xb, 0);
                unsigned long long ka = ko_okey(xa + 0.0), kb = ko_okey(xb + 0.0);
                if (ka > kb) {
                    const unsigned long long t = ka;
                    ka = kb;
                    kb = t;
                }
                kA[i] = ka;
                mode[i] = M_CAPTURE;
                bool is_tail = false;
                if (ka == kb) {   // the whole bracket is one tie group: a levels plan answers it, or nothing here does
                    mode[i] = M_COUNT;
                    kB[i] = ka + 1;
                    ties = 1;
                    soft = 1;
                } else if (hi_i >= valid_s) {
                    is_tail = true;
                    kB[i] = KO_KEY_PINF + 1;
                } else {
                    kB[i] = (kb >= KO_KEY_PINF) ? KO_KEY_PINF : kb + 1;
                }
                if (kB[i] <= kA[i] || (mode[i] == M_COUNT && kA[i] >= KO_KEY_PINF)) defer = 1;  // +inf corner
                if (mode[i] == M_CAPTURE) {
                    int below_a, na_a, below_b, na_b;
                    planner_count(r, pa >> 10, below_a, na_a);
                    planner_count(r, pb >> 10, below_b, na_b);
                    if (na_a - below_a >= 4 || na_b - below_b >= 4) ties = 1;   // repeated values at a bracket end
                    const int ea = below_a, eb = is_tail ? valid_s - 1 : na_b - 1;  // whole tie groups at the ends
                    const double mass = (double)(eb - ea + 3) / (double)valid_s;
                    const double per_thread = fmin(mass, 1.0) * (double)L / (double)P_NS;
                    need[i] = (int)ceil(per_thread) + 1;
                    total += need[i];
                    ncap++;
                }
                if (is_tail) tail = 1; else nb++;
                last_is_tail = is_tail;
            }
        }
        if (nb > 2 || total > P_SLOTS) soft = 1;
        if (tail && !last_is_tail) soft = 1;
        // First attempt: repeated values at a bracket end send the series to a levels plan.  Second attempt: one
        // quantile has all the slots, so a capture plan that fits is taken first -- it also finds a value the
        // pre-sample never saw, which is what a levels plan fails on.
        if ((soft || (ties && !a.wide)) && !defer) {
            // Levels plan, preferred whenever it applies (tie-heavy brackets overflow the capture slots): list the distinct values the pre-sample shows inside the intervals.  Distinctness is
            // judged on the packed keys; whatever this misses only makes the finish warps defer the series.
            unsigned prev_last = __shfl_up_sync(0xffffffffu, r[31], 1);
            if (lane == 0) prev_last = 0xFFFFFFFFu;
            unsigned seen = 0xFFFFFFFFu;   // packed position of the last level listed by an earlier interval
#pragma unroll
            for (int i = 0; i < P_NI; ++i) {
                if (i < ni) {
                    const int lo_i = lo.v[i], hi_i = hi.v[i];
                    const int a0 = lo_i < 0 ? 0 : lo_i, b0 = hi_i >= valid_s ? valid_s - 1 : hi_i;
                    unsigned selmask = 0;
#pragma unroll
                    for (int e = 0; e < 32; ++e) {
                        const int p = (lane << 5) | e;
                        const unsigned prevk = e ? r[e > 0 ? e - 1 : 0] : prev_last;
                        const unsigned mine = r[e] >> 10;
                        const bool sl = p >= a0 && p <= b0 && (p == a0 ? mine != seen : mine != (prevk >> 10));
                        selmask |= sl ? (1u << e) : 0u;
                    }
                    seen = planner_at(r, b0, lane) >> 10;
                    int cnt = __popc(selmask), inc = cnt;
#pragma unroll
                    for (int o = 1; o < 32; o <<= 1) {
                        const int t = __shfl_up_sync(0xffffffffu, inc, o);
                        if (lane >= o) inc += t;
                    }
                    int ofs = nlev + inc - cnt;
#pragma unroll
                    for (int e = 0; e < 32; ++e) {
                        if ((selmask >> e) & 1u) {
                            if (ofs <= P_LEVELS) pp->vidx[ofs] = (int)(r[e] & 1023u);
                            ofs++;
                        }
                    }
                    nlev += __shfl_sync(0xffffffffu, inc, 31);
                }
            }
            __syncwarp();
            if (nlev < 1 || nlev > P_LEVELS || L > P_MAX_L_LEVELS) {
                defer = soft;   // too many levels: capture plan if that fits, else slow path
                if (soft && lane == 0) KO_COUNT(2, a.wide);
            } else {
                const double xv = (lane < nlev) ? v[sample_pos(pp->vidx[lane], L)] : 0.0;
                const unsigned long long key = ko_okey(xv + 0.0);
                const unsigned long long pk = __shfl_up_sync(0xffffffffu, key, 1);
                const bool bad = lane < nlev && (!(xv == xv) || key >= KO_KEY_PINF || (lane > 0 && key <= pk));
                if (__any_sync(0xffffffffu, bad)) {
                    defer = soft;
                } else {
                    if (lane < nlev) {
                        pp->thr[2 * lane] = ko_key2d(key);
                        pp->thr[2 * lane + 1] = ko_key2d(key + 1);
                    }
                    if (lane >= 2 * nlev && lane <= P_NTHR) pp->thr[lane] = INFINITY;
                    vals = 1;
                }
            }
        }
        if (!defer && !vals && lane == 0) {
            int slot = 0;
            const int spare = P_SLOTS - total;
            int big = 0;   // the interval with most candidates also takes what an even split leaves over
#pragma unroll
            for (int i = 1; i < P_NI; ++i)
                if (i < ni && need[i] > need[big]) big = i;
#pragma unroll
            for (int i = 0; i < P_NI; ++i) {
                if (i < ni) {
                    pp->mode[i] = mode[i];
                    pp->kA[i] = kA[i];
                    pp->kB[i] = kB[i];
                    pp->tA[i] = ko_key2d(kA[i]);
                    pp->tB[i] = ko_key2d(kB[i] > KO_KEY_PINF ? KO_KEY_PINF : kB[i]);
                    pp->slot0[i] = slot;
                    pp->nslots[i] = 0;
                    if (mode[i] == M_CAPTURE) {
                        pp->nslots[i] = need[i] + spare / ncap + (i == big ? spare % ncap : 0);
                        slot += pp->nslots[i];
                    }
                }
            }
        }
    }
    if (lane == 0) {
        pp->sid = sid;
        pp->defer = defer;
        pp->L = L;
        pp->head = head;
        pp->body = body;
        pp->ntiles = (body + P_TILE - 1) / P_TILE;
        pp->v = v;
        pp->last = lastv;
        pp->smin = smin;
        pp->smax = smax;
        pp->ni = ni;
        pp->nb = nb;
        pp->tail = tail;
        pp->vals = vals;
        pp->nlev = nlev;
        pp->qsel = qsel;
        pp->item = item;
#pragma unroll
        for (int j = 0; j < NI_MAX; ++j) pp->q_iv[j] = q_iv.v[j];
    }
    __syncwarp();
}

// ------------------------------------------------------------------------------------------ plan kernel
// One warp per work item, ahead of series_pipe: the plans are about 5 000 instructions of straight-line code per
// series, which inside the pipelined kernel kept evicting the stream and finish loops from the instruction caches.
#ifndef KO_PLAN_WARPS
#define KO_PLAN_WARPS 8
#endif
constexpr int PLAN_WARPS = KO_PLAN_WARPS;
#ifndef KO_PLAN_GRID
#define KO_PLAN_GRID 8   // CTAs per SM in the grid: finer than the two resident ones, so that constant and tie-heavy series do not leave warps idle
#endif
__global__ void __launch_bounds__(PLAN_WARPS * 32, 16 / PLAN_WARPS) series_plan(SeriesArgs a) {
    __shared__ PipePlan sp[PLAN_WARPS];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const long long n_items = (long long)a.counters[a.work_cnt] * (a.wide ? a.nq : 1);
    const long long hi = n_items < a.w_begin + a.plan_cap ? n_items : a.w_begin + a.plan_cap;
    const long long gw = (long long)blockIdx.x * PLAN_WARPS + warp, nw = (long long)gridDim.x * PLAN_WARPS;
    PipePlan* pp = &sp[warp];
    if (blockIdx.x == 0 && threadIdx.x == 0) a.counters[a.work_cur] = 0;   // cursor of the series_pipe launch that follows
    for (long long w = a.w_begin + gw; w < hi; w += nw) {
        const int item = a.wide ? (int)(w / a.nq) : (int)w;
        const int qsel = a.wide ? (int)(w - (long long)item * a.nq) : -1;
        const int entry = a.work_list[item];
        const bool packed = a.wide && a.pack;
        const int sid = packed ? (entry & 0x07FFFFFF) : (entry & 0x7FFFFFFF);
        if (packed && !((entry >> (27 + qsel)) & 1)) {   // this quantile was answered by the first attempt
            if (lane == 0) {
                pp->sid = sid;
                pp->defer = 2;
            }
        } else {
            planner_build(pp, a, sid, qsel, item, lane);
        }
        __syncwarp();
        const uint4* src = reinterpret_cast<const uint4*>(pp);
        uint4* dst = reinterpret_cast<uint4*>(reinterpret_cast<PipePlan*>(a.plans) + (w - a.w_begin));
        for (int t = lane; t < (int)(sizeof(PipePlan) / 16); t += 32) dst[t] = src[t];
        __syncwarp();
    }
    // wide attempt: what does not fit the plan window goes straight to the slow path's list (once per series)
    if (a.wide && lane == 0)
        for (long long w = hi + gw; w < n_items; w += nw) {
            const int item = (int)(w / a.nq);
            const int entry = atomicOr(const_cast<int*>(&a.work_list[item]), (int)0x80000000);
            if (entry >= 0) {
                const int pos = atomicAdd(&a.counters[a.out_cnt], 1);
                a.out_list[pos] = a.pack ? (entry & 0x07FFFFFF) : entry;
            }
        }
}

// ------------------------------------------------------------------------------------------ stream
struct PAcc {
    double sum, mn, mx;
    unsigned nvalid;
    unsigned cA[2];
    uint32_t p[P_NI], pend[P_NI];
};

struct PStreamCtx {
    uint32_t pool32;       // shared address of this series' pool
    uint32_t ovq32;        // shared address of its overflow queue
    uint32_t ovcnt32;      // shared address of PipeResult::ovcnt
    uint32_t ovn32;        // ... of PipeResult::ovn
    uint32_t ovtag32;      // ... of PipeResult::ovtag
    uint32_t trash32;
};

__device__ __forceinline__ unsigned lds_u8(uint32_t addr) {
    unsigned c;
    asm volatile("ld.shared.u8 %0, [%1];" : "=r"(c) : "r"(addr) : "memory");
    return c;
}
__device__ __forceinline__ void sts_u8(uint32_t addr, unsigned c) {
    asm volatile("st.shared.u8 [%0], %1;" ::"r"(addr), "r"(c) : "memory");
}
// Helpers of the stream loop.  The compare chains are PTX so that the compiler does not rewrite them into a
// min/max tree, and the rare updates are out-of-line calls so that they stay behind a real branch instead of
// being predicated into every iteration.
// t is NaN, or one of a..d is below mn or above mx: one predicate for the three rare events of a step
__device__ __forceinline__ bool rare4(double t, double a, double b, double c, double d, double mn, double mx) {
    unsigned r;
    asm("{\n.reg .pred p, q, n;\nsetp.neu.f64 n, %1, %1;\n"
        "setp.lt.f64 p, %2, %6;\nsetp.gt.f64 q, %2, %7;\n"
        "setp.lt.or.f64 p, %3, %6, p;\nsetp.gt.or.f64 q, %3, %7, q;\n"
        "setp.lt.or.f64 p, %4, %6, p;\nsetp.gt.or.f64 q, %4, %7, q;\n"
        "setp.lt.or.f64 p, %5, %6, p;\nsetp.gt.or.f64 q, %5, %7, q;\n"
        "or.pred p, p, q;\nor.pred p, p, n;\nselp.u32 %0, 1, 0, p;\n}"
        : "=r"(r) : "d"(t), "d"(a), "d"(b), "d"(c), "d"(d), "d"(mn), "d"(mx));
    return r != 0;
}
struct Stats4 {
    double sum, mn, mx;
    unsigned nv;
};
__device__ __noinline__ Stats4 rare_nan4(Stats4 st, double a, double b, double c, double d) {
    const double v[4] = {a, b, c, d};
#pragma unroll
    for (int e = 0; e < 4; ++e) {
        if (v[e] == v[e]) {
            st.sum += v[e];
            st.nv++;
        }
        st.mn = fmin(st.mn, v[e]);
        st.mx = fmax(st.mx, v[e]);
    }
    return st;
}

// One sample against a bracket / the upper tail, as predicated PTX (the compiler's own lowering of the same C++
// spends two instructions per conditional count).  p is the shared-memory address of the thread's next free
// private slot, pend the end of its slots; a sample that finds them full sets BIT in ovm and is queued by the
// caller (rare), so that the common path has no branch per sample and no "is there room" test per step.
template <int BIT>
__device__ __forceinline__ void cap_bracket(double x, double tA, double tB, unsigned& cA, uint32_t& p, uint32_t pend,
                                            unsigned& ovm) {
    asm volatile(
        "{\n.reg .pred lt, in, st, ov;\n"
        "setp.lt.f64 lt, %3, %4;\n"
        "@lt add.u32 %0, %0, 1;\n"
        "setp.lt.and.f64 in, %3, %5, !lt;\n"
        "setp.ne.and.u32 st, %1, %6, in;\n"
        "setp.eq.and.u32 ov, %1, %6, in;\n"
        "@st st.shared.f64 [%1], %3;\n"
        "@st add.u32 %1, %1, %7;\n"
        "@ov or.b32 %2, %2, %8;\n}"
        : "+r"(cA), "+r"(p), "+r"(ovm) : "d"(x), "d"(tA), "d"(tB), "r"(pend), "n"(P_NS * 8), "n"(BIT) : "memory");
}
template <int BIT>
__device__ __forceinline__ void cap_tail(double x, double tT, uint32_t& p, uint32_t pend, unsigned& ovm) {
    asm volatile(
        "{\n.reg .pred in, st, ov;\n"
        "setp.ge.f64 in, %2, %3;\n"
        "setp.ne.and.u32 st, %0, %4, in;\n"
        "setp.eq.and.u32 ov, %0, %4, in;\n"
        "@st st.shared.f64 [%0], %2;\n"
        "@st add.u32 %0, %0, %5;\n"
        "@ov or.b32 %1, %1, %6;\n}"
        : "+r"(p), "+r"(ovm) : "d"(x), "d"(tT), "r"(pend), "n"(P_NS * 8), "n"(BIT) : "memory");
}
// A capture that found the thread's private slots full: claim an entry of the shared overflow queue.
__device__ __forceinline__ void pipe_overflow(uint32_t ovn32, uint32_t ovcnt32, uint32_t ovq32, uint32_t ovtag32,
                                           int i, double x) {
    unsigned pos;
    asm volatile("atom.shared.add.u32 %0, [%1], 1;" : "=r"(pos) : "r"(ovn32) : "memory");
    asm volatile("red.shared.add.u32 [%0], 1;" ::"r"(ovcnt32 + 4u * i) : "memory");
    if (pos < (unsigned)P_OVQ) {
        sts_f64(ovq32 + pos * 8u, x);
        sts_u8(ovtag32 + pos, (unsigned)i);
    }
}

template <int NB, bool TAIL>
struct PFast {
    double tA[NB > 0 ? NB : 1], tB[NB > 0 ? NB : 1];
    double tT;
};

template <int NB, bool TAIL>
__device__ __forceinline__ void pipe_elems4(const double (&x)[4], PAcc& acc, const PFast<NB, TAIL>& fr,
                                            const PStreamCtx& c, int lane) {
    // sum / count / min / max, four samples at once.  A NaN among them (staleness marker, rare) takes the
    // per-sample path.  With a pre-sample the running min / max start at the sample extremes, so an update is
    // rare and is tested with four compares; with an upper tail every sample >= tT is captured and the finish
    // warps take the maximum from there.
    {
        const double t = (x[0] + x[1]) + (x[2] + x[3]);
        if (NB > 0 || TAIL) {   // seeded: one test (and one branch) for "a NaN, a new minimum or a new maximum"
            if (!rare4(t, x[0], x[1], x[2], x[3], acc.mn, acc.mx)) {
                acc.sum += t;
                acc.nvalid += 4;
            } else {   // a stale sample or a new record among the four: per sample, in line
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const bool ok = x[e] == x[e];
                    acc.sum += ok ? x[e] : 0.0;
                    acc.nvalid += ok ? 1u : 0u;
                    acc.mn = fmin(acc.mn, x[e]);   // fmin / fmax ignore a NaN operand
                    acc.mx = fmax(acc.mx, x[e]);
                }
            }
        } else if (t == t) {
            acc.sum += t;
            acc.nvalid += 4;
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                if (x[e] < acc.mn) acc.mn = x[e];
                if (x[e] > acc.mx) acc.mx = x[e];
            }
        } else {
            Stats4 st = {acc.sum, acc.mn, acc.mx, acc.nvalid};
            st = rare_nan4(st, x[0], x[1], x[2], x[3]);
            acc.sum = st.sum;
            acc.mn = st.mn;
            acc.mx = st.mx;
            acc.nvalid = st.nv;
        }
    }
    // Brackets and tail: straight-line predicated code; overflowing captures (bit 4*interval + sample of ovm) are
    // queued afterwards.
    unsigned ovm = 0;
#define KO_CAP_BRACKET(i)                                                                                        \
    cap_bracket<(1 << (4 * (i) + 0))>(x[0], fr.tA[i], fr.tB[i], acc.cA[i], acc.p[i], acc.pend[i], ovm);            \
    cap_bracket<(1 << (4 * (i) + 1))>(x[1], fr.tA[i], fr.tB[i], acc.cA[i], acc.p[i], acc.pend[i], ovm);            \
    cap_bracket<(1 << (4 * (i) + 2))>(x[2], fr.tA[i], fr.tB[i], acc.cA[i], acc.p[i], acc.pend[i], ovm);            \
    cap_bracket<(1 << (4 * (i) + 3))>(x[3], fr.tA[i], fr.tB[i], acc.cA[i], acc.p[i], acc.pend[i], ovm);
    if (NB > 0) {
        KO_CAP_BRACKET(0)
    }
    if (NB > 1) {
        KO_CAP_BRACKET(NB > 1 ? 1 : 0)
    }
#undef KO_CAP_BRACKET
    if (TAIL) {
        cap_tail<(1 << (4 * NB + 0))>(x[0], fr.tT, acc.p[NB], acc.pend[NB], ovm);
        cap_tail<(1 << (4 * NB + 1))>(x[1], fr.tT, acc.p[NB], acc.pend[NB], ovm);
        cap_tail<(1 << (4 * NB + 2))>(x[2], fr.tT, acc.p[NB], acc.pend[NB], ovm);
        cap_tail<(1 << (4 * NB + 3))>(x[3], fr.tT, acc.p[NB], acc.pend[NB], ovm);
    }
    while (ovm) {   // one round per set bit (a warp rarely has more than two or three in one lane)
        const int b = __ffs((int)ovm) - 1;
        ovm &= ovm - 1u;
        const int e = b & 3;
        const double xe = e == 0 ? x[0] : (e == 1 ? x[1] : (e == 2 ? x[2] : x[3]));
        pipe_overflow(c.ovn32, c.ovcnt32, c.ovq32, c.ovtag32, b >> 2, xe);
    }
    (void)lane;
}

template <int NB, bool TAIL>
__device__ __forceinline__ void pipe_stream(const PipePlan* pp, PAcc& acc, const PStreamCtx& c, const PipeSmem& sm,
                                            uint32_t& g, int stid, int lane) {
    PFast<NB, TAIL> fr;
#pragma unroll
    for (int i = 0; i < NB; ++i) {
        fr.tA[i] = pp->tA[i];
        fr.tB[i] = pp->tB[i];
    }
    fr.tT = TAIL ? pp->tA[NB] : 0.0;
    const double nan = __longlong_as_double(0x7FF8000000000000LL);
    const int L = pp->L, head = pp->head, body = pp->body, ntiles = pp->ntiles;
    const double* v = pp->v;
    if (stid == 0 && head) {
        const double x[4] = {v[0], nan, nan, nan};
        pipe_elems4<NB, TAIL>(x, acc, fr, c, lane);
    }
    if (stid == 32 && head + body < L) {   // first lane of another warp, so the two stragglers do not serialise
        const double x[4] = {v[L - 1], nan, nan, nan};
        pipe_elems4<NB, TAIL>(x, acc, fr, c, lane);
    }
    const uint32_t full0 = smem_u32(sm.bars() + B_RING_FULL), empty0 = smem_u32(sm.bars() + B_RING_EMPTY);
    const uint32_t ring0 = smem_u32(sm.ring()) + (uint32_t)stid * 16u;
    uint32_t s = g % P_STAGES, par = (g / P_STAGES) & 1u;
    const int nfull = body / P_TILE;   // whole tiles; a last partial one follows when body is not a multiple
    for (int t = 0; t < nfull; ++t) {
        { KO_PT0(); mbar_wait_addr(full0 + s * 8u, par); KO_PT1(16, stid == 0); }
        const uint32_t base = ring0 + s * (uint32_t)(P_TILE * 8);
        double x[4];
        lds_v2f64(base, x[0], x[1]);
        lds_v2f64(base + P_NS * 16, x[2], x[3]);
        // the samples are in registers: the stage goes back to the producer before they are worked on (a warp's
        // shared-memory loads and its barrier arrival are carried out in order)
        __syncwarp();
        if (lane == 0) mbar_arrive_addr(empty0 + s * 8u);
        pipe_elems4<NB, TAIL>(x, acc, fr, c, lane);
        if (++s == P_STAGES) {
            s = 0;
            par ^= 1u;
        }
    }
    if (nfull < ntiles) {
        const int rem2 = (body - nfull * P_TILE) >> 1;
        mbar_wait_addr(full0 + s * 8u, par);
        const uint32_t base = ring0 + s * (uint32_t)(P_TILE * 8);
        double x[4];
        x[0] = x[1] = x[2] = x[3] = nan;
        if (stid < rem2) lds_v2f64(base, x[0], x[1]);
        if (stid + P_NS < rem2) lds_v2f64(base + P_NS * 16, x[2], x[3]);
        __syncwarp();
        if (lane == 0) mbar_arrive_addr(empty0 + s * 8u);
        pipe_elems4<NB, TAIL>(x, acc, fr, c, lane);
    }
    g += (uint32_t)ntiles;
}

// Levels plan: per sample a branch-free binary search over the 31 ascending thresholds gives the number of
// thresholds <= x (NaN: 0), and the thread's private 16-bit counter of that gap is incremented.  Counter
// (gap g, stream thread t) lives at pool + (g * P_NS + t) * 2.
__device__ __forceinline__ unsigned lds_u16(uint32_t addr) {
    unsigned short c;
    asm volatile("ld.shared.u16 %0, [%1];" : "=h"(c) : "r"(addr) : "memory");
    return c;
}
__device__ __forceinline__ void sts_u16(uint32_t addr, unsigned c) {
    asm volatile("st.shared.u16 [%0], %1;" ::"r"(addr), "h"((unsigned short)c) : "memory");
}
__device__ __forceinline__ void pipe_levels4(const double (&x)[4], PAcc& acc, uint32_t thr32, uint32_t col32) {
    {
        const double t = (x[0] + x[1]) + (x[2] + x[3]);
        if (!rare4(t, x[0], x[1], x[2], x[3], acc.mn, acc.mx)) {
            acc.sum += t;
            acc.nvalid += 4;
        } else {   // a stale sample or a new record among the four: per sample, in line
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const bool ok = x[e] == x[e];
                acc.sum += ok ? x[e] : 0.0;
                acc.nvalid += ok ? 1u : 0u;
                acc.mn = fmin(acc.mn, x[e]);
                acc.mx = fmax(acc.mx, x[e]);
            }
        }
    }
    // ga = address of thr[number of thresholds <= x so far]: load, compare, predicated add per step
    uint32_t ga[4] = {thr32, thr32, thr32, thr32};
#define KO_LEVEL_STEP(STEP)                                                                                  \
    _Pragma("unroll") for (int e = 0; e < 4; ++e) {                                                          \
        const double t = lds_f64(ga[e] + (uint32_t)((STEP)-1) * 8u);                                         \
        asm("{\n.reg .pred p;\nsetp.ge.f64 p, %1, %2;\n@p add.u32 %0, %0, %3;\n}"                            \
            : "+r"(ga[e]) : "d"(x[e]), "d"(t), "n"((STEP)*8));                                               \
    }
    KO_LEVEL_STEP(16)
    KO_LEVEL_STEP(8)
    KO_LEVEL_STEP(4)
    KO_LEVEL_STEP(2)
    KO_LEVEL_STEP(1)
#undef KO_LEVEL_STEP
#pragma unroll
    for (int e = 0; e < 4; ++e) {   // one after another: two samples of a thread often share a gap
        const uint32_t addr = col32 + (ga[e] - thr32) * (uint32_t)(P_NS * 2 / 8);
        sts_u16(addr, lds_u16(addr) + 1u);
    }
}

__device__ __forceinline__ void pipe_stream_levels(const PipePlan* pp, PAcc& acc, uint32_t pool32, const PipeSmem& sm,
                                                uint32_t& g, int stid, int lane) {
    const uint32_t thr32 = smem_u32(&pp->thr[0]);
    const uint32_t col32 = pool32 + (uint32_t)stid * 2u;
#pragma unroll 4
    for (int b = 0; b <= P_NTHR; ++b) sts_u16(col32 + (uint32_t)b * (P_NS * 2), 0u);
    const double nan = __longlong_as_double(0x7FF8000000000000LL);
    const int L = pp->L, head = pp->head, body = pp->body, ntiles = pp->ntiles;
    const double* v = pp->v;
    if (stid == 0 && head) {
        const double x[4] = {v[0], nan, nan, nan};
        pipe_levels4(x, acc, thr32, col32);
    }
    if (stid == 32 && head + body < L) {
        const double x[4] = {v[L - 1], nan, nan, nan};
        pipe_levels4(x, acc, thr32, col32);
    }
    const uint32_t full0 = smem_u32(sm.bars() + B_RING_FULL), empty0 = smem_u32(sm.bars() + B_RING_EMPTY);
    const uint32_t ring0 = smem_u32(sm.ring()) + (uint32_t)stid * 16u;
    uint32_t s = g % P_STAGES, par = (g / P_STAGES) & 1u;
    const int nfull = body / P_TILE;   // whole tiles, then the partial one
    for (int t = 0; t < nfull; ++t) {
        mbar_wait_addr(full0 + s * 8u, par);
        const uint32_t base = ring0 + s * (uint32_t)(P_TILE * 8);
        double x[4];
        lds_v2f64(base, x[0], x[1]);
        lds_v2f64(base + P_NS * 16, x[2], x[3]);
        // the samples are in registers: the stage goes back to the producer before they are worked on (a warp's
        // shared-memory loads and its barrier arrival are carried out in order)
        __syncwarp();
        if (lane == 0) mbar_arrive_addr(empty0 + s * 8u);
        pipe_levels4(x, acc, thr32, col32);
        if (++s == P_STAGES) {
            s = 0;
            par ^= 1u;
        }
    }
    if (nfull < ntiles) {
        const int rem2 = (body - nfull * P_TILE) >> 1;
        mbar_wait_addr(full0 + s * 8u, par);
        const uint32_t base = ring0 + s * (uint32_t)(P_TILE * 8);
        double x[4];
        x[0] = x[1] = x[2] = x[3] = nan;
        if (stid < rem2) lds_v2f64(base, x[0], x[1]);
        if (stid + P_NS < rem2) lds_v2f64(base + P_NS * 16, x[2], x[3]);
        __syncwarp();
        if (lane == 0) mbar_arrive_addr(empty0 + s * 8u);
        pipe_levels4(x, acc, thr32, col32);
    }
    g += (uint32_t)ntiles;
}

// ------------------------------------------------------------------------------------------ finish
// The finish warps turn the captured candidates of one series into its order statistics.
//
// Selection without shared-memory atomics on the hot path (a warp-wide shared atomic costs ~64 cycles): the
// bracket of an interval is cut into 64 sub-brackets, linear in key space; every finish thread counts its share of
// the candidates into PRIVATE byte counters (plain load / add / store), the counters are summed with dp4a, the
// sub-bracket holding the wanted rank is found with a warp scan, its few dozen candidates are collected, and
// the exact order statistic among them is found by rank counting.
//
// The capturing intervals of a series (1..3) are worked on side by side: the P_NF finish threads split into one
// group per interval, each group synchronising on its own named barrier.  Everything the groups need to agree on
// before that (valid count, ranks, which bracket holds which rank) is computed redundantly by every thread from
// the totals the stream warps left in PipeResult, so no barrier precedes the first scan.
__device__ __forceinline__ unsigned lds_u32(uint32_t addr) {
    unsigned c;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(c) : "r"(addr) : "memory");
    return c;
}
__device__ __forceinline__ void sts_u32(uint32_t addr, unsigned c) {
    asm volatile("st.shared.u32 [%0], %1;" ::"r"(addr), "r"(c) : "memory");
}
__device__ __forceinline__ void group_sync(int grp, int nthreads) {
    asm volatile("bar.sync %0, %1;" ::"r"(3 + grp), "r"(nthreads) : "memory");
}

#ifdef KO_PHASE_TIMING
#define KO_FT(slot)                                                         \
    do {                                                                    \
        if (blockIdx.x == 0 && ftid == 0) {                                 \
            const unsigned long long _t = clock64();                        \
            g_phase[slot] += _t - _ft;                                      \
            _ft = _t;                                                       \
        }                                                                   \
    } while (0)
#else
#define KO_FT(slot) do { } while (0)
#endif

// What every finish thread knows about its own interval after the prologue.
struct FinIv {
    int iv;                      // interval index
    unsigned long long kA;       // bracket [kA, kB) in key space
    int ish;                     // (key - kA) >> ish is the sub-bracket, in [0, 64)
    int j0, j1;                  // quantiles (ascending positions) to select in this interval, -1 = none
    unsigned rho0, rho1;         // their ranks inside the bracket (1-based)
};

// Visit the candidates of interval fi.iv that belong to group thread gt: the private slots of the stream threads
// gt, gt + G, ... taken two columns at a time (shared-memory latency is > 100 cycles while the stream warps and
// the TMA engine keep the banks busy, so the next row is fetched before the current one is processed), then a
// share of the overflow queue.
template <int NG, typename F4, typename F1>
__device__ __forceinline__ void fin_for_rows(uint32_t pool32, uint32_t ovq32, const PipePlan* pp,
                                             const PipeResult* res, int iv, int gt, F4 frow, F1 fone) {
    constexpr int G = fin_group_size(NG);
    constexpr int NCOL = (P_NS + G - 1) / G;      // columns of the busiest thread (stream threads gt, gt + G, ...)
    constexpr int NPAIR = (NCOL + 1) / 2;         // taken two columns and two rows per step
    const uint32_t base0 = pool32 + (uint32_t)(((size_t)pp->slot0[iv] * P_NS + gt) * 8);
#pragma unroll
    for (int pr = 0; pr < NPAIR; ++pr) {
        const int c0 = gt + 2 * pr * G, c1 = c0 + G;
        const uint32_t base = base0 + (uint32_t)(2 * pr * G) * 8u;
        if ((2 * pr + 1) * G >= P_NS) {   // (compile time) an odd column left over: four of its rows per step
            const int n0 = c0 < P_NS ? res->pcount[iv][c0 < P_NS ? c0 : 0] : 0;
            double y0 = 0.0, y1 = 0.0, y2 = 0.0, y3 = 0.0;
            if (0 < n0) y0 = lds_f64(base);
            if (1 < n0) y1 = lds_f64(base + P_NS * 8);
            if (2 < n0) y2 = lds_f64(base + 2 * P_NS * 8);
            if (3 < n0) y3 = lds_f64(base + 3 * P_NS * 8);
            for (int s = 0; s < n0; s += 4) {
                const double x0 = y0, x1 = y1, x2 = y2, x3 = y3;
                const uint32_t nrow = base + (uint32_t)(s + 4) * (P_NS * 8);
                if (s + 4 < n0) y0 = lds_f64(nrow);
                if (s + 5 < n0) y1 = lds_f64(nrow + P_NS * 8);
                if (s + 6 < n0) y2 = lds_f64(nrow + 2 * P_NS * 8);
                if (s + 7 < n0) y3 = lds_f64(nrow + 3 * P_NS * 8);
                frow(x0, s < n0, x1, s + 1 < n0, x2, s + 2 < n0, x3, s + 3 < n0);
            }
            continue;
        }
        const int n0 = c0 < P_NS ? res->pcount[iv][c0 < P_NS ? c0 : 0] : 0;
        const int n1 = c1 < P_NS ? res->pcount[iv][c1 < P_NS ? c1 : 0] : 0;
        const int nmax = max(n0, n1);
        double y0 = 0.0, y1 = 0.0, y2 = 0.0, y3 = 0.0;
        if (0 < n0) y0 = lds_f64(base);
        if (0 < n1) y1 = lds_f64(base + G * 8);
        if (1 < n0) y2 = lds_f64(base + P_NS * 8);
        if (1 < n1) y3 = lds_f64(base + P_NS * 8 + G * 8);
        for (int s = 0; s < nmax; s += 2) {
            const double x0 = y0, x1 = y1, x2 = y2, x3 = y3;
            const uint32_t nrow = base + (uint32_t)(s + 2) * (P_NS * 8);
            if (s + 2 < n0) y0 = lds_f64(nrow);
            if (s + 2 < n1) y1 = lds_f64(nrow + G * 8);
            if (s + 3 < n0) y2 = lds_f64(nrow + P_NS * 8);
            if (s + 3 < n1) y3 = lds_f64(nrow + P_NS * 8 + G * 8);
            frow(x0, s < n0, x1, s < n1, x2, s + 1 < n0, x3, s + 1 < n1);
        }
    }
    // the overflow queue is shared by the intervals of the series: entries carry their interval as a tag
    const int ovn = (int)min(res->ovn, (unsigned)P_OVQ);
    const uint32_t tag32 = smem_u32(&res->ovtag[0]);
    for (int q = gt; q < ovn; q += G)
        if ((int)lds_u8(tag32 + (uint32_t)q) == iv) fone(lds_f64(ovq32 + (uint32_t)q * 8u));
}

// The group phase: NG capturing intervals, G = P_NF / NG threads each.  `use` is uniform inside a group.
//
// Per wanted rank the candidates are narrowed to one of 64 sub-brackets by counting; a sub-bracket that still
// holds more than P_LIST candidates (skewed data, heavy tie groups) is cut into 64 again, until the candidates
// fit the list or the sub-bracket is a single key (then that key is the answer).
template <int NG>
__device__ __forceinline__ void fin_groups(const PipePlan* pp, const PipeResult* res, PipeSel* sel,
                                           PipeSel::Out* so, uint32_t pool32, uint32_t ovq32, const FinIv& fi,
                                           int grp, int gt, int ftid) {
    constexpr int G = fin_group_size(NG);
    constexpr int GW = G / 4;            // 32-bit words of one sub-bracket's counters inside the group
    const bool use = fi.j0 >= 0;
    const uint32_t cnt32 = smem_u32(&sel->cnt8[0][0]);
    const uint32_t col = cnt32 + (uint32_t)ftid;
    const uint32_t dummy = smem_u32(&sel->dummy[0][0]) + (uint32_t)ftid;
#ifdef KO_PHASE_TIMING
    unsigned long long _ft = clock64();
#endif
    // totals of the group's byte counters per sub-bracket (thread b sums sub-bracket b with dp4a), then, with the
    // counters dead, clear them for the next use while the group's first warp finds the sub-bracket of each
    // wanted rank by a scan over the 64 totals (two per lane).  Three group barriers.
    auto totals_and_walk = [&](int ja, unsigned rhoa, int jb, unsigned rhob) {
        group_sync(grp, G);
        if (gt < 64) {
            const uint32_t w = cnt32 + (uint32_t)(gt * P_CROW + grp * G);
            unsigned t = 0;
#pragma unroll 8
            for (int x = 0; x < GW; ++x) t = __dp4a(lds_u32(w + 4u * x), 0x01010101u, t);
            sel->tot[grp][gt] = t;
        }
        group_sync(grp, G);
#pragma unroll
        for (int t = 0; t < 16; ++t) {
            const int idx = gt + t * G;            // 64 * GW words, G threads
            const int bin = idx / GW, wd = idx - bin * GW;
            sts_u32(cnt32 + (uint32_t)(bin * P_CROW + grp * G + wd * 4), 0u);
        }
        if (gt < 32) {
            const int lane = gt;
            const unsigned h0 = sel->tot[grp][2 * lane], h1 = sel->tot[grp][2 * lane + 1];
            const unsigned s = h0 + h1;
            unsigned incl = s;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const unsigned t = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += t;
            }
            const unsigned excl = incl - s;
#pragma unroll
            for (int jj = 0; jj < 2; ++jj) {
                const int j = jj ? jb : ja;
                const unsigned rho = jj ? rhob : rhoa;
                if (j >= 0) {
                    const bool m0 = excl < rho && rho <= excl + h0;
                    const bool m1 = excl + h0 < rho && rho <= incl;
                    const unsigned hit = __ballot_sync(0xffffffffu, m0 || m1);
                    if (m0 || m1) {
                        sel->bin[j] = 2 * lane + (m1 ? 1 : 0);
                        sel->rho[j] = rho - (m1 ? excl + h0 : excl);
                        sel->nlist[j] = m1 ? h1 : h0;
                    }
                    if (hit == 0u && lane == 0) {   // cannot happen while the counts are consistent
                        atomicOr(&so->fail, 1 << j);
                        sel->bin[j] = 0;
                        sel->nlist[j] = 0;
                        KO_COUNT(5, 0);
                    }
                }
            }
        }
        group_sync(grp, G);
    };
    unsigned long long lo[2] = {0, 0};   // sub-bracket [lo, lo + 2^sh) of each wanted rank
    int sh[2] = {0, 0};
    bool direct[2] = {false, false};
    if (use) {
        const unsigned long long kA = fi.kA;
        const int ish = fi.ish;
        // ---- first cut: private byte counters per sub-bracket (zero on entry), every candidate of the interval
        {
            auto addr_of = [&](double xv, bool valid, int which) -> uint32_t {
                const unsigned long long k = ko_okey(xv + 0.0);
                return valid ? col + (uint32_t)((k - kA) >> ish) * (uint32_t)P_CROW : dummy + (uint32_t)which * P_NF;
            };
            fin_for_rows<NG>(pool32, ovq32, pp, res, fi.iv, gt,
                             [&](double x0, bool v0, double x1, bool v1, double x2, bool v2, double x3, bool v3) {
                                 // four counters are loaded together and stored in order; a later sample that hits
                                 // the counter of an earlier one carries that one's increment
                                 const uint32_t a0 = addr_of(x0, v0, 0), a1 = addr_of(x1, v1, 1);
                                 const uint32_t a2 = addr_of(x2, v2, 2), a3 = addr_of(x3, v3, 3);
                                 const unsigned c0 = lds_u8(a0), c1 = lds_u8(a1), c2 = lds_u8(a2), c3 = lds_u8(a3);
                                 const unsigned d1 = c1 + (a1 == a0 ? 1u : 0u);
                                 const unsigned d2 = c2 + (a2 == a0 ? 1u : 0u) + (a2 == a1 ? 1u : 0u);
                                 const unsigned d3 = c3 + (a3 == a0 ? 1u : 0u) + (a3 == a1 ? 1u : 0u) + (a3 == a2 ? 1u : 0u);
                                 sts_u8(a0, c0 + 1);
                                 sts_u8(a1, d1 + 1);
                                 sts_u8(a2, d2 + 1);
                                 sts_u8(a3, d3 + 1);
                             },
                             [&](double xv) {
                                 const uint32_t a0 = addr_of(xv, true, 0);
                                 sts_u8(a0, lds_u8(a0) + 1);
                             });
        }
        KO_FT(24);
        totals_and_walk(fi.j0, fi.rho0, fi.j1, fi.rho1);
        KO_FT(25);
        // ---- further cuts, one wanted rank at a time, only while its sub-bracket holds too many candidates
#pragma unroll
        for (int jj = 0; jj < 2; ++jj) {
            const int j = jj ? fi.j1 : fi.j0;
            if (j < 0) continue;
            lo[jj] = kA + ((unsigned long long)sel->bin[j] << ish);
            sh[jj] = ish;
            while (sel->nlist[j] > (unsigned)P_LIST && sh[jj] > 0) {   // uniform in the group
                const unsigned long long l0 = lo[jj];
                const int wsh = sh[jj], bsh = wsh > 6 ? wsh - 6 : 0;
                const unsigned rho = sel->rho[j];
                auto bump = [&](double xv) {
                    const unsigned long long d = ko_okey(xv + 0.0) - l0;
                    if ((d >> wsh) == 0) {
                        const uint32_t ad = col + (uint32_t)(d >> bsh) * (uint32_t)P_CROW;
                        sts_u8(ad, lds_u8(ad) + 1);
                    }
                };
                fin_for_rows<NG>(pool32, ovq32, pp, res, fi.iv, gt,
                                 [&](double x0, bool v0, double x1, bool v1, double x2, bool v2, double x3, bool v3) {
                                     if (v0) bump(x0);
                                     if (v1) bump(x1);
                                     if (v2) bump(x2);
                                     if (v3) bump(x3);
                                 },
                                 bump);
                totals_and_walk(j, rho, -1, 0u);
                lo[jj] = l0 + ((unsigned long long)sel->bin[j] << bsh);
                sh[jj] = bsh;
            }
            if (sel->nlist[j] > (unsigned)P_LIST) {   // a single key holds the rank
                direct[jj] = true;
                if (gt == 0) so->val[j] = ko_key2d(lo[jj]);
            }
        }
        KO_FT(26);
        // ---- collect the candidates of each wanted sub-bracket (at most two ranks per interval)
        {
            const int ja = direct[0] ? -1 : fi.j0, jb = direct[1] ? -1 : fi.j1;
            const unsigned long long la = lo[0], lb = lo[1];
            const int sa = sh[0], sb = sh[1];
            const uint32_t cnt0 = smem_u32(&so->list_cnt[ja < 0 ? 0 : ja]), cnt1 = smem_u32(&so->list_cnt[jb < 0 ? 0 : jb]);
            const uint32_t l0 = smem_u32(&sel->list[ja < 0 ? 0 : ja][0]), l1 = smem_u32(&sel->list[jb < 0 ? 0 : jb][0]);
            // the sub-brackets as ranges of doubles (doubles order like their keys): [dla, dha) and [dlb, dhb); an
            // upper end beyond +inf is a NaN, which no comparison passes
            const unsigned long long ha = la + (1ull << sa), hb = lb + (1ull << sb);
            const double qnan = __longlong_as_double(0x7FF8000000000000LL);
            const double dla = ko_key2d(la), dlb = ko_key2d(lb);
            const double dha = (ha < la || ha > KO_KEY_PINF) ? qnan : ko_key2d(ha);
            const double dhb = (hb < lb || hb > KO_KEY_PINF) ? qnan : ko_key2d(hb);
            auto take = [&](double xv) {
                const double xc = xv + 0.0;
                if (ja >= 0 && xc >= dla && !(xc >= dha)) {
                    unsigned pos;
                    asm volatile("atom.shared.add.u32 %0, [%1], 1;" : "=r"(pos) : "r"(cnt0) : "memory");
                    if (pos < (unsigned)P_LIST) sts_f64(l0 + pos * 8u, xc);
                }
                if (jb >= 0 && xc >= dlb && !(xc >= dhb)) {
                    unsigned pos;
                    asm volatile("atom.shared.add.u32 %0, [%1], 1;" : "=r"(pos) : "r"(cnt1) : "memory");
                    if (pos < (unsigned)P_LIST) sts_f64(l1 + pos * 8u, xc);
                }
            };
            if (ja >= 0 || jb >= 0)
                fin_for_rows<NG>(pool32, ovq32, pp, res, fi.iv, gt,
                                 [&](double x0, bool v0, double x1, bool v1, double x2, bool v2, double x3, bool v3) {
                                     if (v0) take(x0);
                                     if (v1) take(x1);
                                     if (v2) take(x2);
                                     if (v3) take(x3);
                                 },
                                 take);
        }
        group_sync(grp, G);
        KO_FT(27);
        // ---- exact order statistic inside each short list by rank counting (doubles compare like their keys).
        // A list holds a few dozen candidates, i.e. work for two warps: the second rank of the interval is taken by
        // the threads counted from the other end of the group, so that both ranks are counted at the same time.
#pragma unroll
        for (int jj = 0; jj < 2; ++jj) {
            const int j = jj ? fi.j1 : fi.j0;
            if (j >= 0 && !direct[jj]) {
                const int n = (int)min(so->list_cnt[j], (unsigned)P_LIST);
                const unsigned rho = sel->rho[j];
                const uint32_t l32 = smem_u32(&sel->list[j][0]);
                for (int idx = jj ? G - 1 - gt : gt; idx < n; idx += G) {
                    const double e = lds_f64(l32 + (uint32_t)idx * 8u);
                    unsigned lt = 0, le = 0;
#pragma unroll 4
                    for (int m = 0; m < n; ++m) {
                        const double v = lds_f64(l32 + (uint32_t)m * 8u);
                        asm("{\n.reg .pred p, q;\nsetp.lt.f64 p, %2, %3;\nsetp.le.f64 q, %2, %3;\n"
                            "@p add.u32 %0, %0, 1;\n@q add.u32 %1, %1, 1;\n}"
                            : "+r"(lt), "+r"(le) : "d"(v), "d"(e));
                    }
                    if (lt < rho && rho <= le) so->val[j] = e;   // all writers hold the same value
                }
            }
        }
        KO_FT(28);
    }
}

// Run by one thread (lane 0 of the prologue warp) once all totals of the series are in *res.
__device__ __noinline__ void fin_prologue(const PipePlan* pp, PipeResult* res, int nq, double q0, double q1, double q2,
                                          double q3, int wide) {
    (void)wide;
    FinPro* pro = &res->pro;
    const int ni = pp->ni, nb = pp->nb, qsel = pp->qsel;
    const double nan = __longlong_as_double(0x7FF8000000000000LL);
    const unsigned n = res->t_nvalid;
    double s = 0.0, mn = INFINITY, mx = -INFINITY;
    for (int w = 0; w < P_SW; ++w) {   // in warp order: reproducible
        s += res->sum[w];
        mn = fmin(mn, res->mn[w]);
        mx = fmax(mx, res->mx[w]);
    }
    pro->n = n;
    pro->sum = s;
    pro->mn = mn;
    pro->mx = mx;
    unsigned cAt[P_NI], nin[P_NI];
    unsigned long long kA[P_NI], kB[P_NI];
    bool capm[P_NI];
    const bool ovbad = res->ovn > (unsigned)P_OVQ;
    int ncapi = 0;
#pragma unroll
    for (int i = 0; i < P_NI; ++i) {
        cAt[i] = nin[i] = 0;
        kA[i] = 0;
        kB[i] = 1;
        capm[i] = false;
        int ish = 0, ord = -1;
        if (i < ni) {
            nin[i] = res->t_ncap[i] + res->ovcnt[i] + res->t_cin[i];
            cAt[i] = (i < nb) ? res->t_cA[i] : n - nin[i];   // upper tail: everything outside is below
            kA[i] = pp->kA[i];
            kB[i] = pp->kB[i];
            if (kB[i] > KO_KEY_PINF) kB[i] = ko_okey(mx + 0.0) + 1;   // upper tail: nothing lies above the maximum
            const unsigned long long span = kB[i] - kA[i] - 1;        // kB > kA
            const int bits = span ? 64 - __clzll((long long)span) : 0;
            ish = bits > 6 ? bits - 6 : 0;
            capm[i] = pp->mode[i] == M_CAPTURE;
            if (capm[i]) ord = ncapi++;
        }
        pro->kA[i] = kA[i];
        pro->ish[i] = ish;
        pro->ord[i] = ord;
    }
    pro->ncapi = ncapi;
    int failmask = 0;
    const double qs[NI_MAX] = {q0, q1, q2, q3};
#pragma unroll
    for (int j = 0; j < NI_MAX; ++j) {
        double dval = nan;
        int act = -1;
        unsigned rho = 0;
        if (j < nq && n > 0 && (qsel < 0 || j == qsel)) {
            const int i = pp->q_iv[j];
            const unsigned r = nearest_rank_dev(n, qs[j]);
            unsigned cA = cAt[0], nn = nin[0];
            unsigned long long ka = kA[0], kb = kB[0];
            bool cm = capm[0];
#pragma unroll
            for (int t = 1; t < P_NI; ++t)
                if (i == t) {
                    cA = cAt[t];
                    nn = nin[t];
                    ka = kA[t];
                    kb = kB[t];
                    cm = capm[t];
                }
            if (r <= cA || r > cA + nn) {
                failmask |= 1 << j;       // the bracket missed the rank
                KO_COUNT(3, wide);
            } else if (!cm || kb - ka == 1) {
                dval = ko_key2d(ka);      // a single value: answered by the counts
            } else if (ovbad) {
                failmask |= 1 << j;       // more candidates than shared memory holds
                KO_COUNT(4, wide);
            } else {
                act = i;
                rho = r - cA;
            }
        }
        pro->dval[j] = dval;
        pro->act_iv[j] = act;
        pro->rho[j] = rho;
    }
    pro->failmask = failmask;
}

// Returns (meaningful on finish thread 0) the mask of quantiles that a later attempt has to answer; 0 = all done.
__device__ __forceinline__ int pipe_finish(const SeriesArgs& a, const PipePlan* pp, PipeResult* res, PipeSel* sel,
                                            PipeSel::Out* so, const unsigned char* pool, const double* ovq, int ftid,
                                            int fwarp, int lane) {
    (void)fwarp;
    const int ni = pp->ni, nq = a.nq, nb = pp->nb, qsel = pp->qsel;
    const double nan = __longlong_as_double(0x7FF8000000000000LL);
#ifdef KO_PHASE_TIMING
    unsigned long long _ft = clock64();
#endif
    // ---- what the prologue warp worked out (fin_prologue)
    const FinPro* pro = &res->pro;
    const unsigned n = pro->n;
    if (ftid == 32) {   // the plain statistics do not depend on the selection: written here, off the critical path
        const int sid = pp->sid;
        const double s = pro->sum;
        if (a.out_sum) a.out_sum[sid] = s;
        if (a.out_mean) a.out_mean[sid] = n > 0 ? s / (double)n : 0.0;
        if (a.out_min) a.out_min[sid] = n > 0 ? pro->mn + 0.0 : nan;
        if (a.out_max) a.out_max[sid] = n > 0 ? pro->mx + 0.0 : nan;
        if (a.out_last) a.out_last[sid] = pp->last;
        if (a.out_count) a.out_count[sid] = (int64_t)n;
    }
    unsigned long long kA[P_NI];
    int ish[P_NI], ord[P_NI];
#pragma unroll
    for (int i = 0; i < P_NI; ++i) {
        kA[i] = pro->kA[i];
        ish[i] = pro->ish[i];
        ord[i] = pro->ord[i];
    }
    const int ncapi = pro->ncapi;
    int failmask = pro->failmask;   // quantiles this attempt cannot answer
    double dval[NI_MAX];
    int act_iv[NI_MAX];
    unsigned rho[NI_MAX];
#pragma unroll
    for (int j = 0; j < NI_MAX; ++j) {
        dval[j] = pro->dval[j];
        act_iv[j] = pro->act_iv[j];
        rho[j] = pro->rho[j];
    }
    (void)ni;
    (void)nb;
    (void)lane;
    KO_FT(13);
    bool any_act = false;
#pragma unroll
    for (int j = 0; j < NI_MAX; ++j) any_act = any_act || act_iv[j] >= 0;
    if (any_act) {
        // my group and its interval
        const int gsz = fin_group_size(ncapi);
        const int grp = ftid / gsz, gt = ftid - grp * gsz;
        const bool member = grp < ncapi;   // three groups of 64 leave the last warps without one
        FinIv fi;
        fi.iv = 0;
        fi.kA = 0;
        fi.ish = 0;
        fi.j0 = fi.j1 = -1;
        fi.rho0 = fi.rho1 = 0;
#pragma unroll
        for (int i = 0; i < P_NI; ++i)
            if (ord[i] == grp) {
                fi.iv = i;
                fi.kA = kA[i];
                fi.ish = ish[i];
            }
#pragma unroll
        for (int j = 0; j < NI_MAX; ++j)
            if (act_iv[j] >= 0 && act_iv[j] == fi.iv) {
                if (fi.j0 < 0) {
                    fi.j0 = j;
                    fi.rho0 = rho[j];
                } else {
                    fi.j1 = j;
                    fi.rho1 = rho[j];
                }
            }
        const uint32_t pool32 = smem_u32(pool), ovq32 = smem_u32(ovq);
        if (!member) {
        } else if (ncapi == 1) {
            fin_groups<1>(pp, res, sel, so, pool32, ovq32, fi, grp, gt, ftid);
        } else if (ncapi == 2) {
            fin_groups<2>(pp, res, sel, so, pool32, ovq32, fi, grp, gt, ftid);
        } else {
            fin_groups<3>(pp, res, sel, so, pool32, ovq32, fi, grp, gt, ftid);
        }
    }
    KO_FT(14);
    fin_sync();
    KO_FT(7);
    if (ftid == 0) {
        failmask |= so->fail;
        double out[NI_MAX];
#pragma unroll
        for (int j = 0; j < NI_MAX; ++j) {
            out[j] = act_iv[j] >= 0 ? so->val[j] : dval[j];
            if (act_iv[j] >= 0 && !(out[j] == out[j])) {
                if (!((failmask >> j) & 1)) KO_COUNT(6, a.wide);
                failmask |= 1 << j;   // no candidate carried the rank
            }
        }
        if (a.out_quant) {   // the quantiles that failed are written by a later attempt
            const int sid = pp->sid;
#pragma unroll
            for (int j = 0; j < NI_MAX; ++j)
                if (j < nq && (qsel < 0 || j == qsel) && !((failmask >> j) & 1))
                    a.out_quant[(int64_t)sid * nq + a.qcol[j]] = out[j];
        }
    }
    return failmask;
}

// Finish of a levels plan: total the per-thread gap counters, then every rank must fall on a listed level, i.e.
// (#valid samples < v) < rank <= (#valid samples <= v); otherwise the series goes to the slow path.
__device__ __forceinline__ int pipe_finish_levels(const SeriesArgs& a, const PipePlan* pp, PipeResult* res, PipeSel* sel,
                                                PipeSel::Out* so, const unsigned char* pool, int ftid, int fwarp,
                                                int lane) {
    const int nq = a.nq;
    const double nan = __longlong_as_double(0x7FF8000000000000LL);
    const uint32_t pool32 = smem_u32(pool);
    for (int b = fwarp; b <= P_NTHR; b += P_FW) {
        unsigned t = 0;
#pragma unroll
        for (int wd = 0; wd < P_NS / 64; ++wd) {
            unsigned w;
            asm volatile("ld.shared.u32 %0, [%1];" : "=r"(w) : "r"(pool32 + (uint32_t)(b * P_NS * 2 + (wd * 32 + lane) * 4)));
            t += (w & 0xFFFFu) + (w >> 16);
        }
        t = warp_sum_u(t);
        if (lane == 0) sel->tot[0][b] = t;
    }
    fin_sync();
    if (ftid == 0) {
        double s = 0.0, mn = INFINITY, mx = -INFINITY;
        unsigned n = 0;
        for (int w = 0; w < P_SW; ++w) {
            s += res->sum[w];
            mn = fmin(mn, res->mn[w]);
            mx = fmax(mx, res->mx[w]);
        }
        n = res->t_nvalid;
        unsigned tot = 0;
        for (int b = 0; b <= P_NTHR; ++b) tot += sel->tot[0][b];
        // stale samples (and the padding of the last tile) all sit in gap 0
        unsigned cum = sel->tot[0][0] - (tot - n);   // valid samples below thr[0]
        int fail = 0;
        unsigned rk[NI_MAX];
        double val[NI_MAX];
#pragma unroll
        for (int j = 0; j < NI_MAX; ++j) {
            rk[j] = (j < nq && n > 0 && (pp->qsel < 0 || j == pp->qsel)) ? nearest_rank_dev(n, a.q[j]) : 0u;
            val[j] = nan;
        }
        const int nlev = pp->nlev;
        for (int d = 0; d < nlev; ++d) {
            const unsigned below = cum;                       // < v_d
            const unsigned upto = cum + sel->tot[0][2 * d + 1];   // <= v_d
#pragma unroll
            for (int j = 0; j < NI_MAX; ++j)
                if (rk[j] > below && rk[j] <= upto) {
                    val[j] = pp->thr[2 * d];
                    rk[j] = 0u;
                }
            cum = upto + sel->tot[0][2 * d + 2];             // < v_{d+1}
        }
#pragma unroll
        for (int j = 0; j < NI_MAX; ++j)
            if (rk[j] != 0u) {
                fail |= 1 << j;   // the rank fell on a value that is not listed
                KO_COUNT(7, a.wide);
            }
        so->fail = fail;
        {
            const int sid = pp->sid;
            if (a.out_sum) a.out_sum[sid] = s;
            if (a.out_mean) a.out_mean[sid] = n > 0 ? s / (double)n : 0.0;
            if (a.out_min) a.out_min[sid] = n > 0 ? mn + 0.0 : nan;
            if (a.out_max) a.out_max[sid] = n > 0 ? mx + 0.0 : nan;
            if (a.out_last) a.out_last[sid] = pp->last;
            if (a.out_count) a.out_count[sid] = (int64_t)n;
            if (a.out_quant) {
#pragma unroll
                for (int j = 0; j < NI_MAX; ++j)
                    if (j < nq && (pp->qsel < 0 || j == pp->qsel) && !((fail >> j) & 1))
                        a.out_quant[(int64_t)sid * nq + a.qcol[j]] = val[j];
            }
        }
    }
    fin_sync();
    return so->fail;
}

// ------------------------------------------------------------------------------------------ kernel
__global__ void __launch_bounds__(P_THREADS, 1) series_pipe(SeriesArgs a) {
    extern __shared__ __align__(128) unsigned char psmem[];
    PipeSmem sm;
    sm.base = psmem;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (tid == 0) {
        for (int s = 0; s < P_STAGES; ++s) {
            mbar_init(sm.bars() + B_RING_FULL + s, 1);
            mbar_init(sm.bars() + B_RING_EMPTY + s, P_SW);
        }
        for (int s = 0; s < P_PLANS; ++s) {
            mbar_init(sm.bars() + B_PLAN_FULL + s, 1);
            mbar_init(sm.bars() + B_PLAN_FREE + s, P_SW + 3);   // stream warps + finish + producer + prologue
        }
        for (int s = 0; s < 2; ++s) {
            mbar_init(sm.bars() + B_POOL_FULL + s, 1);      // the prologue warp, after ...
            mbar_init(sm.bars() + B_HANDOFF + s, P_SW);     // ... every stream warp has handed off
            mbar_init(sm.bars() + B_POOL_FREE + s, 1);
            PipeResult* r0 = sm.res(s);
            r0->t_nvalid = 0;
            r0->ovn = 0;
            for (int i = 0; i < P_NI; ++i) r0->ovcnt[i] = r0->t_cA[i] = r0->t_cin[i] = r0->t_ncap[i] = 0;
        }
        {
            PipeSel* sl = sm.sel();
            for (int o = 0; o < 2; ++o) {
                sl->o[o].fail = 0;
                for (int j = 0; j < NI_MAX; ++j) {
                    sl->o[o].list_cnt[j] = 0;
                    sl->o[o].val[j] = __longlong_as_double(0x7FF8000000000000LL);
                }
            }
        }
        mbar_fence_init();
    }
    {   // the finish warps' byte counters start (and are left) zero
        uint32_t* w = reinterpret_cast<uint32_t*>(&sm.sel()->cnt8[0][0]);
        for (int t = tid; t < 64 * P_CROW / 4; t += P_THREADS) w[t] = 0;
    }
    __syncthreads();
    const long long n_long = (long long)a.counters[a.work_cnt] * (a.wide ? a.nq : 1);

    if (warp < P_PW) {
        // ================================================================ PLAN LOADER
        // Work items come from a global cursor (reset by series_plan); their plans, written by series_plan, are
        // fetched into the plan slots with bulk copies.
        if (lane == 0) {
            const long long hi = n_long < a.w_begin + a.plan_cap ? n_long : a.w_begin + a.plan_cap;
            for (uint32_t k = 0;; ++k) {
                const uint32_t slot = k & (P_PLANS - 1);
                mbar_wait_sleep(sm.bars() + B_PLAN_FREE + slot, ((k >> P_PLAN_LOG) & 1u) ^ 1u);
                // items are handed out one at a time (series differ in cost: levels plans, second attempts)
                const long long w = a.w_begin + (long long)atomicAdd(&a.counters[a.work_cur], 1);
                PipePlan* pp = sm.plan(slot);
                if (w >= hi) {
                    pp->sid = -1;
                    mbar_arrive(sm.bars() + B_PLAN_FULL + slot);
                    break;
                }
                mbar_expect_tx(sm.bars() + B_PLAN_FULL + slot, (uint32_t)sizeof(PipePlan));
                tma_bulk_g2s(pp, reinterpret_cast<const PipePlan*>(a.plans) + (w - a.w_begin),
                             (uint32_t)sizeof(PipePlan), sm.bars() + B_PLAN_FULL + slot);
            }
        }
    } else if (warp == P_PW) {
        // ================================================================ PRODUCER
        if (lane == 0) {
            uint32_t g = 0;
            for (uint32_t k = 0;; ++k) {
                const uint32_t slot = k & (P_PLANS - 1);
                mbar_wait_sleep(sm.bars() + B_PLAN_FULL + slot, (k >> P_PLAN_LOG) & 1u);
                const PipePlan* pp = sm.plan(slot);
                if (pp->sid < 0) break;
                if (!pp->defer) {
                    const int ntiles = pp->ntiles, body = pp->body;
                    const double* vbody = pp->v + pp->head;
                    // The ring holds 3 tiles, which is less than HBM latency and warp skew need.  L2 is the rest of
                    // the ring: every tile is asked into L2 P_PF tiles before its copy, across the end of the
                    // series into the next one when that plan has arrived.
                    const PipePlan* np = nullptr;
                    bool looked = false;
                    for (int t = 0; t < ntiles; ++t) {
                        const uint32_t gg = g + (uint32_t)t;
                        const uint32_t s = gg % P_STAGES;
                        mbar_wait(sm.bars() + B_RING_EMPTY + s, ((gg / P_STAGES) & 1u) ^ 1u);
                        const uint32_t bytes = (uint32_t)min(P_TILE, body - t * P_TILE) * 8u;
                        mbar_expect_tx(sm.bars() + B_RING_FULL + s, bytes);
                        tma_bulk_g2s(sm.ring() + (size_t)s * P_TILE, vbody + (size_t)t * P_TILE, bytes,
                                     sm.bars() + B_RING_FULL + s);
                        const int tp = t + P_PF;
                        if (tp < ntiles) {
                            bulk_prefetch_l2(vbody + (size_t)tp * P_TILE, (uint32_t)min(P_TILE, body - tp * P_TILE) * 8u);
                        } else {
                            if (!looked) {
                                looked = true;
                                const uint32_t k1 = k + 1, slot1 = k1 & (P_PLANS - 1);
                                if (mbar_test(sm.bars() + B_PLAN_FULL + slot1, (k1 >> P_PLAN_LOG) & 1u)) {
                                    const PipePlan* q = sm.plan(slot1);
                                    if (q->sid >= 0 && !q->defer) np = q;
                                }
                            }
                            const int u = tp - ntiles;
                            if (np && u < np->ntiles)
                                bulk_prefetch_l2(np->v + np->head + (size_t)u * P_TILE,
                                                 (uint32_t)min(P_TILE, np->body - u * P_TILE) * 8u);
                        }
                    }
                    g += (uint32_t)ntiles;
                }
                mbar_arrive(sm.bars() + B_PLAN_FREE + slot);
            }
        }
    } else if (warp < P_PW + 1 + P_SW) {
        // ================================================================ STREAM
        const int stid = tid - (P_PW + 1) * 32;
        const int swarp = warp - (P_PW + 1);
        uint32_t g = 0, c = 0;
        for (uint32_t k = 0;; ++k) {
            const uint32_t slot = k & (P_PLANS - 1);
            { KO_PT0(); mbar_wait_sleep(sm.bars() + B_PLAN_FULL + slot, (k >> P_PLAN_LOG) & 1u); KO_PT1(11, stid == 0); }
            const PipePlan* pp = sm.plan(slot);
            if (pp->sid < 0) break;
            if (!pp->defer) {
                const uint32_t pool = c & 1u;
                mbar_wait_sleep(sm.bars() + B_POOL_FREE + pool, ((c >> 1) & 1u) ^ 1u);
                PipeResult* res = sm.res(pool);
                const int ni = pp->ni;
                PAcc acc;
                acc.sum = 0.0;
                acc.mn = pp->smin;
                acc.mx = pp->smax;
                acc.nvalid = 0;
                acc.cA[0] = acc.cA[1] = 0;
                uint32_t p0[P_NI];
                PStreamCtx cx;
                cx.pool32 = smem_u32(sm.pool(pool));
                cx.ovq32 = smem_u32(sm.ovq(pool));
                cx.ovcnt32 = smem_u32(&res->ovcnt[0]);
                cx.ovn32 = smem_u32(&res->ovn);
                cx.ovtag32 = smem_u32(&res->ovtag[0]);
                cx.trash32 = smem_u32(&res->trash);
#pragma unroll
                for (int i = 0; i < P_NI; ++i) {
                    const bool on = i < ni;
                    p0[i] = on ? cx.pool32 + (uint32_t)(((size_t)pp->slot0[i] * P_NS + stid) * 8) : 0u;
                    acc.p[i] = p0[i];
                    acc.pend[i] = on ? p0[i] + (uint32_t)pp->nslots[i] * P_NS * 8 : 0u;
                }
                const int nb = pp->nb;
                KO_PT0();
                if (pp->vals) {
                    pipe_stream_levels(pp, acc, cx.pool32, sm, g, stid, lane);
                } else if (pp->tail) {
                    if (nb == 0) pipe_stream<0, true>(pp, acc, cx, sm, g, stid, lane);
                    else if (nb == 1) pipe_stream<1, true>(pp, acc, cx, sm, g, stid, lane);
                    else pipe_stream<2, true>(pp, acc, cx, sm, g, stid, lane);
                } else {
                    if (nb == 0) pipe_stream<0, false>(pp, acc, cx, sm, g, stid, lane);
                    else if (nb == 1) pipe_stream<1, false>(pp, acc, cx, sm, g, stid, lane);
                    else pipe_stream<2, false>(pp, acc, cx, sm, g, stid, lane);
                }
                KO_PT1(9, stid == 0);
                // hand the partials to the finish warps
                const double ws = warp_sum(acc.sum), wmn = warp_min(acc.mn), wmx = warp_max(acc.mx);
                const unsigned wn = warp_sum_u(acc.nvalid);
                const unsigned ca0 = warp_sum_u(acc.cA[0]), ca1 = warp_sum_u(acc.cA[1]);
                unsigned pc[P_NI], wpc[P_NI];
#pragma unroll
                for (int i = 0; i < P_NI; ++i) {
                    pc[i] = (acc.p[i] - p0[i]) / (P_NS * 8);
                    wpc[i] = warp_sum_u(pc[i]);
                    res->pcount[i][stid] = (unsigned short)pc[i];
                }
                if (lane == 0) {
                    res->sum[swarp] = ws;
                    res->mn[swarp] = wmn;
                    res->mx[swarp] = wmx;
                    red_shared_add(&res->t_nvalid, wn);
                    red_shared_add(&res->t_cA[0], ca0);
                    red_shared_add(&res->t_cA[1], ca1);
#pragma unroll
                    for (int i = 0; i < P_NI; ++i) red_shared_add(&res->t_ncap[i], wpc[i]);
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(sm.bars() + B_HANDOFF + pool);
                ++c;
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(sm.bars() + B_PLAN_FREE + slot);
        }
    } else if (warp == P_PW + 1 + P_SW + P_FW) {
        // ================================================================ PROLOGUE
        // One lane between the stream warps and the finish warps: once every stream warp has handed a series off,
        // it works out what all finish threads need first (totals, nearest ranks, which interval answers which
        // quantile) and only then opens the pool to them.  The stream warps are the slower side and go straight
        // on to the next series; the finish warps work a series behind and have the time.
        if (lane == 0) {
            uint32_t c = 0;
            for (uint32_t k = 0;; ++k) {
                const uint32_t slot = k & (P_PLANS - 1);
                mbar_wait_sleep(sm.bars() + B_PLAN_FULL + slot, (k >> P_PLAN_LOG) & 1u);
                const PipePlan* pp = sm.plan(slot);
                if (pp->sid < 0) break;
                if (!pp->defer) {
                    const uint32_t pool = c & 1u;
                    mbar_wait_hint(sm.bars() + B_HANDOFF + pool, (c >> 1) & 1u);
                    if (!pp->vals) fin_prologue(pp, sm.res(pool), a.nq, a.q[0], a.q[1], a.q[2], a.q[3], a.wide);
                    mbar_arrive(sm.bars() + B_POOL_FULL + pool);
                    ++c;
                }
                mbar_arrive(sm.bars() + B_PLAN_FREE + slot);
            }
        }
    } else {
        // ================================================================ FINISH
        const int ftid = tid - (P_PW + 1) * 32 - P_NS, fwarp = warp - (P_PW + 1) - P_SW;
        uint32_t c = 0;
        for (uint32_t k = 0;; ++k) {
            const uint32_t slot = k & (P_PLANS - 1);
            mbar_wait_sleep(sm.bars() + B_PLAN_FULL + slot, (k >> P_PLAN_LOG) & 1u);
            const PipePlan* pp = sm.plan(slot);
            if (pp->sid < 0) break;
            int todo = 0;   // quantiles left for a later attempt
            if (pp->defer == 1) todo = pp->qsel < 0 ? (a.nq > 0 ? (1 << a.nq) - 1 : 1) : (1 << pp->qsel);
            if (!pp->defer) {
                const uint32_t pool = c & 1u;
                { KO_PT0(); mbar_wait_sleep(sm.bars() + B_POOL_FULL + pool, (c >> 1) & 1u); KO_PT1(12, ftid == 0); }
                PipeSel::Out* so = &sm.sel()->o[pool];   // scratch copy of this series (alternates like the pools)
                { KO_PT0(); todo = pp->vals ? pipe_finish_levels(a, pp, sm.res(pool), sm.sel(), so, sm.pool(pool), ftid, fwarp, lane)
                                          : pipe_finish(a, pp, sm.res(pool), sm.sel(), so, sm.pool(pool), sm.ovq(pool), ftid, fwarp, lane); KO_PT1(10, ftid == 0); }
                // Both finish routines end behind a barrier of the finish warps, after which only thread 0 still
                // reads this series' pool, hand-off and scratch copy: it resets them and releases the pool while
                // the other threads go on to the next series (other pool, other scratch copy) without waiting.
                if (ftid == 0) {
                    PipeResult* r0 = sm.res(pool);
                    r0->t_nvalid = 0;
                    r0->ovn = 0;
                    for (int i = 0; i < P_NI; ++i) r0->ovcnt[i] = r0->t_cA[i] = r0->t_cin[i] = r0->t_ncap[i] = 0;
                    so->fail = 0;
                    for (int j = 0; j < NI_MAX; ++j) {
                        so->list_cnt[j] = 0;
                        so->val[j] = __longlong_as_double(0x7FF8000000000000LL);
                    }
                    mbar_arrive(sm.bars() + B_POOL_FREE + pool);
                }
                ++c;
            }
            if (ftid == 0) {
                if (todo) {  // next attempt: wide brackets per quantile, then the exact slow path
                    bool first = true;
                    if (a.wide)   // several quantiles of one series may fail: list it once
                        first = (atomicOr(const_cast<int*>(&a.work_list[pp->item]), (int)0x80000000) >= 0);
                    if (first) {
                        const int pos = atomicAdd(&a.counters[a.out_cnt], 1);
                        a.out_list[pos] = (!a.wide && a.pack) ? (pp->sid | (todo << 27)) : pp->sid;
                    }
                }
            }
            if (ftid == 0) mbar_arrive(sm.bars() + B_PLAN_FREE + slot);
        }
    }
}


bool g_attr_done = false;

// Host-pointer calls run one launch sequence per staged chunk on the same scratch: the counters a caller wants to
// see (series that needed the second attempt / the slow path, device error flag) are added up here per chunk.
__global__ void series_tally(const int* counters, int* tally) {
    tally[0] += counters[4];
    tally[1] += counters[6];
    tally[2] |= counters[2];
}

constexpr int PLAN_CAP_CTX = 16384;      // plan window of the context's scratch (9 MB)
constexpr int PLAN_CAP_STATIC = 2048;    // ... of the static scratch of ko_series_rollup_static (1.1 MB)
// counters + the two series lists, rounded so that the plans behind them are 256-byte aligned
constexpr size_t series_lists_bytes(int64_t n_series) { return ((size_t)(8 + 2 * n_series) * sizeof(int) + 255) / 256 * 256; }

// The kernels of the long-series path, enqueued on `st`:
//   1. first attempt, in windows of plan_cap series: series_plan (brackets from a pre-sample) then series_pipe
//      (fast path: one pass per series);
//   2. second attempt over what step 1 could not finish, one work item per (series, unanswered quantile), with
//      brackets twice as wide -- a bracket that missed its rank or overflowed is settled here;
//   3. series_long, the exact multi-pass slow path, over what is still left (series with almost no valid samples,
//      brackets that fit neither the capture slots nor a levels plan).
// One window covers n_series <= plan_cap.  Beyond that the number of long series is read back from the device
// (one stream synchronisation) so that only the windows that hold work are launched.
static cudaError_t launch_long_path(SeriesArgs a, int num_sms, int64_t n_series, void* plans, int plan_cap,
                                    cudaStream_t st, int* launches, cudaEvent_t ev_p0 = nullptr,
                                    cudaEvent_t ev_p1 = nullptr) {
    int pgrid = num_sms;
    if ((int64_t)pgrid > n_series) pgrid = (int)n_series;
    int grid = 2 * num_sms;
    if ((int64_t)grid > n_series) grid = (int)n_series;
    a.pack = n_series < (1LL << 27) ? 1 : 0;
    a.plans = plans;
    a.plan_cap = plan_cap;
    int64_t n_long = n_series;
    if (n_series > (int64_t)plan_cap) {
        int h = 0;
        cudaError_t e = cudaMemcpyAsync(&h, a.counters, sizeof(int), cudaMemcpyDeviceToHost, st);
        if (e == cudaSuccess) e = cudaStreamSynchronize(st);
        if (e != cudaSuccess) return e;
        n_long = h;
    }
    a.work_list = a.long_list;
    a.work_cnt = 0;
    a.work_cur = 1;
    a.out_list = a.defer_list;
    a.out_cnt = 4;
    a.wide = 0;
    for (int64_t w0 = 0; w0 < n_long; w0 += plan_cap) {
        const int64_t cnt = n_long - w0 < plan_cap ? n_long - w0 : plan_cap;
        a.w_begin = w0;
        const int pg = (int)((cnt + PLAN_WARPS - 1) / PLAN_WARPS);
        series_plan<<<pg < KO_PLAN_GRID * num_sms ? pg : KO_PLAN_GRID * num_sms, PLAN_WARPS * 32, 0, st>>>(a);
        if (w0 == 0 && ev_p0) cudaEventRecord(ev_p0, st);
        series_pipe<<<(int64_t)pgrid < cnt ? pgrid : (int)cnt, P_THREADS, PIPE_SMEM_BYTES, st>>>(a);
        if (w0 == 0 && ev_p1) cudaEventRecord(ev_p1, st);
        *launches += 2;
    }
    if (n_long <= 0 && ev_p0) {   // nothing long: an empty interval
        cudaEventRecord(ev_p0, st);
        cudaEventRecord(ev_p1, st);
    }
    if (a.nq > 0) {
        a.work_list = a.defer_list;
        a.work_cnt = 4;
        a.out_list = a.long_list;   // the first list is no longer needed
        a.out_cnt = 6;
        a.wide = 1;
        a.work_cur = 5;
        a.w_begin = 0;
        series_plan<<<num_sms, PLAN_WARPS * 32, 0, st>>>(a);
        series_pipe<<<pgrid, P_THREADS, PIPE_SMEM_BYTES, st>>>(a);
        *launches += 2;
        a.work_list = a.long_list;
        a.work_cnt = 6;
    } else {
        a.work_list = a.defer_list;
        a.work_cnt = 4;
    }
    a.work_cur = 7;
    series_long<<<grid, NTHREADS, LONG_SMEM_BYTES, st>>>(a);
    *launches += 1;
    return cudaSuccess;
}

}  // namespace

// Enqueue the three kernels on `stream`.  scratch must hold 8 ints + n_series ints.
int ko_series_launch(ko_ctx* ctx, const double* d_values, const int64_t* d_offsets, int64_t n_series,
                     const double* q, int nq, double* d_sum, double* d_mean, double* d_min, double* d_max,
                     double* d_last, int64_t* d_count, double* d_quant, int* d_scratch, cudaStream_t stream) {
    KO_REQUIRE(nq >= 0 && nq <= KO_MAX_QUANTILES, "series rollup: nq must be 0..KO_MAX_QUANTILES");
    KO_REQUIRE(n_series < 2147483647LL, "series rollup: too many series in one call");
    SeriesArgs a;
    a.values = d_values;
    a.offsets = d_offsets;
    a.n_series = n_series;
    a.out_sum = d_sum;
    a.out_mean = d_mean;
    a.out_min = d_min;
    a.out_max = d_max;
    a.out_last = d_last;
    a.out_count = d_count;
    a.out_quant = d_quant;
    a.counters = d_scratch;
    a.long_list = d_scratch + 8;
    a.defer_list = d_scratch + 8 + n_series;
    a.nq = nq;
    // quantiles ascending (stable insertion sort), remembering the caller's column
    for (int j = 0; j < NI_MAX; ++j) {
        a.q[j] = 0.0;
        a.qcol[j] = 0;
    }
    for (int j = 0; j < nq; ++j) {
        KO_REQUIRE(q[j] >= 0.0 && q[j] <= 1.0, "series rollup: quantile outside [0,1]");
        int k = j;
        while (k > 0 && a.q[k - 1] > q[j]) {
            a.q[k] = a.q[k - 1];
            a.qcol[k] = a.qcol[k - 1];
            --k;
        }
        a.q[k] = q[j];
        a.qcol[k] = j;
    }
    if (!g_attr_done) {
        KO_CUDA(cudaFuncSetAttribute(series_long, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)LONG_SMEM_BYTES));
        KO_CUDA(cudaFuncSetAttribute(series_small, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     8 * SMALL_MAX * 8 + 8 * HBINS * 4));
        KO_CUDA(cudaFuncSetAttribute(series_pipe, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)PIPE_SMEM_BYTES));
        g_attr_done = true;
    }
    KO_CUDA(cudaMemsetAsync(d_scratch, 0, 8 * sizeof(int), stream));
    const int cb = 256;
    classify_series<<<(unsigned)((n_series + cb - 1) / cb), cb, 0, stream>>>(a);
    const size_t small_smem = nq > 0 ? (size_t)8 * SMALL_MAX * 8 + (size_t)8 * HBINS * 4 : 0;
    series_small<<<(unsigned)((n_series + 7) / 8), 256, small_smem, stream>>>(a);
    KO_CUDA(cudaEventRecord(ctx->ev_k0, stream));
    int launches = 2;
    void* plans = reinterpret_cast<unsigned char*>(d_scratch) + series_lists_bytes(n_series);
    KO_CUDA(launch_long_path(a, ctx->num_sms, n_series, plans, PLAN_CAP_CTX, stream, &launches, ctx->ev_p0,
                             ctx->ev_p1));
    KO_CUDA(cudaEventRecord(ctx->ev_k1, stream));
    ctx->k_timed = true;
    KO_CUDA(cudaGetLastError());
    ctx->counters.kernel_launches += launches;
    return KO_OK;
}

size_t ko_series_scratch_bytes(int64_t n_series) {
    return series_lists_bytes(n_series) + (size_t)PLAN_CAP_CTX * sizeof(PipePlan);
}

extern "C" int ko_series_rollup_dev(ko_ctx* ctx, const double* d_values, const int64_t* d_offsets,
                                    int64_t n_series, const double* q, int nq, double* d_out_sum,
                                    double* d_out_mean, double* d_out_min, double* d_out_max,
                                    double* d_out_last, int64_t* d_out_count, double* d_out_quantiles,
                                    void* stream) {
    KO_REQUIRE(ctx && ctx->alive, "ko_series_rollup_dev: ctx not initialised");
    KO_REQUIRE(n_series >= 0, "ko_series_rollup_dev: negative n_series");
    if (n_series == 0) return KO_OK;
    KO_REQUIRE(d_values && d_offsets, "ko_series_rollup_dev: NULL input");
    KO_REQUIRE(nq == 0 || q, "ko_series_rollup_dev: q is NULL");
    std::lock_guard<std::mutex> lk(ctx->mu);
    KO_CUDA(cudaSetDevice(ctx->device));
    int rc = ko_buf_reserve(&ctx->scratch, ko_series_scratch_bytes(n_series));
    if (rc) return rc;
    return ko_series_launch(ctx, d_values, d_offsets, n_series, q, nq, d_out_sum, d_out_mean, d_out_min,
                            d_out_max, d_out_last, d_out_count, d_out_quantiles, (int*)ctx->scratch.p,
                            (cudaStream_t)stream);
}

// Host-pointer entry point: chunked, double-buffered H2D of the sample column overlapped with the kernels.
extern "C" int ko_series_rollup(ko_ctx* ctx, const double* values, const int64_t* offsets, int64_t n_series,
                                const double* q, int nq, double* out_sum, double* out_mean, double* out_min,
                                double* out_max, double* out_last, int64_t* out_count, double* out_quantiles) {
    KO_REQUIRE(ctx && ctx->alive, "ko_series_rollup: ctx not initialised");
    KO_REQUIRE(n_series >= 0, "ko_series_rollup: negative n_series");
    if (n_series == 0) return KO_OK;
    KO_REQUIRE(offsets, "ko_series_rollup: offsets is NULL");
    KO_REQUIRE(nq >= 0 && nq <= KO_MAX_QUANTILES, "ko_series_rollup: nq must be 0..KO_MAX_QUANTILES");
    KO_REQUIRE(nq == 0 || q, "ko_series_rollup: q is NULL");
    for (int64_t i = 0; i < n_series; ++i)
        KO_REQUIRE(offsets[i + 1] >= offsets[i] && offsets[i + 1] - offsets[i] < 2147483647LL,
                   "ko_series_rollup: offsets must be non-decreasing with series shorter than 2^31");
    KO_REQUIRE(values || offsets[n_series] == offsets[0], "ko_series_rollup: values is NULL");
    std::lock_guard<std::mutex> lk(ctx->mu);
    KO_CUDA(cudaSetDevice(ctx->device));

    const size_t off_bytes = (size_t)(n_series + 1) * 8;
    const int ncols = 6 + nq;  // sum mean min max last count quant[nq]
    const size_t col = ko_align_up((size_t)n_series * 8, 256);
    int rc;
    if ((rc = ko_buf_reserve(&ctx->in_small, off_bytes))) return rc;
    if ((rc = ko_buf_reserve(&ctx->out_small, col * 6 + ko_align_up((size_t)n_series * 8 * (nq ? nq : 1), 256))))
        return rc;
    if ((rc = ko_buf_reserve(&ctx->scratch, ko_series_scratch_bytes(n_series) + 64))) return rc;   // + the tally
    (void)ncols;
    int64_t* d_off = (int64_t*)ctx->in_small.p;
    unsigned char* ob = (unsigned char*)ctx->out_small.p;
    double* d_sum = (double*)(ob + 0 * col);
    double* d_mean = (double*)(ob + 1 * col);
    double* d_min = (double*)(ob + 2 * col);
    double* d_max = (double*)(ob + 3 * col);
    double* d_last = (double*)(ob + 4 * col);
    int64_t* d_cnt = (int64_t*)(ob + 5 * col);
    double* d_q = (double*)(ob + 6 * col);

    // chunk plan: whole series per chunk, about 256 MB each
    const size_t target = (size_t)256 << 20;
    size_t max_chunk = 0;
    {
        int64_t s0 = 0;
        while (s0 < n_series) {
            int64_t s1 = s0;
            size_t bytes = 0;
            while (s1 < n_series && (bytes == 0 || bytes < target)) {
                bytes += (size_t)(offsets[s1 + 1] - offsets[s1]) * 8;
                ++s1;
            }
            if (bytes > max_chunk) max_chunk = bytes;
            s0 = s1;
        }
    }
    const bool multi = max_chunk < (size_t)(offsets[n_series] - offsets[0]) * 8;
    if ((rc = ko_buf_reserve(&ctx->stage[0], max_chunk + 256))) return rc;
    if (multi && (rc = ko_buf_reserve(&ctx->stage[1], max_chunk + 256))) return rc;

    KO_CUDA(cudaMemcpyAsync(d_off, offsets, off_bytes, cudaMemcpyHostToDevice, ctx->stream));
    ctx->counters.h2d_bytes += (int64_t)off_bytes;
    int* d_tally = reinterpret_cast<int*>(reinterpret_cast<unsigned char*>(ctx->scratch.p) +
                                          ko_series_scratch_bytes(n_series));   // 4 ints behind the scratch
    KO_CUDA(cudaMemsetAsync(d_tally, 0, 4 * sizeof(int), ctx->stream));
    int64_t s0 = 0;
    int c = 0;
    while (s0 < n_series) {
        int64_t s1 = s0;
        size_t bytes = 0;
        while (s1 < n_series && (bytes == 0 || bytes < target)) {
            bytes += (size_t)(offsets[s1 + 1] - offsets[s1]) * 8;
            ++s1;
        }
        const int b = c & 1;
        if (c >= 2) KO_CUDA(cudaStreamWaitEvent(ctx->copy_stream, ctx->ev_done[b], 0));
        if (bytes)
            KO_CUDA(cudaMemcpyAsync(ctx->stage[b].p, values + offsets[s0], bytes, cudaMemcpyHostToDevice,
                                    ctx->copy_stream));
        ctx->counters.h2d_bytes += (int64_t)bytes;
        KO_CUDA(cudaEventRecord(ctx->ev_copied[b], ctx->copy_stream));
        KO_CUDA(cudaStreamWaitEvent(ctx->stream, ctx->ev_copied[b], 0));
        // the kernels index values with absolute offsets: bias the base pointer by the chunk start
        const double* base = reinterpret_cast<const double*>(reinterpret_cast<uintptr_t>(ctx->stage[b].p) -
                                                              (uintptr_t)offsets[s0] * 8u);
        const int64_t ns = s1 - s0;
        rc = ko_series_launch(ctx, base, d_off + s0, ns, q, nq, out_sum ? d_sum + s0 : nullptr,
                              out_mean ? d_mean + s0 : nullptr, out_min ? d_min + s0 : nullptr,
                              out_max ? d_max + s0 : nullptr, out_last ? d_last + s0 : nullptr,
                              out_count ? d_cnt + s0 : nullptr,
                              (out_quantiles && nq) ? d_q + s0 * nq : nullptr, (int*)ctx->scratch.p, ctx->stream);
        if (rc) {
            cudaStreamSynchronize(ctx->stream);
            cudaStreamSynchronize(ctx->copy_stream);
            return rc;
        }
        series_tally<<<1, 1, 0, ctx->stream>>>((const int*)ctx->scratch.p, d_tally);
        ctx->counters.kernel_launches += 1;
        KO_CUDA(cudaEventRecord(ctx->ev_done[b], ctx->stream));
        // the scratch (work list / counters) is shared by consecutive chunks: they run in stream order
        s0 = s1;
        ++c;
    }
    const size_t nb = (size_t)n_series * 8;
    if (out_sum) KO_CUDA(cudaMemcpyAsync(out_sum, d_sum, nb, cudaMemcpyDeviceToHost, ctx->stream));
    if (out_mean) KO_CUDA(cudaMemcpyAsync(out_mean, d_mean, nb, cudaMemcpyDeviceToHost, ctx->stream));
    if (out_min) KO_CUDA(cudaMemcpyAsync(out_min, d_min, nb, cudaMemcpyDeviceToHost, ctx->stream));
    if (out_max) KO_CUDA(cudaMemcpyAsync(out_max, d_max, nb, cudaMemcpyDeviceToHost, ctx->stream));
    if (out_last) KO_CUDA(cudaMemcpyAsync(out_last, d_last, nb, cudaMemcpyDeviceToHost, ctx->stream));
    if (out_count) KO_CUDA(cudaMemcpyAsync(out_count, d_cnt, nb, cudaMemcpyDeviceToHost, ctx->stream));
    if (out_quantiles && nq)
        KO_CUDA(cudaMemcpyAsync(out_quantiles, d_q, nb * nq, cudaMemcpyDeviceToHost, ctx->stream));
    int flags[4] = {0, 0, 0, 0};   // over all chunks: second attempts, slow-path series, device error flag
    KO_CUDA(cudaMemcpyAsync(flags, d_tally, sizeof(flags), cudaMemcpyDeviceToHost, ctx->stream));
    KO_CUDA(cudaStreamSynchronize(ctx->stream));
    int nout = (out_sum != 0) + (out_mean != 0) + (out_min != 0) + (out_max != 0) + (out_last != 0) +
               (out_count != 0) + ((out_quantiles && nq) ? nq : 0);
    ctx->counters.d2h_bytes += (int64_t)nb * nout + (int64_t)sizeof(flags);
    ctx->counters.series_fallback = flags[0];  // series the first attempt of the fast path could not finish
    if (flags[2]) {
        ko_set_error("ko_series_rollup: device reported error flag %d", flags[2]);
        return KO_ECUDA;
    }
    return KO_OK;
}

// ---------------------------------------------------------------------------------------------------------
// ko_series_rollup_static — the same three kernels with their scratch (work cursor, flags, list of long
// series) in a fixed array that lives in the library image instead of in memory obtained at run time.
//
// Who needs this: a caller that may not allocate device memory at the moment it wants a rollup.  Examples are
// a process whose GPU memory has been handed over completely to another component's pool (a trainer or an
// inference server sharing the GPU with the monitoring worker), a request thread that must not run into the
// device-wide synchronisation that cudaMalloc / cudaFree imply, and code running between two points where the
// CUDA allocator is off limits.  ko_series_rollup_dev grows ctx->scratch with cudaMalloc when a bigger batch
// arrives; this entry point never does: no cudaMalloc, no cudaFree, no ctx (ko_init creates streams/events).
//
// Price: every GPU that loads libkorollup.so holds the array until the process exits
// (4 * (8 + 2 * KO_STATIC_MAX_SERIES) bytes = 512 KiB), batches above KO_STATIC_MAX_SERIES series are refused, and
// all calls on one device share the array: it keeps the previous call's contents (the counters are cleared on
// the caller's stream at the start of every call), so two calls whose kernels may overlap on the GPU — different
// streams, or threads launching concurrently — race on it.  Serialise them on one stream.
namespace {

static __device__ __align__(256) int g_static_scratch[(series_lists_bytes(KO_STATIC_MAX_SERIES) +
                                                        (size_t)PLAN_CAP_STATIC * sizeof(PipePlan)) / sizeof(int)];

constexpr int STATIC_MAX_DEVICES = 64;
struct StaticDevice {
    int* scratch = nullptr;  // DEVICE address of g_static_scratch on that GPU
    int num_sms = 0;
    bool ready = false;
};
StaticDevice g_static_dev[STATIC_MAX_DEVICES];
std::mutex g_static_mu;

// Once per device, with that device current: g_static_scratch as written in host code is only a host-side
// handle of the symbol, not an address a kernel can use, so the device address is looked up here.
int static_setup(int device, StaticDevice** out) {
    KO_REQUIRE(device >= 0 && device < STATIC_MAX_DEVICES, "ko_series_rollup_static: device ordinal out of range");
    std::lock_guard<std::mutex> lk(g_static_mu);
    KO_CUDA(cudaSetDevice(device));
    StaticDevice* sd = &g_static_dev[device];
    if (!sd->ready) {
        int* ptr = nullptr;
        KO_CUDA(cudaGetSymbolAddress((void**)&ptr, g_static_scratch));
        KO_REQUIRE(ptr != nullptr, "ko_series_rollup_static: cudaGetSymbolAddress returned NULL");
        int sms = 0;
        KO_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
        KO_CUDA(cudaFuncSetAttribute(series_long, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)LONG_SMEM_BYTES));
        KO_CUDA(cudaFuncSetAttribute(series_small, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     8 * SMALL_MAX * 8 + 8 * HBINS * 4));
        KO_CUDA(cudaFuncSetAttribute(series_pipe, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)PIPE_SMEM_BYTES));
        sd->scratch = ptr;
        sd->num_sms = sms;
        sd->ready = true;
    }
    *out = sd;
    return KO_OK;
}

}  // namespace

extern "C" int ko_series_rollup_static(int device, const double* d_values, const int64_t* d_offsets,
                                       int64_t n_series, const double* q, int nq, double* d_out_sum,
                                       double* d_out_mean, double* d_out_min, double* d_out_max, double* d_out_last,
                                       int64_t* d_out_count, double* d_out_quantiles, void* stream) {
    KO_REQUIRE(n_series >= 0, "ko_series_rollup_static: negative n_series");
    if (n_series > KO_STATIC_MAX_SERIES) {
        ko_set_error("ko_series_rollup_static: %lld series exceed the fixed scratch of KO_STATIC_MAX_SERIES=%d; "
                     "split the batch or use ko_series_rollup_dev",
                     (long long)n_series, KO_STATIC_MAX_SERIES);
        return KO_EINVAL;
    }
    KO_REQUIRE(nq >= 0 && nq <= KO_MAX_QUANTILES, "ko_series_rollup_static: nq must be 0..KO_MAX_QUANTILES");
    KO_REQUIRE(nq == 0 || q, "ko_series_rollup_static: q is NULL");
    StaticDevice* sd = nullptr;
    int rc = static_setup(device, &sd);  // first call on a device: symbol lookup + kernel attributes
    if (rc) return rc;
    if (n_series == 0) return KO_OK;
    KO_REQUIRE(d_values && d_offsets, "ko_series_rollup_static: NULL input");
    SeriesArgs a;
    a.values = d_values;
    a.offsets = d_offsets;
    a.n_series = n_series;
    a.out_sum = d_out_sum;
    a.out_mean = d_out_mean;
    a.out_min = d_out_min;
    a.out_max = d_out_max;
    a.out_last = d_out_last;
    a.out_count = d_out_count;
    a.out_quant = d_out_quantiles;
    a.counters = sd->scratch;
    a.long_list = sd->scratch + 8;
    a.defer_list = sd->scratch + 8 + n_series;
    a.nq = nq;
    for (int j = 0; j < NI_MAX; ++j) {
        a.q[j] = 0.0;
        a.qcol[j] = 0;
    }
    for (int j = 0; j < nq; ++j) {  // ascending quantiles, remembering the caller's column
        KO_REQUIRE(q[j] >= 0.0 && q[j] <= 1.0, "ko_series_rollup_static: quantile outside [0,1]");
        int k = j;
        while (k > 0 && a.q[k - 1] > q[j]) {
            a.q[k] = a.q[k - 1];
            a.qcol[k] = a.qcol[k - 1];
            --k;
        }
        a.q[k] = q[j];
        a.qcol[k] = j;
    }
    cudaStream_t st = (cudaStream_t)stream;
    // the array still holds the last call's cursor / flags / list: the eight counters must start at zero
    KO_CUDA(cudaMemsetAsync(sd->scratch, 0, 8 * sizeof(int), st));
    classify_series<<<(unsigned)((n_series + 255) / 256), 256, 0, st>>>(a);
    const size_t small_smem = nq > 0 ? (size_t)8 * SMALL_MAX * 8 + (size_t)8 * HBINS * 4 : 0;
    series_small<<<(unsigned)((n_series + 7) / 8), 256, small_smem, st>>>(a);
    int launches = 0;
    void* plans = reinterpret_cast<unsigned char*>(sd->scratch) + series_lists_bytes(KO_STATIC_MAX_SERIES);
    KO_CUDA(launch_long_path(a, sd->num_sms, n_series, plans, PLAN_CAP_STATIC, st, &launches));
    KO_CUDA(cudaGetLastError());
    return KO_OK;
}

#ifdef KO_PHASE_TIMING
extern "C" int ko_debug_phases(unsigned long long* out16, int reset) {
    KO_CUDA(cudaMemcpyFromSymbol(out16, g_phase, sizeof(unsigned long long) * 48));
    if (reset) {
        unsigned long long z[48] = {0};
        KO_CUDA(cudaMemcpyToSymbol(g_phase, z, sizeof(z)));
    }
    return KO_OK;
}
#endif
