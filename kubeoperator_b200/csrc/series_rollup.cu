// Window statistics over columnar f64 series: sum / count / mean / min / max / last + exact nearest-rank
// quantiles.  BUILD-DEFINED semantics (oracle/rollup_oracle.py series_stats); the reference only ships
// PromQL strings for this (core/apps/kubeops_api/prometheus_client.py:90-111), see include/korollup.h.
//
// HBM-bound by design: 8 algorithmic bytes per sample, read once.  Kernels of one call:
//   classify_series   - builds the work lists of long and of medium series (scratch), finds the longest, validates
//                       offsets
//   series_small      - series of <= 256 samples, 16 per warp: four windows of <= 32 samples side by side (quantiles
//                       by rank counting), 33..256 samples in registers with a one-pass histogram select
//   series_medium     - series of 257..1024 samples, from classify's list: one warp per series, keys in shared
//                       memory, MSB radix select
//   series_flow       - series_flow.cu: persistent, one CTA of 32 warps per SM, warp-specialised; plan (pre-sample,
//                       key interval per quantile), TMA ring, scan (statistics + "count below the interval, keep
//                       what is inside it") and exact selection among the kept candidates in ONE pass over HBM
//   series_flow again - second attempt with wide intervals over the few quantiles the first could not answer
//   series_long       - exact multi-pass slow path (histogram refinement in key space) over what is left
// Every interval is verified by exact counts, so the result is always the oracle's value.
#include <math.h>
#include <stdio.h>
#include <stdlib.h>

#include "series_common.cuh"

// series_flow.cu
cudaError_t ko_flow_set_attributes();
cudaError_t ko_flow_launch(const void* series_args, int grid, cudaStream_t st);

namespace {

constexpr int SMALL_MAX = 1024;      // series up to this length take the warp paths
constexpr int SMALL_SHORT = 256;     // ... up to this one series_small, above it series_medium (through a work list)
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
constexpr int HBINS = 256;
constexpr double CSIG = 3.0;         // bracket half width in binomial sigmas of the pre-sample

enum { M_COUNT = 0, M_CAPTURE = 1, M_HIST = 2 };
enum { A_NONE = 0, A_SELECT = 1 };


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

__device__ __forceinline__ void cons_sync() { asm volatile("bar.sync 1, %0;" ::"n"(NCONS) : "memory"); }


// ------------------------------------------------------------------------------------ classify kernel
// Appends `item` of every lane with `want` set to list[...] behind the counter: one atomic per warp.
__device__ __forceinline__ void warp_append(int* counter, int* list, bool want, int item) {
    const unsigned m = __ballot_sync(0xffffffffu, want);
    if (m == 0u) return;
    const int lane = threadIdx.x & 31;
    int base = 0;
    if (lane == __ffs(m) - 1) base = atomicAdd(counter, __popc(m));
    base = __shfl_sync(0xffffffffu, base, __ffs(m) - 1);
    if (want) list[base + __popc(m & ((1u << lane) - 1u))] = item;
}
__global__ void classify_series(SeriesArgs a) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;   // (blockDim.x is a multiple of 32: whole warps)
    int64_t L = 0;
    if (i < a.n_series) {
        L = a.offsets[i + 1] - a.offsets[i];
        if (L < 0 || L >= 2147483647LL) {
            atomicExch(&a.counters[2], 1);
            L = 0;
        }
    }
    const bool is_long = L > SMALL_MAX, is_medium = !is_long && L > SMALL_SHORT;
    warp_append(&a.counters[0], a.long_list, is_long, (int)i);
    // (the list of medium series lives where the first series_flow launch later writes its deferred items:
    // series_medium is done by then)
    warp_append(&a.counters[16], a.defer_list, is_medium, (int)i);
    int Lmax = is_long ? (int)L : 0;
    Lmax = __reduce_max_sync(0xffffffffu, Lmax);
    if ((threadIdx.x & 31) == 0 && Lmax > 0) atomicMax(&a.counters[8], Lmax);   // series_flow: may two series share the pool
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
__device__ __forceinline__ void small_sorted_quantiles(const SeriesArgs& a, unsigned long long (&r)[K], unsigned nv,
                                                       int64_t sid, int lane) {
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

// Windows of 33 .. 256 samples: one pass of a 256-bin histogram over the span [min, max] of the series' keys finds the
// bin of every wanted rank (bins are equal power-of-two ranges of key space, so the map is monotone); the few keys of
// that bin are compacted with ballots and the one with the wanted rank among them is picked by counting.  A bin with
// more than 32 keys (tie levels, clusters) sends the series to the register sort above.  scratch: 256 bins + 32 keys.
constexpr int SMALL_SCRATCH_BYTES = HBINS * 4 + 32 * 8;
template <int K>
__device__ __forceinline__ void small_hist_quantiles(const SeriesArgs& a, unsigned long long (&k)[K], unsigned nv,
                                                     double mn, double mx, int64_t sid, int lane,
                                                     unsigned char* scratch) {
    const double nan = __longlong_as_double(0x7FF8000000000000LL);
    if (nv == 0 || mn == mx) {   // no valid sample / one distinct value (-0 and +0 both give +0, as everywhere)
        if (lane < a.nq) a.out_quant[sid * a.nq + lane] = nv == 0 ? nan : mn + 0.0;
        return;
    }
    unsigned* hist = reinterpret_cast<unsigned*>(scratch);
    unsigned long long* cand = reinterpret_cast<unsigned long long*>(scratch + HBINS * 4);
    const unsigned long long kmin = ko_okey(mn + 0.0), diff = ko_okey(mx + 0.0) - kmin;
    const int sh = max(0, 56 - __clzll((long long)diff));   // (diff >> sh) <= 255
    reinterpret_cast<uint4*>(hist)[lane * 2] = make_uint4(0u, 0u, 0u, 0u);
    reinterpret_cast<uint4*>(hist)[lane * 2 + 1] = make_uint4(0u, 0u, 0u, 0u);
    __syncwarp();
    unsigned bin[K];
#pragma unroll
    for (int e = 0; e < K; ++e) {
        bin[e] = k[e] != KO_KEY_NAN ? (unsigned)((k[e] - kmin) >> sh) : 0xFFFFu;
        if (bin[e] < (unsigned)HBINS) red_shared_add(&hist[bin[e]], 1u);
    }
    __syncwarp();
    const uint4 h0 = reinterpret_cast<const uint4*>(hist)[lane * 2], h1 = reinterpret_cast<const uint4*>(hist)[lane * 2 + 1];
    const unsigned c[8] = {h0.x, h0.y, h0.z, h0.w, h1.x, h1.y, h1.z, h1.w};
    unsigned s = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += c[i];
    unsigned incl = s;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
#pragma unroll 1
    for (int j = 0; j < a.nq; ++j) {
        const unsigned p = nearest_rank_dev(nv, a.q[j]) - 1u;   // 0-based rank
        const unsigned bal = __ballot_sync(0xffffffffu, incl > p);
        const int src = __ffs(bal) - 1;                          // nv > p: some lane has it
        unsigned b = 0, before = 0, cnt = 0;
        if (lane == src) {
            unsigned run = incl - s;
            bool found = false;
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                if (!found && run + c[i] > p) {
                    b = (unsigned)(lane * 8 + i);
                    before = run;
                    cnt = c[i];
                    found = true;
                }
                run += c[i];
            }
        }
        b = __shfl_sync(0xffffffffu, b, src);
        before = __shfl_sync(0xffffffffu, before, src);
        cnt = __shfl_sync(0xffffffffu, cnt, src);
        if (cnt > 32u) {   // warp-uniform: sort the whole series instead (writes every quantile)
            small_sorted_quantiles<K>(a, k, nv, sid, lane);
            return;
        }
        unsigned base = 0;
#pragma unroll
        for (int e = 0; e < K; ++e) {
            const bool hit = bin[e] == b;
            const unsigned m = __ballot_sync(0xffffffffu, hit);
            if (hit) cand[base + __popc(m & ((1u << lane) - 1u))] = k[e];
            base += __popc(m);
        }
        __syncwarp();
        const unsigned rho = p - before;   // 0-based rank inside the bin
        const unsigned long long mine = lane < (int)cnt ? cand[lane] : KO_KEY_NAN;
        int less = 0;
#pragma unroll 1
        for (unsigned i = 0; i < cnt; ++i) less += cand[i] < mine ? 1 : 0;
        // the wanted key is the one with the most keys below it among those with at most rho below (ties share a count)
        const int best = (lane < (int)cnt && (unsigned)less <= rho) ? less : -1;
        const int top = __reduce_max_sync(0xffffffffu, best);
        const int owner = __ffs(__ballot_sync(0xffffffffu, best == top)) - 1;
        const unsigned long long key = __shfl_sync(0xffffffffu, mine, owner);
        if (lane == 0) a.out_quant[sid * a.nq + a.qcol[j]] = ko_key2d(key);
        __syncwarp();   // cand is reused by the next quantile
    }
}

// One series of at most 32 * K samples by one warp: the samples are loaded once, K per lane, and stay in registers
// for the statistics and for the quantiles.
template <int K>
__device__ __forceinline__ void small_one(const SeriesArgs& a, const double* v, int L, int64_t sid, int lane,
                                          unsigned char* scratch) {
    const double nan = __longlong_as_double(0x7FF8000000000000LL);
    double x[K];
#pragma unroll
    for (int e = 0; e < K; ++e) {
        const int i = lane + 32 * e;
        x[e] = i < L ? ko_ld_stream(v + i) : nan;
    }
    double sum = 0.0, mn = INFINITY, mx = -INFINITY, lastv = nan;
    unsigned nv = 0;
    int last_idx = -1;
#pragma unroll
    for (int e = 0; e < K; ++e) {
        if (x[e] == x[e]) {
            sum += x[e];
            nv++;
            last_idx = lane + 32 * e;
            lastv = x[e];
        }
        mn = fmin(mn, x[e]);
        mx = fmax(mx, x[e]);
    }
    sum = warp_sum(sum);
    mn = warp_min(mn);
    mx = warp_max(mx);
    nv = warp_sum_u(nv);
    const int last_all = __reduce_max_sync(0xffffffffu, last_idx);
    const int owner = __ffs(__ballot_sync(0xffffffffu, last_idx == last_all)) - 1;
    lastv = __shfl_sync(0xffffffffu, lastv, owner);
    if (lane == 0) {
        if (a.out_sum) a.out_sum[sid] = sum;
        if (a.out_mean) a.out_mean[sid] = nv > 0 ? sum / (double)nv : 0.0;
        if (a.out_min) a.out_min[sid] = nv > 0 ? mn + 0.0 : nan;
        if (a.out_max) a.out_max[sid] = nv > 0 ? mx + 0.0 : nan;
        if (a.out_last) a.out_last[sid] = last_all >= 0 ? lastv : nan;
        if (a.out_count) a.out_count[sid] = (int64_t)nv;
    }
    if (a.nq == 0 || !a.out_quant) return;
    unsigned long long k[K];
#pragma unroll
    for (int e = 0; e < K; ++e) k[e] = (x[e] == x[e]) ? ko_okey(x[e] + 0.0) : KO_KEY_NAN;
    if (K == 1) small_sorted_quantiles<K>(a, k, nv, sid, lane);
    else small_hist_quantiles<K>(a, k, nv, mn, mx, sid, lane, scratch);
}

// Four series of at most 32 samples side by side in one warp, eight lanes each (alert-rule windows: 20 samples =
// 5 min at 15 s).  Lane sl of a group holds samples sl, sl + 8, sl + 16, sl + 24 (loaded by the caller, NaN beyond the
// series).  Quantiles by counting: for every key the number of keys of its series below it; the key wanted for
// 0-based rank p is the one with the largest such count <= p.
__device__ __forceinline__ void small_quad(const SeriesArgs& a, int64_t sid, bool valid, const double (&x)[4], int Lmax,
                                           int lane) {
    const int sl = lane & 7;
    const double nan = __longlong_as_double(0x7FF8000000000000LL);
    double sum = 0.0, mn = INFINITY, mx = -INFINITY, lastv = nan;
    unsigned nv = 0;
    int last_idx = -1;
#pragma unroll
    for (int e = 0; e < 4; ++e) {
        if (x[e] == x[e]) {
            sum += x[e];
            nv++;
            last_idx = sl + 8 * e;
            lastv = x[e];
        }
        mn = fmin(mn, x[e]);
        mx = fmax(mx, x[e]);
    }
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) {   // stays inside the aligned group of eight lanes
        sum += __shfl_xor_sync(0xffffffffu, sum, o);
        mn = fmin(mn, __shfl_xor_sync(0xffffffffu, mn, o));
        mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        nv += __shfl_xor_sync(0xffffffffu, nv, o);
        const int oi = __shfl_xor_sync(0xffffffffu, last_idx, o);
        const double ov = __shfl_xor_sync(0xffffffffu, lastv, o);
        if (oi > last_idx) {
            last_idx = oi;
            lastv = ov;
        }
    }
    if (valid && sl == 0) {
        if (a.out_sum) a.out_sum[sid] = sum;
        if (a.out_mean) a.out_mean[sid] = nv > 0 ? sum / (double)nv : 0.0;
        if (a.out_min) a.out_min[sid] = nv > 0 ? mn + 0.0 : nan;
        if (a.out_max) a.out_max[sid] = nv > 0 ? mx + 0.0 : nan;
        if (a.out_last) a.out_last[sid] = last_idx >= 0 ? lastv : nan;
        if (a.out_count) a.out_count[sid] = (int64_t)nv;
    }
    if (a.nq == 0 || !a.out_quant) return;
    unsigned long long k[4];
    int less[4] = {0, 0, 0, 0};
#pragma unroll
    for (int e = 0; e < 4; ++e) k[e] = (x[e] == x[e]) ? ko_okey(x[e] + 0.0) : KO_KEY_NAN;
    const int ne = (Lmax + 7) >> 3;   // rows of eight that hold samples somewhere in the warp
#pragma unroll
    for (int ee = 0; ee < 4; ++ee) {
        if (ee < ne) {
#pragma unroll
            for (int src = 0; src < 8; ++src) {
                const unsigned long long kj = __shfl_sync(0xffffffffu, k[ee], src, 8);
#pragma unroll
                for (int e = 0; e < 4; ++e) less[e] += kj < k[e] ? 1 : 0;
            }
        }
    }
#pragma unroll 1
    for (int j = 0; j < a.nq; ++j) {
        const int p = nv > 0 ? (int)nearest_rank_dev(nv, a.q[j]) - 1 : -1;
        int best = -1;
        unsigned long long bk = KO_KEY_NAN;
#pragma unroll
        for (int e = 0; e < 4; ++e)   // (stale samples and empty slots have nv keys below them: never <= p)
            if (less[e] <= p && less[e] > best) {
                best = less[e];
                bk = k[e];
            }
#pragma unroll
        for (int o = 4; o > 0; o >>= 1) {
            const int ob = __shfl_xor_sync(0xffffffffu, best, o);
            const unsigned long long ok = __shfl_xor_sync(0xffffffffu, bk, o);
            if (ob > best) {
                best = ob;
                bk = ok;
            }
        }
        if (valid && sl == 0) a.out_quant[sid * a.nq + a.qcol[j]] = nv > 0 ? ko_key2d(bk) : nan;
    }
}

// Statistics of one series of at most SMALL_MAX samples by one warp, keys to shared memory (series_medium).
struct SmallStats {
    double sum, mn, mx;
    unsigned nv;
};
__device__ __forceinline__ SmallStats small_stats(const SeriesArgs& a, const double* v, int L, int64_t sid, int lane,
                                                  unsigned long long* keys) {
    double sum = 0.0, mn = INFINITY, mx = -INFINITY;
    unsigned nv = 0;
    int last_idx = -1;
    const double qnan = __longlong_as_double(0x7FF8000000000000LL);
    for (int i0 = lane; i0 < L; i0 += 128) {   // four loads of a lane in flight
        double x4[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) x4[u] = i0 + 32 * u < L ? ko_ld_stream(v + i0 + 32 * u) : qnan;
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int i = i0 + 32 * u;
            const double x = x4[u];
            const bool ok = (x == x);
            if (ok) {
                sum += x;
                nv++;
                last_idx = i;
            }
            mn = fmin(mn, x);
            mx = fmax(mx, x);
            if (keys && i < L) keys[i] = ok ? ko_okey(x + 0.0) : KO_KEY_NAN;
        }
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
    SmallStats st = {sum, mn, mx, nv};
    return st;
}

// Series of at most SMALL_SHORT samples: SMALL_PER_WARP consecutive series per warp.  When all of them have at most 32
// samples they are computed four at a time side by side (small_quad), the loads of all of them issued before the
// first is used; else one after the other by the whole warp (small_one).  Longer series are skipped here.
constexpr int SMALL_PER_WARP = 16;
__global__ void __launch_bounds__(256, 4) series_small(SeriesArgs a) {
    extern __shared__ __align__(16) unsigned char smem_small[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t sid0 = ((int64_t)blockIdx.x * 8 + warp) * SMALL_PER_WARP;
    if (sid0 >= a.n_series) return;
    // lane l < 17 holds offsets[sid0 + l] (clamped to the batch)
    int64_t off_l = 0;
    {
        const int64_t i = sid0 + lane;
        off_l = a.offsets[i <= a.n_series ? i : a.n_series];
    }
    const int64_t len_l = __shfl_down_sync(0xffffffffu, off_l, 1) - off_l;   // lanes 0..15: the series' lengths
    const bool mine_short = lane >= SMALL_PER_WARP || sid0 + lane >= a.n_series || (len_l >= 0 && len_l <= 32);
    if (__all_sync(0xffffffffu, mine_short)) {
        int Lmax = (lane < SMALL_PER_WARP && sid0 + lane < a.n_series) ? (int)len_l : 0;
        Lmax = __reduce_max_sync(0xffffffffu, Lmax);
        const int g = lane >> 3, sl = lane & 7;
        const double nan = __longlong_as_double(0x7FF8000000000000LL);
        double x[SMALL_PER_WARP / 4][4];
#pragma unroll
        for (int qd = 0; qd < SMALL_PER_WARP / 4; ++qd) {
            const int64_t sid = sid0 + qd * 4 + g;
            const int64_t off = __shfl_sync(0xffffffffu, off_l, qd * 4 + g);
            const int Lg = (int)__shfl_sync(0xffffffffu, len_l, qd * 4 + g);   // (every lane takes part)
            const int L = sid < a.n_series ? Lg : 0;
            const double* v = a.values + off;
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const int i = sl + 8 * e;
                x[qd][e] = i < L ? ko_ld_stream(v + i) : nan;
            }
        }
#pragma unroll
        for (int qd = 0; qd < SMALL_PER_WARP / 4; ++qd) {
            const int64_t sid = sid0 + qd * 4 + g;
            small_quad(a, sid, sid < a.n_series, x[qd], Lmax, lane);
        }
        return;
    }
    unsigned char* scratch = smem_small + (size_t)warp * SMALL_SCRATCH_BYTES;
#pragma unroll 1
    for (int gg = 0; gg < SMALL_PER_WARP; ++gg) {
        const int64_t s1 = sid0 + gg;
        const int64_t o1 = __shfl_sync(0xffffffffu, off_l, gg), L1 = __shfl_sync(0xffffffffu, len_l, gg);
        if (s1 >= a.n_series || L1 < 0 || L1 > SMALL_SHORT) continue;   // warp-uniform
        const int L = (int)L1;
        const double* v = a.values + o1;
        if (L <= 32) small_one<1>(a, v, L, s1, lane, scratch);
        else if (L <= 64) small_one<2>(a, v, L, s1, lane, scratch);
        else if (L <= 128) small_one<4>(a, v, L, s1, lane, scratch);
        else small_one<8>(a, v, L, s1, lane, scratch);
        __syncwarp();
    }
}

// Quantiles of a series whose keys lie in shared memory (257 .. 1024 samples): like small_hist_quantiles - one 256-bin
// histogram over [min, max] for all quantiles, then per quantile the keys of its bin are compacted (<= 32) and
// rank-counted.  Returns false (nothing written) when a bin is more crowded: the caller falls back to the radix select.
constexpr int MEDIUM_SMEM_BYTES = 8 * SMALL_MAX * 8 + 8 * HBINS * 4 + 8 * 32 * 8;
__device__ __forceinline__ bool medium_hist_quantiles(const SeriesArgs& a, const unsigned long long* keys, int L,
                                                      unsigned nv, double mn, double mx, int64_t sid, int lane,
                                                      unsigned* hist, unsigned long long* cand) {
    const unsigned long long kmin = ko_okey(mn + 0.0), diff = ko_okey(mx + 0.0) - kmin;
    const int sh = max(0, 56 - __clzll((long long)diff));   // (diff >> sh) <= 255
    reinterpret_cast<uint4*>(hist)[lane * 2] = make_uint4(0u, 0u, 0u, 0u);
    reinterpret_cast<uint4*>(hist)[lane * 2 + 1] = make_uint4(0u, 0u, 0u, 0u);
    __syncwarp();
    for (int i = lane; i < L; i += 32) {
        const unsigned long long k = keys[i];
        if (k != KO_KEY_NAN) red_shared_add(&hist[(unsigned)((k - kmin) >> sh)], 1u);
    }
    __syncwarp();
    const uint4 h0 = reinterpret_cast<const uint4*>(hist)[lane * 2], h1 = reinterpret_cast<const uint4*>(hist)[lane * 2 + 1];
    const unsigned c[8] = {h0.x, h0.y, h0.z, h0.w, h1.x, h1.y, h1.z, h1.w};
    unsigned s = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += c[i];
    unsigned incl = s;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    unsigned qb[NI_MAX], qbefore[NI_MAX], qcnt[NI_MAX];
    bool crowded = false;
#pragma unroll
    for (int j = 0; j < NI_MAX; ++j) {
        qb[j] = qbefore[j] = qcnt[j] = 0u;
        if (j < a.nq) {
            const unsigned p = nearest_rank_dev(nv, a.q[j]) - 1u;
            const int src = __ffs(__ballot_sync(0xffffffffu, incl > p)) - 1;
            unsigned b = 0, before = 0, cnt = 0;
            if (lane == src) {
                unsigned run = incl - s;
                bool found = false;
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    if (!found && run + c[i] > p) {
                        b = (unsigned)(lane * 8 + i);
                        before = run;
                        cnt = c[i];
                        found = true;
                    }
                    run += c[i];
                }
            }
            qb[j] = __shfl_sync(0xffffffffu, b, src);
            qbefore[j] = __shfl_sync(0xffffffffu, before, src);
            qcnt[j] = __shfl_sync(0xffffffffu, cnt, src);
            crowded = crowded || qcnt[j] > 32u;
        }
    }
    if (crowded) return false;
#pragma unroll
    for (int j = 0; j < NI_MAX; ++j) {
        if (j < a.nq) {
            unsigned base = 0;
            for (int i0 = 0; i0 < L; i0 += 32) {
                const int i = i0 + lane;
                const unsigned long long k = i < L ? keys[i] : KO_KEY_NAN;
                const bool hit = k != KO_KEY_NAN && (unsigned)((k - kmin) >> sh) == qb[j];
                const unsigned m = __ballot_sync(0xffffffffu, hit);
                if (hit) cand[base + __popc(m & ((1u << lane) - 1u))] = k;
                base += __popc(m);
            }
            __syncwarp();
            const unsigned cnt = qcnt[j], rho = nearest_rank_dev(nv, a.q[j]) - 1u - qbefore[j];
            const unsigned long long mine = lane < (int)cnt ? cand[lane] : KO_KEY_NAN;
            int less = 0;
#pragma unroll 1
            for (unsigned i = 0; i < cnt; ++i) less += cand[i] < mine ? 1 : 0;
            const int best = (lane < (int)cnt && (unsigned)less <= rho) ? less : -1;
            const int top = __reduce_max_sync(0xffffffffu, best);
            const int owner = __ffs(__ballot_sync(0xffffffffu, best == top)) - 1;
            const unsigned long long key = __shfl_sync(0xffffffffu, mine, owner);
            if (lane == 0) a.out_quant[sid * a.nq + a.qcol[j]] = ko_key2d(key);
            __syncwarp();   // cand is reused by the next quantile
        }
    }
    return true;
}

__device__ __forceinline__ void series_medium_one(const SeriesArgs& a, int64_t sid, int warp, int lane,
                                                  unsigned char* smem_small) {
    const int64_t off0 = a.offsets[sid];
    const int64_t Ll = a.offsets[sid + 1] - off0;
    if (Ll > SMALL_MAX || Ll <= SMALL_SHORT) return;   // (direct mode: not every series is this kernel's)
    const int L = (int)Ll;
    const double* v = a.values + off0;
    unsigned long long* keys = nullptr;
    unsigned* hist = nullptr;
    if (a.nq > 0) {
        keys = reinterpret_cast<unsigned long long*>(smem_small) + (size_t)warp * SMALL_MAX;
        hist = reinterpret_cast<unsigned*>(smem_small + (size_t)8 * SMALL_MAX * 8) + warp * HBINS;
    }
    const SmallStats st = small_stats(a, v, L, sid, lane, keys);
    if (a.nq > 0 && a.out_quant) {
        __syncwarp();
        if (st.nv > 0 && st.mn != st.mx) {
            unsigned long long* cand = reinterpret_cast<unsigned long long*>(smem_small + (size_t)8 * SMALL_MAX * 8 +
                                                                             (size_t)8 * HBINS * 4) + warp * 32;
            if (medium_hist_quantiles(a, keys, L, st.nv, st.mn, st.mx, sid, lane, hist, cand)) return;
            __syncwarp();
        }
        const unsigned long long kmin = ko_okey(st.mn + 0.0), kmax = ko_okey(st.mx + 0.0);
        for (int j = 0; j < a.nq; ++j) {
            double r = __longlong_as_double(0x7FF8000000000000LL);
            if (st.nv > 0) {
                unsigned rho = nearest_rank_dev(st.nv, a.q[j]);
                r = ko_key2d(warp_radix_select(keys, L, rho, kmin, kmax, hist, lane));
            }
            if (lane == 0) a.out_quant[sid * a.nq + a.qcol[j]] = r;
        }
    }
}

// Series of SMALL_SHORT + 1 .. SMALL_MAX samples: one warp per series of the list classify_series made (a fixed grid
// walks it: a batch without such series costs one idle launch, not a CTA per eight series), keys in shared memory,
// MSB-first radix select.
__global__ void __launch_bounds__(256) series_medium(SeriesArgs a, int direct) {
    extern __shared__ __align__(16) unsigned char smem_small[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (direct) {   // statistics only: no key buffers, CTAs are cheap, one warp per series of the batch
        const int64_t sid = (int64_t)blockIdx.x * 8 + warp;
        if (sid < a.n_series) series_medium_one(a, sid, warp, lane, smem_small);
        return;
    }
    const int n_medium = a.counters[16];
    for (int w = blockIdx.x * 8 + warp; w < n_medium; w += gridDim.x * 8) {
        series_medium_one(a, a.defer_list[w], warp, lane, smem_small);
        __syncwarp();   // the warp's key buffer is reused by its next series
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


// Host-pointer calls run one launch sequence per staged chunk on the same scratch: the counters a caller wants to
// see (series that needed the second attempt / the slow path, device error flag) are added up here per chunk.
__global__ void series_tally(const int* counters, int* tally) {
    tally[0] += counters[4];
    tally[1] += counters[6];
    tally[2] |= counters[2];
}

constexpr int N_COUNTERS = 24;   // [16]: series of SMALL_SHORT + 1 .. SMALL_MAX samples (series_medium's list)
constexpr int SPILL_MAX = 65536;         // candidate keys one series_flow CTA may spill to global scratch
// counters + the two work lists, rounded so that the spill space behind them is 256-byte aligned.  A list has room
// for KO_MAX_QUANTILES entries per series: the first attempt of series_flow defers single quantiles.
constexpr int LIST_PER_SERIES = KO_MAX_QUANTILES;
constexpr size_t series_lists_bytes(int64_t n_series) {
    return ((size_t)(N_COUNTERS + 2 * LIST_PER_SERIES * n_series) * sizeof(int) + 255) / 256 * 256;
}

// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) is per device: remember which devices have been set up.
constexpr int ATTR_MAX_DEVICES = 64;
bool g_attr_dev[ATTR_MAX_DEVICES];
std::mutex g_attr_mu;
int series_set_attributes(int device) {
    KO_REQUIRE(device >= 0 && device < ATTR_MAX_DEVICES, "series rollup: device ordinal out of range");
    std::lock_guard<std::mutex> lk(g_attr_mu);
    if (g_attr_dev[device]) return KO_OK;
    KO_CUDA(cudaFuncSetAttribute(series_long, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)LONG_SMEM_BYTES));
    KO_CUDA(cudaFuncSetAttribute(series_medium, cudaFuncAttributeMaxDynamicSharedMemorySize, MEDIUM_SMEM_BYTES));
    KO_CUDA(ko_flow_set_attributes());
    g_attr_dev[device] = true;
    return KO_OK;
}

// The kernels of the long-series path, enqueued on `st`:
//   1. series_flow over every long series: one pass per series, intervals 2.5 sigma wide (series_flow.cu);
//   2. series_flow again over what step 1 could not finish (a rank outside its interval, ~1 % of the series;
//      a candidate pool that overflowed), with intervals 5 sigma wide and the whole pool for one series;
//   3. series_long, the exact multi-pass slow path, over what is still left (series with almost no valid samples,
//      value spreads below the resolution of the high key, tie structures no interval fits).
// Each kernel reads the length of its list from the counters, so an empty list costs one idle launch.
static cudaError_t launch_long_path(SeriesArgs a, int num_sms, int64_t n_series, void* spill, size_t spill_bytes,
                                    cudaStream_t st, int* launches, cudaEvent_t ev_p0 = nullptr,
                                    cudaEvent_t ev_p1 = nullptr) {
    int fgrid = num_sms;
    if ((int64_t)fgrid > n_series) fgrid = (int)n_series;
    int grid = 2 * num_sms;
    if ((int64_t)grid > n_series) grid = (int)n_series;
    size_t cap = spill_bytes / 8 / (size_t)fgrid;
    if (cap > (size_t)SPILL_MAX) cap = SPILL_MAX;
    a.spill = reinterpret_cast<unsigned long long*>(spill);
    a.spill_cap = (int)(cap & ~(size_t)1);
    a.work_list = a.long_list;
    a.work_cnt = 0;
    a.work_cur = 1;
    a.out_list = a.defer_list;
    a.out_cnt = 4;
    a.wide = 0;
    // the first attempt defers single quantiles: up to KO_MAX_QUANTILES entries per series, and the list has room
    // for n_series entries, so only when that cannot overflow
    a.pack = (n_series < (1LL << 27)) ? 1 : 0;
    cudaError_t e;
    if (getenv("KO_DEBUG_FORCE_LONG")) {   // development: everything through the slow exact path
        a.work_cur = 7;
        series_long<<<grid, NTHREADS, LONG_SMEM_BYTES, st>>>(a);
        *launches += 1;
        return cudaGetLastError();
    }
    if (ev_p0) cudaEventRecord(ev_p0, st);
    if ((e = ko_flow_launch(&a, fgrid, st)) != cudaSuccess) return e;
    if (ev_p1) cudaEventRecord(ev_p1, st);
    *launches += 1;
    const char* dbg = getenv("KO_DEBUG_STAGES");   // development: stop after the first n kernels of the long path
    const int stages = dbg ? atoi(dbg) : 3;
    if (stages < 2) return cudaSuccess;
    if (a.nq > 0) {
        a.work_list = a.defer_list;
        a.work_cnt = 4;
        a.work_cur = 5;
        a.out_list = a.long_list;   // the first list is no longer needed
        a.out_cnt = 6;
        a.wide = 1;
        if ((e = ko_flow_launch(&a, fgrid, st)) != cudaSuccess) return e;
        *launches += 1;
        a.work_list = a.long_list;
        a.work_cnt = 6;
    } else {
        a.work_list = a.defer_list;
        a.work_cnt = 4;
    }
    a.work_cur = 7;
    if (stages < 3) return cudaSuccess;
    series_long<<<grid, NTHREADS, LONG_SMEM_BYTES, st>>>(a);
    *launches += 1;
    return cudaGetLastError();
}

// The two kernels of the series that are not long: each skips the series that are not its own.
static void launch_short_paths(const SeriesArgs& a, int64_t n_series, int nq, cudaStream_t st) {
    series_small<<<(unsigned)((n_series + 8 * SMALL_PER_WARP - 1) / (8 * SMALL_PER_WARP)), 256, 8 * SMALL_SCRATCH_BYTES, st>>>(a);
    const size_t medium_smem = nq > 0 ? (size_t)MEDIUM_SMEM_BYTES : 0;
    unsigned mgrid = (unsigned)((n_series + 7) / 8);
    // with key buffers (74 KB per CTA, three per SM) a fixed grid walks the list: a batch of a million alert windows
    // must not pay for 125 000 CTAs that find nothing to do
    // statistics only and a batch of moderate size: one warp per series of the batch, no list (CTAs without key
    // buffers that find nothing to do are cheap up to a few ten thousand of them)
    const int direct = (nq == 0 && n_series <= 262144) ? 1 : 0;
    const unsigned mcap = nq > 0 ? 12u * 148u : 64u * 148u;
    if (!direct && mgrid > mcap) mgrid = mcap;
    series_medium<<<mgrid, 256, medium_smem, st>>>(a, direct);
}

static int fill_args(SeriesArgs& a, const double* d_values, const int64_t* d_offsets, int64_t n_series,
                     const double* q, int nq, double* d_sum, double* d_mean, double* d_min, double* d_max,
                     double* d_last, int64_t* d_count, double* d_quant, int* d_scratch) {
    KO_REQUIRE(nq >= 0 && nq <= KO_MAX_QUANTILES, "series rollup: nq must be 0..KO_MAX_QUANTILES");
    KO_REQUIRE(n_series < 2147483647LL, "series rollup: too many series in one call");
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
    a.long_list = d_scratch + N_COUNTERS;
    a.defer_list = d_scratch + N_COUNTERS + LIST_PER_SERIES * n_series;
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
    return KO_OK;
}

}  // namespace

// Enqueue the kernels of one series rollup on `stream`.  d_scratch must hold ko_series_scratch_bytes().
int ko_series_launch(ko_ctx* ctx, const double* d_values, const int64_t* d_offsets, int64_t n_series,
                     const double* q, int nq, double* d_sum, double* d_mean, double* d_min, double* d_max,
                     double* d_last, int64_t* d_count, double* d_quant, int* d_scratch, cudaStream_t stream) {
    SeriesArgs a;
    int rc = fill_args(a, d_values, d_offsets, n_series, q, nq, d_sum, d_mean, d_min, d_max, d_last, d_count, d_quant,
                       d_scratch);
    if (rc) return rc;
    if ((rc = series_set_attributes(ctx->device))) return rc;
    KO_CUDA(cudaMemsetAsync(d_scratch, 0, N_COUNTERS * sizeof(int), stream));
    const int cb = 256;
    classify_series<<<(unsigned)((n_series + cb - 1) / cb), cb, 0, stream>>>(a);
    launch_short_paths(a, n_series, nq, stream);
    KO_CUDA(cudaEventRecord(ctx->ev_k0, stream));
    int launches = 3;
    void* spill = reinterpret_cast<unsigned char*>(d_scratch) + series_lists_bytes(n_series);
    KO_CUDA(launch_long_path(a, ctx->num_sms, n_series, spill, (size_t)ctx->num_sms * SPILL_MAX * 8, stream,
                             &launches, ctx->ev_p0, ctx->ev_p1));
    KO_CUDA(cudaEventRecord(ctx->ev_k1, stream));
    ctx->k_timed = true;
    ctx->last_series_stream = stream;
    ctx->last_series_scratch = d_scratch;
    KO_CUDA(cudaGetLastError());
    ctx->counters.kernel_launches += launches;
    return KO_OK;
}

extern "C" int ko_series_status(ko_ctx* ctx, int32_t* out4) {
    KO_REQUIRE(ctx && ctx->alive, "ko_series_status: ctx not initialised");
    KO_REQUIRE(out4, "ko_series_status: out4 is NULL");
    std::lock_guard<std::mutex> lk(ctx->mu);
    out4[0] = out4[1] = out4[2] = out4[3] = 0;
    if (!ctx->last_series_scratch) return KO_OK;
    KO_CUDA(cudaSetDevice(ctx->device));
    int h[N_COUNTERS];
    KO_CUDA(cudaMemcpyAsync(h, ctx->last_series_scratch, sizeof(h), cudaMemcpyDeviceToHost, ctx->last_series_stream));
    KO_CUDA(cudaStreamSynchronize(ctx->last_series_stream));
    out4[0] = h[0];
    out4[1] = h[4];
    out4[2] = h[6];
    out4[3] = h[2];
    if (getenv("KO_DEBUG_WHY"))
        fprintf(stderr, "series_flow deferrals: plan %d below %d above %d overflow %d const %d nointerval %d select %d\n",
                h[9], h[10], h[11], h[12], h[13], h[14], h[15]);
    return KO_OK;
}

size_t ko_series_scratch_bytes(int64_t n_series, int num_sms) {
    return series_lists_bytes(n_series) + (size_t)num_sms * SPILL_MAX * 8;
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
    int rc = ko_buf_reserve(&ctx->scratch, ko_series_scratch_bytes(n_series, ctx->num_sms));
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
    KO_REQUIRE(offsets[0] >= 0, "ko_series_rollup: offsets[0] is negative");
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
    if ((rc = ko_buf_reserve(&ctx->scratch, ko_series_scratch_bytes(n_series, ctx->num_sms) + 64))) return rc;   // + the tally
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
                                          ko_series_scratch_bytes(n_series, ctx->num_sms));   // 4 ints behind the scratch
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
// ko_series_rollup_static — the same kernels with their scratch (work cursors, flags, the lists of long and
// deferred series, spill space for quantile candidates) in a fixed array that lives in the library image instead of
// in memory obtained at run time.
//
// Who needs this: a caller that may not allocate device memory at the moment it wants a rollup.  Examples are
// a process whose GPU memory has been handed over completely to another component's pool (a trainer or an
// inference server sharing the GPU with the monitoring worker), a request thread that must not run into the
// device-wide synchronisation that cudaMalloc / cudaFree imply, and code running between two points where the
// CUDA allocator is off limits.  ko_series_rollup_dev grows ctx->scratch with cudaMalloc when a bigger batch
// arrives; this entry point never does: no cudaMalloc, no cudaFree, no ctx (ko_init creates streams/events).
//
// Price: every GPU that loads libkorollup.so holds the array until the process exits - 6.6 MiB: the counters and
// the two work lists for KO_STATIC_MAX_SERIES series (4 entries per series and list, 2 MiB) and 4.6 MiB of spill
// space for quantile candidates.  Batches above KO_STATIC_MAX_SERIES series are refused with KO_EINVAL and a message
// that names the limit.  All calls on one device share the array: it keeps the previous call's contents (the
// counters are cleared on the caller's stream at the start of every call), so two calls whose kernels may overlap on
// the GPU - different streams, or threads launching concurrently - race on it.  Serialise them on one stream.
// The spill space is small (4 096 candidate keys per CTA, 16 times less than ko_series_rollup_dev allocates): series
// whose candidates do not fit it - 30-day series - are computed by the exact multi-pass kernel instead of the one-pass kernel, with the same results.
namespace {

constexpr size_t STATIC_SPILL_BYTES = 148 * 4096 * 8;   // 4 096 candidate keys per CTA on a 148-SM GPU
static __device__ __align__(256) int g_static_scratch[(series_lists_bytes(KO_STATIC_MAX_SERIES) + STATIC_SPILL_BYTES) /
                                                       sizeof(int)];

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
        int rc = series_set_attributes(device);
        if (rc) return rc;
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
    rc = fill_args(a, d_values, d_offsets, n_series, q, nq, d_out_sum, d_out_mean, d_out_min, d_out_max, d_out_last,
                   d_out_count, d_out_quantiles, sd->scratch);
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    // the array still holds the last call's cursors / flags / lists: the counters must start at zero
    KO_CUDA(cudaMemsetAsync(sd->scratch, 0, N_COUNTERS * sizeof(int), st));
    classify_series<<<(unsigned)((n_series + 255) / 256), 256, 0, st>>>(a);
    launch_short_paths(a, n_series, nq, st);
    int launches = 0;
    // the lists are laid out for the call's n_series; the spill space starts behind the largest possible lists
    void* spill = reinterpret_cast<unsigned char*>(sd->scratch) + series_lists_bytes(KO_STATIC_MAX_SERIES);
    KO_CUDA(launch_long_path(a, sd->num_sms, n_series, spill, STATIC_SPILL_BYTES, st, &launches));
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
