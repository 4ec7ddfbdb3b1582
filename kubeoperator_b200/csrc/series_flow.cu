// series_flow — window statistics + exact nearest-rank quantiles of long series (> 1024 samples) in ONE pass over
// HBM.  BUILD-DEFINED semantics (oracle/rollup_oracle.py series_stats; include/korollup.h): the reference only
// ships PromQL strings for range windows (core/apps/kubeops_api/prometheus_client.py:90-111).
//
// One persistent CTA per SM, 32 warps, warp-specialised; three series are in flight per CTA:
//   PLANNER   (8 / 4 warps)  series k+1: strided pre-sample of up to 8192 points (aligned runs of four), two rounds of
//                        256-bin histograms over the 32-bit "high keys" (top half of the order-preserving image of
//                        an f64) give, per wanted quantile, a key interval [A, A+W] that holds its rank with
//                        2.5 sigma; heavy tie levels at the interval ends are listed so that they are counted instead
//                        of stored.
//   PRODUCER  (1 lane)   series k: cp.async.bulk (TMA, SASS UBLKCP) of 16 KB groups into a 7-group ring, completion
//                        on mbarrier transaction counts, groups asked into L2 ahead of the ring.
//   STREAM    (15 / 19 warps) series k: each warp owns whole 4 KB stages.  Per sample: sum, and - as double
//                        comparisons on the FP64 pipe with the results kept in predicates - the running-extreme /
//                        NaN filter, "count below the interval" and the in-interval test per interval.  The
//                        in-interval flags stay in a register (16 per lane and stage); after the stage's scan every
//                        lane reserves room for its flagged samples in the candidate pool (one shared-memory atomic;
//                        shared memory first, global spill behind it), copies them there as 64-bit keys, and the
//                        stage goes back to the producer.
//   FINISH    (8 warps)  series k-1: combines the warps' partial statistics, checks every interval by exact counts
//                        and selects the wanted order statistics among the candidates by 256-bin histogram refinement
//                        in 64-bit key space; writes the outputs.
// Whatever the plan cannot guarantee (a rank outside its interval, an overflowing pool, a pre-sample with almost no
// valid points) is detected from exact counts and the series goes to the next attempt's list: a second series_flow
// launch with 5-sigma intervals, then series_long.  The results are therefore always the oracle's.
#include <stdio.h>

#include "series_common.cuh"

namespace {

// 32 warps: 1 producer, F_NF finish, and stream + planner warps split by the launch's mode: series of more than
// F_DOUBLE_MAXL samples (one series per candidate pool) are bound by the scan and leave the planner idle most of the
// time, shorter ones keep eight planner warps busy.
constexpr int F_NS_LONG = 19, F_NP_LONG = 4;   // stream / planner warps when the batch holds a series > F_DOUBLE_MAXL
constexpr int F_NS_SHORT = 15, F_NP_SHORT = 8; // ... otherwise
constexpr int F_NS_MAX = 19;
constexpr int F_NF = 8;                        // finish warps
constexpr int F_WARPS = 32;
static_assert(1 + F_NP_LONG + F_NF + F_NS_LONG == F_WARPS && 1 + F_NP_SHORT + F_NF + F_NS_SHORT == F_WARPS, "role split");
constexpr int F_THREADS = F_WARPS * 32;        // 1024
// The scheduler prefers the warp with the highest id: the serial roles (whose latency everything else waits for)
// come last, the stream warps (throughput, always something to do) first.
// warps [0, ns) stream, [ns, ns + F_NF) finish, [ns + F_NF, 31) planner, 31 producer
constexpr int F_W_PROD = F_WARPS - 1;          // the producer warp
constexpr int F_STAGE = 512;                   // samples per ring stage (4 KB): 4 warp steps of 128 samples
constexpr int F_STAGE_BYTES = F_STAGE * 8;
constexpr int F_GS = 4;                        // stages per ring group: one TMA copy and one pair of barriers (16 KB)
constexpr int F_GROUP_BYTES = F_GS * F_STAGE_BYTES;
constexpr int F_NGROUPS = 7;                   // ring groups when two series share the candidate pool (112 KB)
// Long-series mode: 19 stream warps hold 19 of the ring's stages, which left 9 stages for copies in flight.  The
// candidate pool is of little use there (until the finish warps release it, most of a series' candidates go to the
// spill), so the ring takes 32 KB of it: 9 groups = 36 stages, and 3 072 pool entries remain.
constexpr int F_NGROUPS_LONG = 9;
constexpr int F_NGROUPS_MAX = 9;
constexpr int F_NSTAGES = F_NGROUPS * F_GS;
constexpr int F_PF = 8;                        // groups asked into L2 ahead of the ring
constexpr int F_ND = 4;                        // series slots (descriptor + plan + partials)
constexpr int F_SMAX = 8192;                   // largest pre-sample (30-day series)
constexpr int F_POOL = 7168;                   // candidate keys held in shared memory
constexpr int F_NH = 8;                        // heavy tie levels per plan
constexpr int F_LIST = 64;                     // candidates selected by direct rank counting
constexpr int F_NT = 2 * NI_MAX;               // interval ends refined in the planner's second round
constexpr int F_DOUBLE_MAXL = 49152;           // longest series for which two series share the candidate pool
constexpr int F_MIN_SAMPLE = 64;               // fewer valid pre-sample points: the series is deferred
constexpr uint32_t HK_PINF = 0xFFF00000u;      // high key of +inf; anything above is a (positive) NaN
constexpr uint32_t HK_NINF = 0x000FFFFFu;      // high key of -inf; anything below is a (negative) NaN
constexpr uint32_t HK_NONE = 0xFFFFFFFFu;      // a sample that takes no part (stale or outside the series)
constexpr uint32_t HK_UNUSED = 0xFFFFFFFEu;    // A of an unused interval (W = 0): no key equals it, HK_NONE is above

enum { FB_FULL = 0, FB_EMPTY = FB_FULL + F_NGROUPS_MAX, FB_DESC = FB_EMPTY + F_NGROUPS_MAX, FB_PLAN = FB_DESC + F_ND,
       FB_FREE = FB_PLAN + F_ND, FB_SDONE = FB_FREE + F_ND, FB_START = FB_SDONE + F_ND, FB_COUNT = FB_START + F_ND };

struct alignas(16) FlowSlot {
    // descriptor (planner thread 0)
    int sid;                 // < 0: no more work
    int L;
    int lead;                // elements in front of the series inside its first stage (0 or 1: 16-byte alignment)
    int nstages;
    uint32_t g0;             // ring stage counter of the series' first stage (a multiple of F_GS)
    uint32_t total_bytes;    // TMA bytes of the whole series (a multiple of 16)
    int tail_fix;            // the last element is stored by hand (it ends the whole array on an odd element)
    int depth1;              // long series: the planner stays one series ahead only (L2 footprint of its gather)
    int qsel;                // second attempt: the one quantile this work item is about, or -1 for all of them
    const double* src;       // 16-byte aligned address of stage 0
    long long start;         // element index of the series in values
    // plan (planner)
    int ni;                  // capture intervals, ascending and disjoint
    int defer;               // the plan gives up: the series goes to the next attempt
    int konst;               // the pre-sample holds one single value: no capture, the finish checks min == max
    int nheavy;
    uint32_t A[NI_MAX], W[NI_MAX];   // interval i holds the samples with A <= high key <= A + W
    double Ad[NI_MAX], Bd[NI_MAX];   // the same interval as doubles: Ad <= x <= Bd (the scan compares doubles)
    int qiv[NI_MAX];         // interval of quantile j
    double mn_seed, mx_seed; // valid samples of the series (or +inf / -inf): start of the running extremes
    unsigned long long heavy[F_NH];  // keys of heavy tie levels (zero canonical)
    // the same levels in a 16-slot hash table (the scan looks every flagged sample up): slot heavy_slot(key, hmul) holds
    // the key (or ~0) and htab_h its index in heavy[]; hmul == 0: no collision-free multiplier found, linear search
    unsigned long long htab[16];
    unsigned char htab_h[16];
    uint32_t hmul;
};

struct alignas(16) FlowPartial {
    double sum, mn, mx;
    unsigned nan;
    unsigned c[NI_MAX];
    unsigned pad[3];
};

struct PlanScratch {      // planner scratch
    unsigned long long kmin, kmax;
    unsigned nvalid;
    int work;
    int tbin[F_NT];          // round-1 bin of interval end t (A ends even, B ends odd), -1: open end
    unsigned trank[F_NT];    // sample rank looked for at that end (1-based, relative to its bin after round 1)
    unsigned tsub[F_NT];     // sub-bin found in round 2
    unsigned tcum[F_NT];     // sample keys below that sub-bin (whole sample)
    unsigned tcnt[F_NT];     // sample keys inside that sub-bin
    unsigned long long hmin[F_NT], hmax[F_NT];   // smallest / largest sample key of that sub-bin (tie level check)
    unsigned hklo[F_NT], hkhi[F_NT];             // smallest / largest HIGH key among the sample points of that sub-bin
    unsigned hifirst[F_NT], hilast[F_NT];        // first / last sample point (index) of that sub-bin
};

struct FinState {
    double sum, mn, mx, last;
    unsigned nvalid;
    unsigned cb[NI_MAX];
    int fail;
    int why;                 // first reason of a failure: 0 plan deferred, 1 rank below, 2 rank above, 3 pool overflow,
                             // 4 constant sample but not constant series, 5 no interval, 6 selection
    int pending;
    int st[NI_MAX];          // per quantile: 0 histogram round, 1 collect, 2 done, 3 failed
    int sh[NI_MAX];
    unsigned rho[NI_MAX];
    unsigned long long lo[NI_MAX];
    uint32_t hi32[NI_MAX];   // last high key of the quantile's interval: nothing above it takes part in its rounds
    unsigned long long ans[NI_MAX];
    unsigned list_n[NI_MAX];
};
enum { Q_HIST = 0, Q_COLLECT = 1, Q_DONE = 2, Q_FAIL = 3 };

// Development aid (-DKO_FLOW_PROGRESS): where every role of every CTA currently is, readable while the kernel hangs.
#ifdef KO_FLOW_PROGRESS
__device__ unsigned g_flow_prog[256 * 8];
__device__ unsigned g_flow_wide;
__device__ unsigned g_flow_trap[16];
__device__ __forceinline__ void flow_trap(unsigned who, unsigned k, unsigned grp, unsigned issued, unsigned extra) {
    if (atomicAdd(&g_flow_trap[0], 1u) == 0u) {
        g_flow_trap[1] = blockIdx.x; g_flow_trap[2] = who; g_flow_trap[3] = k; g_flow_trap[4] = grp;
        g_flow_trap[5] = issued; g_flow_trap[6] = extra;
    }
}
#define FP(role, k, phase) (reinterpret_cast<volatile unsigned*>(g_flow_prog)[blockIdx.x * 8 + (role)] = (g_flow_wide << 31) | ((unsigned)(k) << 12) | (unsigned)(phase))
#else
#define FP(role, k, phase) do { } while (0)
#endif

// Development aid (-DKO_FLOW_TIMING): cycles CTA 0 spends in each phase of each role, read by ko_debug_flow_phases.
#ifdef KO_FLOW_TIMING
__device__ unsigned long long g_flow_phase[64];
#define FT_BEGIN(cond)                     \
    unsigned long long _ft = clock64();    \
    const bool _ftc = (cond)
#define FT_MARK(slot)                                            \
    do {                                                         \
        if (_ftc) {                                              \
            const unsigned long long _t = clock64();             \
            atomicAdd(&g_flow_phase[slot], _t - _ft);            \
            _ft = _t;                                            \
        }                                                        \
    } while (0)
#else
#define FT_BEGIN(cond) do { } while (0)
#define FT_MARK(slot) do { } while (0)
#endif

constexpr size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }
constexpr size_t FO_RING = 0;
constexpr size_t FO_POOL = FO_RING + (size_t)F_NSTAGES * F_STAGE_BYTES;
constexpr size_t FO_HK = FO_POOL + (size_t)F_POOL * 8;
constexpr size_t FO_HIST1 = FO_HK + (size_t)F_SMAX * 4;
constexpr size_t FO_CUM = FO_HIST1 + 256 * 4;
constexpr size_t FO_HIST2 = FO_CUM + 272 * 4;
constexpr size_t FO_LUT = FO_HIST2 + (size_t)F_NT * 256 * 4;
constexpr size_t FO_FBINS = FO_LUT + 256;
constexpr size_t FO_FLISTK = FO_FBINS + (size_t)NI_MAX * 256 * 4;
constexpr size_t FO_FLISTM = FO_FLISTK + (size_t)NI_MAX * F_LIST * 8;
constexpr size_t FO_SLOTS = FO_FLISTM + (size_t)NI_MAX * F_LIST * 4;
constexpr size_t FO_PART = FO_SLOTS + align_up(F_ND * sizeof(FlowSlot), 16);
constexpr size_t FO_PS = FO_PART + (size_t)F_ND * F_NS_MAX * sizeof(FlowPartial);
constexpr size_t FO_FS = FO_PS + align_up(sizeof(PlanScratch), 16);
constexpr size_t FO_PSTATE = FO_FS + align_up(sizeof(FinState), 16);   // csm[F_ND], cgl[F_ND], ovf[F_ND], hcnt[F_ND][F_NH], rel[2], issued
constexpr int F_PSTATE_WORDS = F_ND * (3 + F_NH) + 2 + 2;
constexpr size_t FO_BARS = FO_PSTATE + align_up(F_PSTATE_WORDS * 4, 16);
constexpr size_t FLOW_SMEM_BYTES = FO_BARS + (size_t)FB_COUNT * 8;
static_assert(FLOW_SMEM_BYTES <= 232448, "series_flow must fit the 227 KB of one SM");
static_assert(sizeof(FlowPartial) == 64, "partials are 64 bytes");

struct FlowSmem {
    unsigned char* base;
    __device__ __forceinline__ unsigned char* ring() const { return base + FO_RING; }
    __device__ __forceinline__ unsigned long long* pool() const { return reinterpret_cast<unsigned long long*>(base + FO_POOL); }
    __device__ __forceinline__ uint32_t* hkeys() const { return reinterpret_cast<uint32_t*>(base + FO_HK); }
    __device__ __forceinline__ unsigned* hist1() const { return reinterpret_cast<unsigned*>(base + FO_HIST1); }
    __device__ __forceinline__ unsigned* cum() const { return reinterpret_cast<unsigned*>(base + FO_CUM); }
    __device__ __forceinline__ unsigned* hist2() const { return reinterpret_cast<unsigned*>(base + FO_HIST2); }
    __device__ __forceinline__ unsigned char* lut() const { return base + FO_LUT; }
    __device__ __forceinline__ unsigned* fbins() const { return reinterpret_cast<unsigned*>(base + FO_FBINS); }
    __device__ __forceinline__ unsigned long long* flistk() const { return reinterpret_cast<unsigned long long*>(base + FO_FLISTK); }
    __device__ __forceinline__ unsigned* flistm() const { return reinterpret_cast<unsigned*>(base + FO_FLISTM); }
    __device__ __forceinline__ FlowSlot* slot(int d) const { return reinterpret_cast<FlowSlot*>(base + FO_SLOTS) + d; }
    __device__ __forceinline__ FlowPartial* part(int d, int w) const {
        return reinterpret_cast<FlowPartial*>(base + FO_PART) + d * F_NS_MAX + w;
    }
    __device__ __forceinline__ PlanScratch* ps() const { return reinterpret_cast<PlanScratch*>(base + FO_PS); }
    __device__ __forceinline__ FinState* fs() const { return reinterpret_cast<FinState*>(base + FO_FS); }
    __device__ __forceinline__ unsigned* csm() const { return reinterpret_cast<unsigned*>(base + FO_PSTATE); }   // [F_ND] pool cursor
    __device__ __forceinline__ unsigned* cgl() const { return csm() + F_ND; }                                   // [F_ND] spill cursor
    __device__ __forceinline__ unsigned* ovf() const { return cgl() + F_ND; }                                   // [F_ND]
    __device__ __forceinline__ unsigned* hcnt() const { return ovf() + F_ND; }                                  // [F_ND][F_NH]
    __device__ __forceinline__ unsigned* rel() const { return hcnt() + F_ND * F_NH; }   // releases of pool window 0 / 1
    // Ring groups the producer has issued so far.  TMA copies land in any order and a parity wait only tells a phase
    // from its predecessor, so a consumer first makes sure that "its" use of a ring slot has been issued (which
    // implies that the slot's previous use is complete) and only then waits for the barrier's parity.
    __device__ __forceinline__ volatile unsigned* issued() const { return rel() + 2; }
    __device__ __forceinline__ uint64_t* bar(int i) const { return reinterpret_cast<uint64_t*>(base + FO_BARS) + i; }
};

// Where one series' candidates live: a window of the shared-memory pool and a window of the CTA's global spill,
// both filled through cursors shared by the stream warps (one atomic per lane and stage).  Cursors, tie-level counters and
// the spill window belong to the series' slot (k % F_ND: a slot is reused only after the finish warps are done
// with it).  The shared-memory pool is halved when series are short enough (two series in flight), otherwise it is
// one window; either way the next series may only enter a window once the finish warps have released it (a release
// counter in shared memory) -- until then its candidates go to the spill.
struct PoolWin {
    uint32_t s_base, s_cap;     // entries pool[s_base ... s_base + s_cap)
    uint32_t g_base, g_cap;     //         spill[g_base ... g_base + g_cap)
    int d;                      // slot: which cursors / spill window
    int win;                    // which shared-memory window (0 / 1)
    uint32_t use;               // releases of that window that must have happened before this series may enter it
};
__device__ __forceinline__ PoolWin pool_window(bool single, uint32_t k, int spill_cap) {
    PoolWin w;
    w.d = (int)(k % F_ND);
    w.g_cap = (uint32_t)spill_cap / F_ND;
    w.g_base = (uint32_t)w.d * w.g_cap;
    if (single) {
        w.win = 0;
        w.use = k;
        w.s_base = (uint32_t)(F_NGROUPS_LONG - F_NGROUPS) * F_GROUP_BYTES / 8u;   // behind the ring's extra groups
        w.s_cap = F_POOL - w.s_base;
    } else {
        w.win = (int)(k & 1u);
        w.use = k >> 1;
        w.s_cap = F_POOL / 2;
        w.s_base = (uint32_t)w.win * (F_POOL / 2);
    }
    return w;
}

// Ring slot and use count of ring group grp: grp / ng by a multiplication (exact for grp < 2^32 / 9).
__device__ __forceinline__ void ring_slot(uint32_t grp, bool single, uint32_t& s, uint32_t& u) {
    const uint32_t ng = single ? (uint32_t)F_NGROUPS_LONG : (uint32_t)F_NGROUPS;
    const uint32_t magic = single ? 0xFFFFFFFFu / F_NGROUPS_LONG + 1u : 0xFFFFFFFFu / F_NGROUPS + 1u;
    u = __umulhi(grp, magic);
    s = grp - u * ng;
}
__device__ __forceinline__ void wait_group_full(const FlowSmem& sm, uint32_t grp, bool single) {
    volatile unsigned* iss = sm.issued();
    while (*iss <= grp) __nanosleep(20);
    uint32_t s, u;
    ring_slot(grp, single, s, u);
    mbar_wait(sm.bar(FB_FULL + (int)s), u & 1u);
}
__device__ __forceinline__ void plan_sync(int nthreads) { asm volatile("bar.sync 3, %0;" ::"r"(nthreads) : "memory"); }
__device__ __forceinline__ void fin_sync() { asm volatile("bar.sync 2, %0;" ::"n"(F_NF * 32) : "memory"); }

__device__ __forceinline__ uint64_t createpolicy_evict_first() {
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
    return p;
}
// TMA bulk copy with an L2 eviction hint: the streamed samples are read once and should not push out the lines
// the planner gathered for the next series.
__device__ __forceinline__ void tma_bulk_g2s_hint(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar,
                                                  uint64_t policy) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::
            "r"(dst), "l"(src), "r"(bytes), "r"(bar), "l"(policy)
        : "memory");
}
__device__ __forceinline__ unsigned long long ld_cg_u64(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.global.cg.u64 %0, [%1];" : "=l"(v) : "l"(p));
    return v;
}
__device__ __forceinline__ void sts_f64(uint32_t addr, double x) {
    asm volatile("st.shared.f64 [%0], %1;" ::"r"(addr), "d"(x) : "memory");
}
__device__ __forceinline__ void red_shared_max_u64(const void* p, unsigned long long v) {
    asm volatile("red.shared.max.u64 [%0], %1;" ::"r"(smem_u32(p)), "l"(v) : "memory");
}
__device__ __forceinline__ void red_shared_min_u64(const void* p, unsigned long long v) {
    asm volatile("red.shared.min.u64 [%0], %1;" ::"r"(smem_u32(p)), "l"(v) : "memory");
}

__device__ __forceinline__ void ld_v4f64(const double* p, double& x, double& y, double& z, double& w) {   // 32 bytes
    asm volatile("ld.global.v4.f64 {%0, %1, %2, %3}, [%4];" : "=d"(x), "=d"(y), "=d"(z), "=d"(w) : "l"(p));
}
// Locate the bin whose cumulative count first reaches rho (1-based) in a 256-bin histogram; one full warp.
// Returns false when the histogram holds fewer than rho entries.  total = all entries.
__device__ __forceinline__ bool warp_find_bin(const unsigned* hist, unsigned rho, int lane, int& bin, unsigned& before,
                                              unsigned& cnt, unsigned& total) {
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
        const unsigned t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    total = __shfl_sync(0xffffffffu, incl, 31);
    const unsigned ballot = __ballot_sync(0xffffffffu, incl >= rho);
    if (ballot == 0u || rho == 0u) return false;
    const int src = __ffs(ballot) - 1;
    int d = 0;
    unsigned b = 0, n = 0;
    if (lane == src) {
        unsigned run = incl - s;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (n == 0u && run + c[i] >= rho) {
                d = i;
                b = run;
                n = c[i];
            }
            if (n == 0u) run += c[i];
        }
        d += lane * 8;
    }
    bin = __shfl_sync(0xffffffffu, d, src);
    before = __shfl_sync(0xffffffffu, b, src);
    cnt = __shfl_sync(0xffffffffu, n, src);
    return true;
}

// ------------------------------------------------------------------------------------------------ planner
__device__ __forceinline__ uint32_t heavy_slot(unsigned long long key, uint32_t mul) {
    const uint32_t t = (uint32_t)(key >> 32) ^ ((uint32_t)key * 0x85EBCA6Bu);
    return (t * mul) >> 28;
}
__device__ __forceinline__ int sample_size(int L) {
    int s = 256;
    while (s < F_SMAX && s * 16 <= L) s <<= 1;   // about L/16 .. L/32 points, 256 .. 8192
    return s;
}
// The pre-sample is taken as S/4 aligned runs of four neighbouring samples: one 32-byte request fetches four points,
// and what bounds the gather is the number of memory requests an SM gets through (about one in five cycles beside
// the bulk stream), not their size.  For samples that are independent from one 15 s step to the next this is as good
// as S separate points; for a smooth series the four are correlated and the intervals come out narrower than the
// sigma they claim -- that costs such a series a second attempt now and then (the exact counts notice), never a
// wrong result.
__device__ __forceinline__ int quad_pos(int j, int L, int nquad, int lead4) {
    int p = (int)(((2LL * j + 1) * (long long)L) / (2LL * nquad));
    p -= (lead4 + p) & 3;
    if (p < 0) p += 4;
    if (p + 3 >= L) p -= 4;
    return p;
}

// All planner threads (ptid in [0, NT)).  Fills the plan fields of the slot.
__device__ __forceinline__ void planner_plan(const SeriesArgs& a, const FlowSmem& sm, FlowSlot* sl, int ptid, bool single,
                                             uint32_t kser, int NT) {
    PlanScratch* ps = sm.ps();
    const int lane = ptid & 31;
    const int L = sl->L;
    const double* v = a.values + sl->start;
    const int nq = a.nq;
    const int qsel = sl->qsel;
    FT_BEGIN(blockIdx.x == 0 && ptid == 0);
    const double pinf = __longlong_as_double(0x7FF0000000000000LL);
    if (ptid == 0) {
        sl->ni = 0;
        sl->defer = 0;
        sl->konst = 0;
        sl->nheavy = 0;
        sl->mn_seed = pinf;
        sl->mx_seed = -pinf;
        ps->kmin = ~0ull;
        ps->kmax = 0ull;
        ps->nvalid = 0;
    }
    if (nq == 0) {
        plan_sync(NT);
        return;
    }
    const int S = sample_size(L);
    const int nquad = S / 4;
    const int qper = (nquad + NT - 1) / NT;   // runs of four per thread: 1 (a 256-point sample uses half the threads) .. 8
    const int lead4 = (int)((reinterpret_cast<uintptr_t>(v) >> 3) & 3u);
    uint32_t* sk = sm.hkeys();
    plan_sync(NT);
    // ---- gather: four 32-byte loads of a thread in flight together; the high keys go to shared memory (that is all
    // the histogram rounds need), the exact extremes are kept as 64-bit keys on the way
    {
        unsigned long long kmn = ~0ull, kmx = 0ull;
        unsigned nv = 0;
#pragma unroll 1
        for (int b = 0; b < qper; b += 4) {
            double x[16];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int j = ptid + (b + u) * NT;
                x[4 * u] = x[4 * u + 1] = x[4 * u + 2] = x[4 * u + 3] = 0.0;
                if (b + u < qper && j < nquad)
                    ld_v4f64(v + quad_pos(j, L, nquad, lead4), x[4 * u], x[4 * u + 1], x[4 * u + 2], x[4 * u + 3]);
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int j = ptid + (b + u) * NT;
                if (b + u < qper && j < nquad) {
                    uint32_t hk4[4];
#pragma unroll
                    for (int h = 0; h < 4; ++h) {
                        const double xx = x[4 * u + h];
                        const bool ok = (xx == xx);
                        const unsigned long long k = ko_okey(xx + 0.0);
                        hk4[h] = ok ? (uint32_t)(k >> 32) : HK_NONE;
                        if (ok) {
                            kmn = k < kmn ? k : kmn;
                            kmx = k > kmx ? k : kmx;
                            nv++;
                        }
                    }
                    *reinterpret_cast<uint4*>(sk + 4 * j) = make_uint4(hk4[0], hk4[1], hk4[2], hk4[3]);
                }
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const unsigned long long a0 = __shfl_xor_sync(0xffffffffu, kmn, o);
            const unsigned long long a1 = __shfl_xor_sync(0xffffffffu, kmx, o);
            kmn = a0 < kmn ? a0 : kmn;
            kmx = a1 > kmx ? a1 : kmx;
        }
        nv = __reduce_add_sync(0xffffffffu, nv);
        if (lane == 0 && nv) {
            red_shared_min_u64(&ps->kmin, kmn);
            red_shared_max_u64(&ps->kmax, kmx);
            red_shared_add(&ps->nvalid, nv);
        }
        for (int i = ptid; i < 256; i += NT) sm.hist1()[i] = 0;
        for (int i = ptid; i < F_NT * 256; i += NT) sm.hist2()[i] = 0;
        for (int i = ptid; i < 256 / 4; i += NT) reinterpret_cast<uint32_t*>(sm.lut())[i] = 0;
    }
    plan_sync(NT);
    FT_MARK(10);   // gather
    const unsigned ns = ps->nvalid;
    const unsigned long long kmin64 = ps->kmin, kmax64 = ps->kmax;
    const uint32_t hkmin = (uint32_t)(kmin64 >> 32), hkmax = (uint32_t)(kmax64 >> 32);
    if (ns < (unsigned)F_MIN_SAMPLE || kmin64 == kmax64 || hkmin == hkmax) {
        // almost no valid points or a spread below the resolution of the high key (not for this kernel: deferred),
        // or one single value (no capture; the finish warps check min == max over the whole series)
        if (ptid == 0) {
            if (ns >= (unsigned)F_MIN_SAMPLE && kmin64 == kmax64) sl->konst = 1;
            else sl->defer = 1;
#ifdef KO_FLOW_DEBUG
            if (sl->defer) printf("plan deferred: sid %d L %d ns %u kmin %llx kmax %llx\n", sl->sid, L, ns, kmin64, kmax64);
#endif
            if (ns) {
                sl->mn_seed = ko_key2d(kmin64);
                sl->mx_seed = ko_key2d(kmax64);
            }
        }
        plan_sync(NT);
        return;
    }
    // ---- round 1: 256 bins over [hkmin, hkmax]
    const uint32_t R = hkmax - hkmin;
    const int sh1 = max(0, 32 - __clz((int)R) - 8);
    const int sh2 = sh1 >= 8 ? sh1 - 8 : 0;
    const uint32_t submask = (1u << (sh1 - sh2)) - 1u;
    // (four keys per thread and round through one 16-byte load: four independent chains instead of one)
#pragma unroll 1
    for (int q = ptid; q < nquad; q += NT) {
        const uint4 k4 = *reinterpret_cast<const uint4*>(sk + 4 * q);
        if (k4.x != HK_NONE) red_shared_add(&sm.hist1()[(k4.x - hkmin) >> sh1], 1u);
        if (k4.y != HK_NONE) red_shared_add(&sm.hist1()[(k4.y - hkmin) >> sh1], 1u);
        if (k4.z != HK_NONE) red_shared_add(&sm.hist1()[(k4.z - hkmin) >> sh1], 1u);
        if (k4.w != HK_NONE) red_shared_add(&sm.hist1()[(k4.w - hkmin) >> sh1], 1u);
    }
    plan_sync(NT);
    if (ptid < 32) {   // exclusive prefix of the 256 bins -> cum[0..256]
        unsigned c[8];
        unsigned sum8 = 0;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            c[i] = sm.hist1()[lane * 8 + i];
            sum8 += c[i];
        }
        unsigned incl = sum8;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        unsigned run = incl - sum8;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            sm.cum()[lane * 8 + i] = run;
            run += c[i];
        }
        if (lane == 31) sm.cum()[256] = run;
        __syncwarp();
        // interval ends in sample ranks; lane t handles end t (A ends even, B ends odd)
        int tb = -1;
        if (lane < F_NT) {
            const int j = lane >> 1;
            unsigned tr = 0;
            if (j < nq && (qsel < 0 || j == qsel)) {
                const double z = a.wide ? 5.0 : 2.5;
                const double p = a.q[j];
                const double n = (double)ns;
                double fpc = 1.0 - (double)S / (double)L;
                fpc = fpc < 0.05 ? 0.05 : fpc;
                const double t = p * n;
                const double dl = z * sqrt(p * (1.0 - p) * n * fpc) + 2.0;
                const long long r = (lane & 1) ? (long long)ceil(t + dl) + 1 : (long long)floor(t - dl);
                if (r >= 1 && r <= (long long)ns) {
                    // the bin holding sample rank r: cum[b] < r <= cum[b + 1]
                    int lo = 0, hi = 255;
                    while (lo < hi) {
                        const int mid = (lo + hi + 1) >> 1;
                        if (sm.cum()[mid] < (unsigned)r) lo = mid;
                        else hi = mid - 1;
                    }
                    tb = lo;
                    tr = (unsigned)r - sm.cum()[lo];
                }
            }
            ps->tbin[lane] = tb;
            ps->trank[lane] = tr;
            ps->tsub[lane] = 0;
            ps->tcum[lane] = 0;
            ps->tcnt[lane] = 0;
            ps->hmin[lane] = ~0ull;
            ps->hmax[lane] = 0ull;
            ps->hklo[lane] = ~0u;
            ps->hkhi[lane] = 0u;
            ps->hifirst[lane] = ~0u;
            ps->hilast[lane] = 0u;
        }
        // the look-up table "which ends live in this bin": byte-wide OR through 32-bit atomics
        if (lane < F_NT && tb >= 0)
            atomicOr(reinterpret_cast<unsigned*>(sm.lut()) + (tb >> 2), (1u << lane) << ((tb & 3) * 8));
    }
    plan_sync(NT);
    FT_MARK(11);   // round 1 + targets
    // ---- round 2: 256 sub-bins inside the bins that hold an interval end
#pragma unroll 1
    for (int q = ptid; q < nquad; q += NT) {
        const uint4 k4 = *reinterpret_cast<const uint4*>(sk + 4 * q);
        const uint32_t kk[4] = {k4.x, k4.y, k4.z, k4.w};
        unsigned mm[4];
#pragma unroll
        for (int h = 0; h < 4; ++h) mm[h] = kk[h] != HK_NONE ? sm.lut()[(kk[h] - hkmin) >> sh1] : 0u;   // four look-ups in flight
#pragma unroll
        for (int h = 0; h < 4; ++h) {
            unsigned m = mm[h];
            const uint32_t sub = ((kk[h] - hkmin) >> sh2) & submask;
            while (m) {   // rare: the key lies in a bin that holds an interval end
                const int t = __ffs(m) - 1;
                m &= m - 1;
                red_shared_add(&sm.hist2()[t * 256 + sub], 1u);
            }
        }
    }
    plan_sync(NT);
    // planner warp w resolves the ends w, w + (number of planner warps), ...
    unsigned heavy_any = 0;
    for (int t = ptid >> 5; t < F_NT; t += NT >> 5) {
        const int tb = ps->tbin[t];
        if (tb >= 0) {
            int bin;
            unsigned before, cnt, total;
            const bool ok = warp_find_bin(sm.hist2() + t * 256, ps->trank[t], lane, bin, before, cnt, total);
            if (lane == 0) {
                if (!ok) {   // cannot happen: rank r lies in this bin by construction
                    bin = (t & 1) ? (int)submask : 0;
                    before = (t & 1) ? total : 0;
                    cnt = 0;
                }
                ps->tsub[t] = (unsigned)bin;
                ps->tcum[t] = sm.cum()[tb] + before;
                ps->tcnt[t] = cnt;
            }
        }
    }
    plan_sync(NT);
    FT_MARK(12);   // round 2
    // ---- heavy tie levels: a sub-bin at an interval end that holds many sample points may be one repeated value
    // (an integer-valued gauge, a clamp).  It is taken for one when all its sample points have the same high key and
    // its first and its last point have the same low word too (the low words are not kept in shared memory: two
    // points are fetched again, not all of them).  A wrong guess costs candidates, never correctness: samples that do
    // not equal the listed key are stored like any other.
    const unsigned heavy_min = 16;
#pragma unroll
    for (int t = 0; t < F_NT; ++t) heavy_any |= (ps->tbin[t] >= 0 && ps->tcnt[t] >= heavy_min) ? (1u << t) : 0u;
    if (heavy_any) {
#pragma unroll 1
        for (int idx = ptid; idx < S; idx += NT) {
            const uint32_t k = sk[idx];
            if (k == HK_NONE) continue;
            const uint32_t d = k - hkmin;
            unsigned m = sm.lut()[d >> sh1] & heavy_any;
            const uint32_t sub = (d >> sh2) & submask;
            while (m) {
                const int t = __ffs(m) - 1;
                m &= m - 1;
                if (ps->tsub[t] == sub) {
                    atomicMin(&ps->hklo[t], k);
                    atomicMax(&ps->hkhi[t], k);
                    atomicMin(&ps->hifirst[t], (unsigned)idx);
                    atomicMax(&ps->hilast[t], (unsigned)idx);
                }
            }
        }
        plan_sync(NT);
        if (ptid < F_NT && ((heavy_any >> ptid) & 1u) && ps->hklo[ptid] == ps->hkhi[ptid]) {
            const int i0 = (int)ps->hifirst[ptid], i1 = (int)ps->hilast[ptid];
            const unsigned long long k0 = ko_okey(v[quad_pos(i0 >> 2, L, nquad, lead4) + (i0 & 3)] + 0.0);
            const unsigned long long k1 = ko_okey(v[quad_pos(i1 >> 2, L, nquad, lead4) + (i1 & 3)] + 0.0);
            ps->hmin[ptid] = k0 < k1 ? k0 : k1;
            ps->hmax[ptid] = k0 < k1 ? k1 : k0;
        }
        plan_sync(NT);
    }
    FT_MARK(13);   // heavy levels
    // ---- thread 0: intervals (quantiles ascend, so do their interval ends), heavy list, capacity check
    if (ptid == 0) {
        int ni = 0;
        uint32_t cA = 0, cB = 0;
        unsigned ca = 0, cb = 0, pts = 0;
#pragma unroll
        for (int j = 0; j < NI_MAX; ++j) {
            if (j < nq && (qsel < 0 || j == qsel)) {
                const int ta = 2 * j, tb = 2 * j + 1;
                uint32_t A = 0u, B = HK_PINF;
                unsigned ra = 0u, rb = ns;
                if (ps->tbin[ta] >= 0) {
                    A = hkmin + ((uint32_t)ps->tbin[ta] << sh1) + (ps->tsub[ta] << sh2);
                    ra = ps->tcum[ta];
                }
                if (ps->tbin[tb] >= 0) {
                    const unsigned long long bb = (unsigned long long)hkmin + ((unsigned long long)ps->tbin[tb] << sh1) +
                                                  ((unsigned long long)ps->tsub[tb] << sh2) + ((1ull << sh2) - 1ull);
                    B = bb > (unsigned long long)HK_PINF ? HK_PINF : (uint32_t)bb;
                    rb = ps->tcum[tb] + ps->tcnt[tb];
                }
                if (B < A) B = A;   // cannot happen (ranks ascend); keep the interval well formed
                if (ni > 0 && (A <= cB || A - cB == 1u)) {   // overlaps or touches the current interval: widen it
                    if (A < cA) {
                        cA = A;
                        ca = ra;
                    }
                    if (B > cB) {
                        cB = B;
                        cb = rb;
                    }
                } else {
                    if (ni > 0) {
                        sl->A[ni - 1] = cA;
                        sl->W[ni - 1] = cB - cA;
                        pts += cb - ca;
                    }
                    ++ni;
                    cA = A;
                    cB = B;
                    ca = ra;
                    cb = rb;
                }
                sl->qiv[j] = ni - 1;
            }
        }
        if (ni > 0) {
            sl->A[ni - 1] = cA;
            sl->W[ni - 1] = cB - cA;
            pts += cb - ca;
        }
        for (int i = ni; i < NI_MAX; ++i) {
            sl->A[i] = HK_UNUSED;
            sl->W[i] = 0u;
        }
        // The scan works on the samples themselves (FP64 pipe), so every interval is also handed over as the smallest
        // and the largest double whose high key lies in [A, A + W].  Key order is value order with -0 = +0, which is
        // exactly how doubles compare; high keys at or beyond the infinities' are clamped to the infinities.
        for (int i = 0; i < NI_MAX; ++i) {
            const uint32_t ka = sl->A[i], kb = ka + sl->W[i];
            const double qnan = __longlong_as_double(0x7FF8000000000000LL);
            sl->Ad[i] = i >= ni ? qnan : (ka <= HK_NINF ? -pinf : ko_key2d((unsigned long long)ka << 32));
            sl->Bd[i] = i >= ni ? qnan : (kb >= HK_PINF ? pinf : ko_key2d(((unsigned long long)kb << 32) | 0xFFFFFFFFull));
        }
        int nh = 0;
        unsigned heavy_pts = 0;
        for (int t = 0; t < F_NT; ++t) {
            if (!((heavy_any >> t) & 1u) || ps->hmin[t] != ps->hmax[t]) continue;
            bool dup = false;
            for (int h = 0; h < nh; ++h) dup = dup || (sl->heavy[h] == ps->hmin[t]);
            if (dup || nh >= F_NH) continue;
            sl->heavy[nh++] = ps->hmin[t];
            heavy_pts += ps->tcnt[t];
        }
        pts = pts > heavy_pts ? pts - heavy_pts : 0u;
        const PoolWin pw = pool_window(single, kser, a.spill_cap);
        const double expect = (double)pts * (double)L / (double)S * 1.25 + 256.0;
        sl->ni = ni;
        sl->nheavy = nh;
        {   // hash table of the listed levels
            uint32_t mul = 0u;
            for (int tr = 0; tr < 16 && mul == 0u && nh > 0; ++tr) {
                const uint32_t m = (0x9E3779B1u + (uint32_t)tr * 0x632BE5ABu) | 1u;
                unsigned used = 0u;
                bool ok = true;
                for (int h = 0; h < nh; ++h) {
                    const unsigned bit = 1u << heavy_slot(sl->heavy[h], m);
                    ok = ok && !(used & bit);
                    used |= bit;
                }
                if (ok) mul = m;
            }
            for (int i = 0; i < 16; ++i) sl->htab[i] = ~0ull;
            if (mul != 0u)
                for (int h = 0; h < nh; ++h) {
                    const uint32_t i = heavy_slot(sl->heavy[h], mul);
                    sl->htab[i] = sl->heavy[h];
                    sl->htab_h[i] = (unsigned char)h;
                }
            sl->hmul = mul;
        }
        sl->mn_seed = ko_key2d(kmin64);
        sl->mx_seed = ko_key2d(kmax64);
        if (expect > (double)(pw.s_cap + pw.g_cap)) {
#ifdef KO_FLOW_DEBUG
            printf("plan deferred (capacity): sid %d L %d pts %u heavy_pts %u S %d ns %u expect %f cap %u ni %d | A %x W %x | A %x W %x | "
                   "A %x W %x | tcum %u %u %u %u %u %u tbin %d %d %d %d %d %d\n", sl->sid, L, pts,
                   heavy_pts, S, ns, expect, pw.s_cap + pw.g_cap, ni, sl->A[0], sl->W[0], sl->A[1], sl->W[1], sl->A[2],
                   sl->W[2], ps->tcum[0], ps->tcum[1], ps->tcum[2], ps->tcum[3], ps->tcum[4], ps->tcum[5], ps->tbin[0],
                   ps->tbin[1], ps->tbin[2], ps->tbin[3], ps->tbin[4], ps->tbin[5]);
#endif
            sl->defer = 1;
            sl->ni = 0;
            sl->nheavy = 0;
        }
    }
    FT_MARK(14);   // thread 0
    plan_sync(NT);
}

// ------------------------------------------------------------------------------------------------ stream
struct StreamAcc {
    double s0, s1, mn, mx;   // mn / mx: the warp's running extremes (equal in all lanes)
    unsigned nan;
    unsigned c[NI_MAX];      // samples below interval i
};

// One sample against NI intervals, hand-written so that every comparison result stays a predicate (the compiler
// turns them into SEL / LOP3 / ISETP sequences on the integer pipe, which is the pipe this kernel must keep free):
// per interval  pb = x < Ad;  @pb c += 1;  q_i = (x <= Bd) && !pb;   then  @(q_0 | ... ) m |= bit.
template <int NI>
__device__ __forceinline__ void sample_intervals(double x, const double (&Ad)[NI_MAX], const double (&Bd)[NI_MAX],
                                                 unsigned (&c)[NI_MAX], uint32_t& m, uint32_t bit) {
    if (NI == 1) {
        asm("{\n\t.reg .pred pb, q0;\n\t"
            "setp.lt.f64 pb, %2, %3;\n\t@pb add.u32 %0, %0, 1;\n\tsetp.le.and.f64 q0, %2, %4, !pb;\n\t"
            "@q0 or.b32 %1, %1, %5;\n\t}"
            : "+r"(c[0]), "+r"(m)
            : "d"(x), "d"(Ad[0]), "d"(Bd[0]), "r"(bit));
    } else if (NI == 2) {
        asm("{\n\t.reg .pred pb, q0, q1;\n\t"
            "setp.lt.f64 pb, %3, %4;\n\t@pb add.u32 %0, %0, 1;\n\tsetp.le.and.f64 q0, %3, %5, !pb;\n\t"
            "setp.lt.f64 pb, %3, %6;\n\t@pb add.u32 %1, %1, 1;\n\tsetp.le.and.f64 q1, %3, %7, !pb;\n\t"
            "or.pred q0, q0, q1;\n\t"
            "@q0 or.b32 %2, %2, %8;\n\t}"
            : "+r"(c[0]), "+r"(c[1]), "+r"(m)
            : "d"(x), "d"(Ad[0]), "d"(Bd[0]), "d"(Ad[1]), "d"(Bd[1]), "r"(bit));
    } else if (NI == 3) {
        asm("{\n\t.reg .pred pb, q0, q1, q2;\n\t"
            "setp.lt.f64 pb, %4, %5;\n\t@pb add.u32 %0, %0, 1;\n\tsetp.le.and.f64 q0, %4, %6, !pb;\n\t"
            "setp.lt.f64 pb, %4, %7;\n\t@pb add.u32 %1, %1, 1;\n\tsetp.le.and.f64 q1, %4, %8, !pb;\n\t"
            "setp.lt.f64 pb, %4, %9;\n\t@pb add.u32 %2, %2, 1;\n\tsetp.le.and.f64 q2, %4, %10, !pb;\n\t"
            "or.pred q0, q0, q1;\n\tor.pred q0, q0, q2;\n\t"
            "@q0 or.b32 %3, %3, %11;\n\t}"
            : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(m)
            : "d"(x), "d"(Ad[0]), "d"(Bd[0]), "d"(Ad[1]), "d"(Bd[1]), "d"(Ad[2]), "d"(Bd[2]), "r"(bit));
    } else if (NI == 4) {
        asm("{\n\t.reg .pred pb, q0, q1, q2, q3;\n\t"
            "setp.lt.f64 pb, %5, %6;\n\t@pb add.u32 %0, %0, 1;\n\tsetp.le.and.f64 q0, %5, %7, !pb;\n\t"
            "setp.lt.f64 pb, %5, %8;\n\t@pb add.u32 %1, %1, 1;\n\tsetp.le.and.f64 q1, %5, %9, !pb;\n\t"
            "setp.lt.f64 pb, %5, %10;\n\t@pb add.u32 %2, %2, 1;\n\tsetp.le.and.f64 q2, %5, %11, !pb;\n\t"
            "setp.lt.f64 pb, %5, %12;\n\t@pb add.u32 %3, %3, 1;\n\tsetp.le.and.f64 q3, %5, %13, !pb;\n\t"
            "or.pred q0, q0, q1;\n\tor.pred q2, q2, q3;\n\tor.pred q0, q0, q2;\n\t"
            "@q0 or.b32 %4, %4, %14;\n\t}"
            : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3]), "+r"(m)
            : "d"(x), "d"(Ad[0]), "d"(Bd[0]), "d"(Ad[1]), "d"(Bd[1]), "d"(Ad[2]), "d"(Bd[2]), "d"(Ad[3]), "d"(Bd[3]), "r"(bit));
    }
}
// mn <= x && x <= mx for four samples (one predicate chain; false for NaN)
__device__ __forceinline__ uint32_t plain4(double x0, double x1, double x2, double x3, double mn, double mx) {
    uint32_t ok;
    asm("{\n\t.reg .pred p;\n\t"
        "setp.ge.f64 p, %1, %5;\n\tsetp.le.and.f64 p, %1, %6, p;\n\t"
        "setp.ge.and.f64 p, %2, %5, p;\n\tsetp.le.and.f64 p, %2, %6, p;\n\t"
        "setp.ge.and.f64 p, %3, %5, p;\n\tsetp.le.and.f64 p, %3, %6, p;\n\t"
        "setp.ge.and.f64 p, %4, %5, p;\n\tsetp.le.and.f64 p, %4, %6, p;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "d"(x0), "d"(x1), "d"(x2), "d"(x3), "d"(mn), "d"(mx));
    return ok;
}

// Two warp steps at once: 8 samples per lane.  NI intervals.  addr: the lane's 16 bytes of the first step.  Slots of a
// stage that lie outside the series have been overwritten with NaN before (patch_stage_edges), so every slot is
// either a sample or NaN here.  Everything per sample is a double comparison (DSETP on the FP64 pipe, which is
// otherwise idle: the integer pipe was the bound of this kernel) feeding a predicated add or a predicate:
//   below interval i:   x < Ad[i]                        -> c[i] += 1
//   inside interval i:  !(x < Ad[i]) && x <= Bd[i]       -> flag bit
//   nothing special:    mn <= x && x <= mx               (false for NaN, a new extreme, an infinity beyond them)
// NaN compares false everywhere, so stale samples are neither below nor inside anything.
template <int NI>
__device__ __forceinline__ uint32_t stream_step2(uint32_t addr, StreamAcc& acc, const double (&Ad)[NI_MAX],
                                                 const double (&Bd)[NI_MAX]) {
    double x[8];
    lds_v2f64(addr, x[0], x[1]);
    lds_v2f64(addr + 512u, x[2], x[3]);
    lds_v2f64(addr + 1024u, x[4], x[5]);
    lds_v2f64(addr + 1536u, x[6], x[7]);
    const bool plain = (plain4(x[0], x[2], x[4], x[6], acc.mn, acc.mx) & plain4(x[1], x[3], x[5], x[7], acc.mn, acc.mx)) != 0u;
    uint32_t m = 0u;   // bit e: sample e of this lane lies in one of the intervals
    if (NI > 0) {
#pragma unroll
        for (int e = 0; e < 8; ++e) sample_intervals<NI>(x[e], Ad, Bd, acc.c, m, 1u << e);
    }
    if (__any_sync(0xffffffffu, !plain)) {
        // Somewhere in the warp: a NaN (stale sample) or a new extreme.  The running extremes are the WARP's (kept
        // equal in all lanes), so new extremes die out like 1 / samples seen.
        bool changed = false;
        if (!plain) {
#pragma unroll
            for (int e = 0; e < 8; ++e) {
                if (x[e] != x[e]) {
                    acc.nan += 1u;
                    x[e] = 0.0;
                } else {
                    if (x[e] < acc.mn) {
                        acc.mn = x[e];
                        changed = true;
                    }
                    if (x[e] > acc.mx) {
                        acc.mx = x[e];
                        changed = true;
                    }
                }
            }
        }
        if (__any_sync(0xffffffffu, changed)) {
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                acc.mn = fmin(acc.mn, __shfl_xor_sync(0xffffffffu, acc.mn, o));
                acc.mx = fmax(acc.mx, __shfl_xor_sync(0xffffffffu, acc.mx, o));
            }
        }
    }
    // (pairwise inside the step, two running sums: short dependent chains without more live registers)
    acc.s0 += (x[0] + x[2]) + (x[4] + x[6]);
    acc.s1 += (x[1] + x[3]) + (x[5] + x[7]);
    return m;
}

// One stage: returns the lane's 16 in-interval flags.  Bit b = t * 8 + e flags the sample at stage offset
// (b >> 1) * 64 + lane * 2 + (b & 1): sample e of step t sits at t * 256 + (e >> 1) * 64 + lane * 2 + (e & 1).
template <int NI>
__device__ __forceinline__ uint32_t stream_stage(uint32_t stage_addr, StreamAcc& acc, const double (&A)[NI_MAX],
                                                 const double (&W)[NI_MAX], int lane) {
    uint32_t m = 0u;
#pragma unroll 1
    for (int t = 0; t < 2; ++t)
        m |= stream_step2<NI>(stage_addr + (uint32_t)t * 2048u + (uint32_t)lane * 16u, acc, A, W) << (t * 8);
    return m;
}
// byte offset (from the lane's first sample of the stage) of the sample that flag bit b stands for
__device__ __forceinline__ uint32_t flag_offset(int b) { return (uint32_t)(b >> 1) * 512u + (uint32_t)(b & 1) * 8u; }

// The flagged samples of one scanned stage go to the series' candidate pool as 64-bit keys: room for all of the
// warp's candidates is reserved with one atomic on the slot's cursor (shared-memory window first - when the finish
// warps have released it - then the global spill window), then every lane copies its own.  Samples on a listed tie
// level are counted instead.  flags: bit 0 = the shared-memory window is ours, bit 1 = something did not fit,
// bit 2 = something went to the spill, bit 3 = short series: wait for the window rather than spill.
__device__ __forceinline__ void store_candidates(const SeriesArgs& a, const FlowSmem& sm, bool single, uint32_t kser,
                                                 uint32_t stage_addr, uint32_t m, int lane, uint32_t& flags) {
    // (everything but the flags is looked up again here: the scan loop around this call has no registers to spare)
    const PoolWin pw = pool_window(single, kser, a.spill_cap);
    const FlowSlot* sl = sm.slot(pw.d);
    unsigned long long* spill = a.spill + (size_t)blockIdx.x * (size_t)a.spill_cap;
    const int nheavy = sl->nheavy;
    if (nheavy > 0) {
        uint32_t rest = m;
        const uint32_t hmul = sl->hmul;
        if (hmul != 0u) {   // one probe of the 16-slot table per flagged sample
#pragma unroll 1
            while (rest) {
                const int b = __ffs(rest) - 1;
                rest &= rest - 1;
                const unsigned long long key = ko_okey(lds_f64(stage_addr + (uint32_t)lane * 16u + flag_offset(b)) + 0.0);
                const uint32_t i = heavy_slot(key, hmul);
                if (sl->htab[i] == key) {
                    red_shared_add(&sm.hcnt()[pw.d * F_NH + sl->htab_h[i]], 1u);
                    m &= ~(1u << b);
                }
            }
        } else {
#pragma unroll 1
            while (rest) {
                const int b = __ffs(rest) - 1;
                rest &= rest - 1;
                const unsigned long long key = ko_okey(lds_f64(stage_addr + (uint32_t)lane * 16u + flag_offset(b)) + 0.0);
                int hit = -1;
#pragma unroll 1
                for (int h = 0; h < nheavy; ++h) hit = sl->heavy[h] == key ? h : hit;
                if (hit >= 0) {
                    red_shared_add(&sm.hcnt()[pw.d * F_NH + hit], 1u);
                    m &= ~(1u << b);
                }
            }
        }
        if (!__any_sync(0xffffffffu, m != 0u)) return;
    }
    const unsigned n = __popc(m);
    if (!(flags & 1u)) {   // shared-memory window not released yet: spill, or wait when the spill is full
        volatile unsigned* rel = reinterpret_cast<volatile unsigned*>(&sm.rel()[pw.win]);
        bool ok = *rel >= pw.use;
        // Short series (their finish is quick and is what bounds the CTA): wait, so that the finish warps find every
        // candidate in shared memory instead of reading the spill back through L2.
        if (!ok && ((flags & 8u) || *reinterpret_cast<volatile unsigned*>(&sm.cgl()[pw.d]) + 512u > pw.g_cap)) {
            while (*rel < pw.use) __nanosleep(200);
            ok = true;
        }
        if (ok) {
            flags |= 1u;
            __threadfence_block();   // (only when the window changes hands)
        }
    }
    if (n == 0u) return;
    // Every lane reserves room for its own candidates: one atomic instruction for the warp (lanes that hit the same
    // cursor are serialised by the hardware, which is cheaper than a prefix sum over the warp plus a broadcast).
    // Shared-memory entries first (when the window is ours), whatever does not fit goes to the spill.
    unsigned sbase = pw.s_cap, gbase_i = 0;
    if (flags & 1u) sbase = atom_shared_add(&sm.csm()[pw.d], n);
    const unsigned take = sbase < pw.s_cap ? min(n, pw.s_cap - sbase) : 0u;
    if (take < n) {
        gbase_i = atom_shared_add(&sm.cgl()[pw.d], n - take);
        flags |= 4u;
    }
    unsigned long long* wb = sm.pool() + pw.s_base + sbase;
    unsigned long long* gb = spill + pw.g_base + gbase_i;
    const uint32_t lane_addr = stage_addr + (uint32_t)lane * 16u;
    const bool to_pool = take == n, to_spill = take == 0u && gbase_i + n <= pw.g_cap;
    if (to_pool || to_spill) {   // all of the lane's candidates go to one place (nearly always)
        unsigned long long dst;
        asm volatile("mov.u64 %0, %1;" : "=l"(dst) : "l"(to_pool ? wb : gb));
#pragma unroll 1
        while (m) {
            const int b = __ffs(m) - 1;
            m &= m - 1;
            st_generic_u64(dst, ko_okey(lds_f64(lane_addr + flag_offset(b)) + 0.0));
            dst += 8ull;
        }
    } else {
        unsigned e = 0;
#pragma unroll 1
        while (m) {
            const int b = __ffs(m) - 1;
            m &= m - 1;
            const unsigned long long key = ko_okey(lds_f64(lane_addr + flag_offset(b)) + 0.0);
            if (e < take) wb[e] = key;
            else if (gbase_i + (e - take) < pw.g_cap) gb[e - take] = key;
            else flags |= 2u;
            ++e;
        }
    }
}

// First / last stage of a series: the slots in front of the series (16-byte alignment) and behind its end hold
// bytes that are not the series'.  They are overwritten with NaN, which the scan skips; the finish warps know how
// many such slots a series has and take them out of the stale-sample count again.
__device__ __forceinline__ void patch_stage_edges(uint32_t stage_addr, int lane, int first_valid, int end_valid) {
    const double nan = __longlong_as_double(0x7FF8000000000000LL);
    for (int i = lane; i < first_valid; i += 32) sts_f64(stage_addr + (uint32_t)i * 8u, nan);
    for (int i = end_valid + lane; i < F_STAGE; i += 32) sts_f64(stage_addr + (uint32_t)i * 8u, nan);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // the next writer of these bytes is a TMA copy
    __syncwarp();
}

// Stage j of a series (ring stage g = g0 + j) belongs to stream warp j % ns and to ring group (g / F_GS) % (ring groups).  A series
// starts on a group boundary; the stage numbers that pad its last group are visited too (wait, release, no work),
// so that every group gets its F_GS releases.
template <int NI>
__device__ __forceinline__ void stream_series(const SeriesArgs& a, const FlowSmem& sm, const FlowSlot* sl, StreamAcc& acc,
                                              int ws, int ns, int lane, bool single, uint32_t kser, uint32_t& flags) {
    double A[NI_MAX], W[NI_MAX];   // interval bounds as doubles (only the first NI are live)
#pragma unroll
    for (int i = 0; i < NI_MAX; ++i) {
        A[i] = sl->Ad[i];
        W[i] = sl->Bd[i];
    }
    const int n = sl->nstages;
    const int npad = (n + F_GS - 1) / F_GS * F_GS;
    const uint32_t ring0 = smem_u32(sm.ring());
    const uint32_t empty0 = smem_u32(sm.bar(FB_EMPTY));
    FT_BEGIN(blockIdx.x == 0 && ws == 0 && lane == 0);
    // (stage j of a series always goes to stream warp j % ns, whatever the CTA did before: the partial sums of a
    // series, and with them the last bits of its sum, do not depend on how the work was distributed among the CTAs -
    // only on ns, i.e. on whether the launch is in the long-series mode)
    for (int j = ws; j < npad; j += ns) {
        const uint32_t g = *reinterpret_cast<const volatile uint32_t*>(&sl->g0) + (uint32_t)j;   // (reloaded: registers)
        const uint32_t grp = g / F_GS;
        uint32_t s, u_unused;
        ring_slot(grp, single, s, u_unused);
        const uint32_t stage_addr = ring0 + s * (uint32_t)F_GROUP_BYTES + (g % F_GS) * (uint32_t)F_STAGE_BYTES;
        if (lane == 0 && ws < 3) FP(5 + ws, 0, 16 + j);
        wait_group_full(sm, grp, single);
        FT_MARK(32);   // stream warp 0: waiting for a full stage
        if (j < n) {
            if (j == 0 || j == n - 1) {
                const int lead = sl->lead;
                const int end_valid = sl->L + lead - j * F_STAGE;
                patch_stage_edges(stage_addr, lane, j == 0 ? lead : 0, end_valid < F_STAGE ? end_valid : F_STAGE);
            }
            const uint32_t m = stream_stage<NI>(stage_addr, acc, A, W, lane);
            FT_MARK(33);   // stream warp 0: scan
            if (NI > 0 && __any_sync(0xffffffffu, m != 0u)) store_candidates(a, sm, single, kser, stage_addr, m, lane, flags);
            FT_MARK(34);   // stream warp 0: storing candidates
        }
        __syncwarp();
        if (lane == 0) mbar_arrive_addr(empty0 + s * 8u);
    }
}

__device__ __forceinline__ double shfl_xor_f64(double v, int o) { return __shfl_xor_sync(0xffffffffu, v, o); }

// ------------------------------------------------------------------------------------------------ finish
// All finish threads (ftid in [0, 128)).  Returns nothing; outputs are written here.
__device__ __forceinline__ void finish_series(const SeriesArgs& a, const FlowSmem& sm, const FlowSlot* sl, int d, int ftid,
                              const PoolWin& pw, const unsigned long long* spill, int ns) {
    FinState* fs = sm.fs();
    const int lane = ftid & 31, fw = ftid >> 5;
    const int nq = a.nq, ni = sl->ni;
    const int L = sl->L, sid = sl->sid;
    const double nan = __longlong_as_double(0x7FF8000000000000LL);
    FT_BEGIN(blockIdx.x == 0 && ftid == 0);
    // ---- A. partial statistics (warp 0, fixed order) and the last valid sample (warp 1)
    if (fw == 0) {
        const FlowPartial* p = sm.part(d, lane < ns ? lane : 0);
        double s = lane < ns ? p->sum : 0.0;
        double mn = p->mn, mx = p->mx;
        unsigned nn = lane < ns ? p->nan : 0u;
        unsigned c[NI_MAX];
#pragma unroll
        for (int i = 0; i < NI_MAX; ++i) c[i] = lane < ns ? p->c[i] : 0u;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            s += shfl_xor_f64(s, o);
            mn = fmin(mn, shfl_xor_f64(mn, o));
            mx = fmax(mx, shfl_xor_f64(mx, o));
        }
        nn = __reduce_add_sync(0xffffffffu, nn);
#pragma unroll
        for (int i = 0; i < NI_MAX; ++i) c[i] = __reduce_add_sync(0xffffffffu, c[i]);
        if (lane == 0) {
            fs->sum = s;
            fs->mn = mn;
            fs->mx = mx;
            // nn counts the NaN slots of all stages: stale samples plus the slots outside the series
            fs->nvalid = (unsigned)L - (nn - ((unsigned)sl->nstages * F_STAGE - (unsigned)L));
#pragma unroll
            for (int i = 0; i < NI_MAX; ++i) fs->cb[i] = c[i];
            fs->fail = sl->defer;
            fs->why = 0;
            fs->pending = 0;
        }
    } else if (ftid == 32) {
        const double* v = a.values + sl->start;
        double lastv = nan;
        for (int i = L - 1; i >= 0; --i) {
            const double x = v[i];
            if (x == x) {
                lastv = x;
                break;
            }
        }
        fs->last = lastv;
    }
    fin_sync();
    FT_MARK(23);   // finish A: partials, last sample
    const unsigned nvalid = fs->nvalid;
    unsigned npool_raw = 0;
    const unsigned nsm_all = ni > 0 ? min(sm.csm()[pw.d], pw.s_cap) : 0u;
    const unsigned ng_all = ni > 0 ? min(sm.cgl()[pw.d], pw.g_cap) : 0u;
    npool_raw = nsm_all + ng_all;
    const bool overflow = ni > 0 && sm.ovf()[pw.d] != 0u;
    const unsigned npool = overflow ? 0u : npool_raw;
    const int nheavy = sl->nheavy;
    // ---- B. per-quantile start state
    if (ftid < NI_MAX) {
        const int j = ftid;
        int st = Q_DONE;
        unsigned long long ans = ~0ull;   // "NaN"
        const int qsel = sl->qsel;
        if (j < nq && qsel >= 0 && j != qsel) {
            // not this work item's quantile: answered by the first attempt or by another item
        } else if (j < nq && nvalid > 0 && !fs->fail) {
            if (sl->konst) {
                if (ko_okey(fs->mn + 0.0) == ko_okey(fs->mx + 0.0)) ans = ko_okey(fs->mn + 0.0);
                else {
                    st = Q_FAIL;
                    fs->why = 4;
                }
            } else if (ni == 0 || overflow) {
                st = Q_FAIL;
                fs->why = overflow ? 3 : 5;
            } else {
                const int iv = sl->qiv[j];
                const unsigned r = nearest_rank_dev(nvalid, a.q[j]);
                const unsigned cb = fs->cb[iv];
                if (r <= cb) {
                    st = Q_FAIL;   // the rank lies below its interval
                    fs->why = 1;
                } else {
                    const unsigned long long range = ((unsigned long long)sl->W[iv] << 32) | 0xFFFFFFFFull;
                    fs->lo[j] = (unsigned long long)sl->A[iv] << 32;
                    fs->hi32[j] = sl->A[iv] + sl->W[iv];
                    fs->sh[j] = max(0, 64 - __clzll((long long)range) - 8);
                    fs->rho[j] = r - cb;
                    st = Q_HIST;
                }
            }
        } else if (j < nq && fs->fail) {
            st = Q_FAIL;
        }
        fs->st[j] = st;
        fs->ans[j] = ans;
        fs->list_n[j] = 0;
        if (st == Q_HIST) atomicAdd(&fs->pending, 1);
        if (st == Q_FAIL) fs->fail = 1;
    }
    fin_sync();
    FT_MARK(24);   // finish B: start state
    // ---- C. rounds: histogram refinement, then direct selection among <= F_LIST candidates
    int guard = 0;
    while (fs->pending > 0 && guard++ < 24) {   // (a quantile that fails does not stop the others)
#ifdef KO_FLOW_TIMING
        if (_ftc) atomicAdd(&g_flow_phase[26], 1000ull);   // rounds (x 1000: the report divides by 1000)
#endif
        for (int i = ftid; i < NI_MAX * 256; i += F_NF * 32) sm.fbins()[i] = 0;
        fin_sync();
        FT_MARK(28);   // finish C: clearing the bins
        int stj[NI_MAX], shj[NI_MAX];
        unsigned long long loj[NI_MAX];
#pragma unroll
        for (int j = 0; j < NI_MAX; ++j) {
            stj[j] = j < nq ? fs->st[j] : Q_DONE;
            shj[j] = fs->sh[j];
            loj[j] = fs->lo[j];
        }
        // 32-bit pre-filter per quantile: the key's high word against [lo >> 32, (end of the range) >> 32]; almost
        // every candidate belongs to one quantile's range only, so the 64-bit arithmetic runs once per key
        uint32_t lo32[NI_MAX], span32[NI_MAX];
#pragma unroll
        for (int j = 0; j < NI_MAX; ++j) {
            const int she = stj[j] <= Q_COLLECT ? (shj[j] & 63) : 0;
            const unsigned long long width = stj[j] == Q_HIST ? (255ull << she) | ((1ull << she) - 1ull)
                                                              : (1ull << she) - 1ull;
            unsigned long long end = loj[j] + width;
            if (end < loj[j]) end = ~0ull;
            // The 256 bins of a round may reach beyond the interval's own end (an interval that is open below starts
            // with bins of 2^56 keys): candidates of the NEXT interval must not be counted there, or a rank that lies
            // above this interval would be answered from the neighbour's candidates instead of being reported.
            uint32_t e32 = (uint32_t)(end >> 32);
            e32 = e32 < fs->hi32[j] ? e32 : fs->hi32[j];
            lo32[j] = (uint32_t)(loj[j] >> 32);
            span32[j] = stj[j] <= Q_COLLECT ? e32 - lo32[j] : 0u;
            if (stj[j] > Q_COLLECT) lo32[j] = 0xFFFFFFFFu;   // no high word passes: (hk - lo32) wraps above span
        }
        auto visit = [&](unsigned long long key, unsigned mult) {
            const uint32_t hk = (uint32_t)(key >> 32);
#pragma unroll
            for (int j = 0; j < NI_MAX; ++j) {
                if (hk - lo32[j] > span32[j]) continue;   // (also when hk < lo32: the difference wraps)
                if (key < loj[j]) continue;
                const unsigned long long dd = (key - loj[j]) >> shj[j];
                if (stj[j] == Q_HIST) {
                    if (dd < 256ull) red_shared_add(&sm.fbins()[j * 256 + (unsigned)dd], mult);
                } else if (dd == 0ull) {
                    const unsigned pos = atom_shared_add(&fs->list_n[j], 1u);
                    if (pos < (unsigned)F_LIST) {
                        sm.flistk()[j * F_LIST + pos] = key;
                        sm.flistm()[j * F_LIST + pos] = mult;
                    }
                }
            }
        };
        {
            const unsigned nsm = npool > 0u ? nsm_all : 0u;
            const unsigned ng = npool > 0u ? ng_all : 0u;
            const unsigned long long* wbase = sm.pool() + pw.s_base;
#pragma unroll 1
            for (unsigned i = ftid; i < nsm; i += F_NF * 32) visit(wbase[i], 1u);
            FT_MARK(29);   // finish C: candidates in shared memory
            if (ng > 0u) {   // spilled part: global loads, four in flight
                const unsigned long long* gbase = spill + pw.g_base;
#pragma unroll 1
                for (unsigned i = ftid; i < ng; i += 4u * (F_NF * 32)) {
                    unsigned long long kk[4];
#pragma unroll
                    for (int u = 0; u < 4; ++u)
                        kk[u] = i + u * (F_NF * 32) < ng ? ld_cg_u64(gbase + i + u * (F_NF * 32)) : ~0ull;
#pragma unroll 1
                    for (int u = 0; u < 4; ++u)
                        if (i + u * (F_NF * 32) < ng) visit(u == 0 ? kk[0] : u == 1 ? kk[1] : u == 2 ? kk[2] : kk[3], 1u);
                }
            }
        }
        if (ftid < nheavy) {
            const unsigned mult = sm.hcnt()[pw.d * F_NH + ftid];
            if (mult) visit(sl->heavy[ftid], mult);
        }
        fin_sync();
        FT_MARK(27);   // finish C: visiting the candidates
        // warp j settles quantile j
        if (fw < nq && fs->st[fw] <= Q_COLLECT) {
            const int j = fw;
            if (fs->st[j] == Q_HIST) {
                int bin;
                unsigned before, cnt, total;
                const bool ok = warp_find_bin(sm.fbins() + j * 256, fs->rho[j], lane, bin, before, cnt, total);
                if (!ok) {   // the rank lies above its interval
                    if (lane == 0) {
                        fs->st[j] = Q_FAIL;
                        fs->fail = 1;
                        fs->why = 2;
                        atomicSub(&fs->pending, 1);
                    }
                } else {
                    const int sh = fs->sh[j];
                    const unsigned long long nlo = fs->lo[j] + ((unsigned long long)bin << sh);
                    // one listed tie level filling the whole bin settles it
                    bool hv = false;
                    if (lane < nheavy) {
                        const unsigned long long hk = sl->heavy[lane];
                        // (inside the bin AND inside the interval: only then was it counted into this bin)
                        hv = hk >= nlo && ((hk - nlo) >> sh) == 0ull && (uint32_t)(hk >> 32) <= fs->hi32[j] &&
                             sm.hcnt()[pw.d * F_NH + lane] == cnt;
                    }
                    const unsigned hb = __ballot_sync(0xffffffffu, hv);
                    if (lane == 0) {
                        fs->lo[j] = nlo;
                        fs->rho[j] -= before;
                        if (sh == 0) {
                            fs->ans[j] = nlo;
                            fs->st[j] = Q_DONE;
                            atomicSub(&fs->pending, 1);
                        } else if (hb) {
                            fs->ans[j] = sl->heavy[__ffs(hb) - 1];
                            fs->st[j] = Q_DONE;
                            atomicSub(&fs->pending, 1);
                        } else if (cnt <= (unsigned)F_LIST) {
                            fs->st[j] = Q_COLLECT;     // keys with (key - lo) >> sh == 0
                            fs->list_n[j] = 0;
                        } else {
                            fs->sh[j] = sh >= 8 ? sh - 8 : 0;
                        }
                    }
                }
            } else {   // Q_COLLECT: rank counting among the listed (key, multiplicity) pairs
                const unsigned n = min(fs->list_n[j], (unsigned)F_LIST);
                const unsigned rho = fs->rho[j];
                const unsigned long long* lk = sm.flistk() + j * F_LIST;
                const unsigned* lm = sm.flistm() + j * F_LIST;
                unsigned long long found = 0ull;
                bool got = false;
                for (unsigned c0 = 0; c0 < n; c0 += 32) {
                    const unsigned c = c0 + lane;
                    unsigned less = 0, eq = 0;
                    const unsigned long long kc = c < n ? lk[c] : 0ull;
                    for (unsigned e = 0; e < n; ++e) {
                        const unsigned long long ke = lk[e];
                        const unsigned me = lm[e];
                        less += ke < kc ? me : 0u;
                        eq += ke == kc ? me : 0u;
                    }
                    const bool hit = c < n && less < rho && rho <= less + eq;
                    const unsigned hb = __ballot_sync(0xffffffffu, hit);
                    if (hb && !got) {
                        found = __shfl_sync(0xffffffffu, kc, __ffs(hb) - 1);
                        got = true;
                    }
                }
                if (lane == 0) {
                    if (got) {
                        fs->ans[j] = found;
                        fs->st[j] = Q_DONE;
                    } else {   // cannot happen: the bin count said the rank is among these
                        fs->st[j] = Q_FAIL;
                        fs->fail = 1;
                        fs->why = 6;
#ifdef KO_FLOW_DEBUG
                        printf("select failed: sid %d L %d q %d n %u list_n %u rho %u sh %d lo %llx nheavy %d\n", sid, L, j, n,
                               fs->list_n[j], rho, fs->sh[j], fs->lo[j], nheavy);
                        for (unsigned e = 0; e < n && e < 8; ++e) printf("   key %llx mult %u\n", lk[e], lm[e]);
#endif
                    }
                    atomicSub(&fs->pending, 1);
                }
            }
        }
        fin_sync();
    }
    FT_MARK(25);   // finish C: rounds
    // ---- D. outputs
    if (ftid == 0) {
        const bool fail = fs->fail != 0 || fs->pending > 0;
        if (a.out_sum) a.out_sum[sid] = fs->sum;
        if (a.out_mean) a.out_mean[sid] = nvalid > 0 ? fs->sum / (double)nvalid : 0.0;
        if (a.out_min) a.out_min[sid] = nvalid > 0 ? fs->mn + 0.0 : nan;
        if (a.out_max) a.out_max[sid] = nvalid > 0 ? fs->mx + 0.0 : nan;
        if (a.out_last) a.out_last[sid] = fs->last;
        if (a.out_count) a.out_count[sid] = (int64_t)nvalid;
        const int qsel = sl->qsel;
        if (fail) atomicAdd(&a.counters[9 + (fs->why & 7)], 1);
        if (fail && (a.wide || !a.pack)) {   // whole series to the next list
            const int pos = atomicAdd(&a.counters[a.out_cnt], 1);
            a.out_list[pos] = sid;
        } else {
            // first attempt: every failed quantile becomes a work item (quantile, series) of the second attempt and
            // the others are answered here.  A pool that overflowed: all quantiles, one item each.  A failure the
            // second attempt cannot mend (almost no valid points, a spread below the key resolution, a constant
            // pre-sample of a series that is not constant): code 6, straight on to the slow path.
            const bool direct = fail && (fs->why == 0 || fs->why == 4 || fs->why == 5);
            const bool all = direct || (fail && fs->why == 3);
            if (direct) {
                const int pos = atomicAdd(&a.counters[a.out_cnt], 1);
                a.out_list[pos] = sid | (6 << 27);
            } else if (fail) {
                for (int j = 0; j < nq; ++j) {
                    if (all || fs->st[j] != Q_DONE) {
                        const int pos = atomicAdd(&a.counters[a.out_cnt], 1);
                        a.out_list[pos] = sid | (j << 27);
                    }
                }
            }
            if (a.out_quant && !all) {
                for (int j = 0; j < nq; ++j) {
                    if ((qsel >= 0 && j != qsel) || (fail && fs->st[j] != Q_DONE)) continue;   // (not ours / deferred)
                    a.out_quant[(int64_t)sid * nq + a.qcol[j]] = fs->ans[j] == ~0ull ? nan : ko_key2d(fs->ans[j]);
                }
            }
        }
    }
    if (ftid < F_NH) sm.hcnt()[pw.d * F_NH + ftid] = 0u;
    if (ftid == 32) {
        sm.csm()[pw.d] = 0u;
        sm.cgl()[pw.d] = 0u;
    }
    if (ftid == 64) sm.ovf()[pw.d] = 0u;
    fin_sync();
}

// ------------------------------------------------------------------------------------------------ kernel
__global__ void __launch_bounds__(F_THREADS, 1) series_flow(SeriesArgs a) {
    extern __shared__ __align__(128) unsigned char flow_smem[];
    FlowSmem sm;
    sm.base = flow_smem;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const bool single = a.wide != 0 || a.counters[8] > F_DOUBLE_MAXL;
    const int ns = single ? F_NS_LONG : F_NS_SHORT, np = single ? F_NP_LONG : F_NP_SHORT;
    const int w_fin = ns, w_plan = ns + F_NF;
    if (tid == 0) {
        for (int s = 0; s < F_NGROUPS_MAX; ++s) {
            mbar_init(sm.bar(FB_FULL + s), 1);
            mbar_init(sm.bar(FB_EMPTY + s), F_GS);   // the four warps that scan the group's stages
        }
        for (int d = 0; d < F_ND; ++d) {
            mbar_init(sm.bar(FB_DESC + d), 1);
            mbar_init(sm.bar(FB_PLAN + d), 1);
            mbar_init(sm.bar(FB_FREE + d), 1);
            mbar_init(sm.bar(FB_SDONE + d), (uint32_t)ns);
            mbar_init(sm.bar(FB_START + d), 1);
        }
        mbar_fence_init();
    }
    if (tid < F_PSTATE_WORDS) sm.csm()[tid] = 0u;
#ifdef KO_FLOW_PROGRESS
    if (tid == 0) g_flow_wide = (unsigned)a.wide;
    if (tid < 8) reinterpret_cast<volatile unsigned*>(g_flow_prog)[blockIdx.x * 8 + tid] = ((unsigned)a.wide << 31) | 0xFFFu;
#endif
    __syncthreads();
    const int n_work = a.counters[a.work_cnt];
    unsigned long long* spill = a.spill + (size_t)blockIdx.x * (size_t)a.spill_cap;

    if (warp == F_W_PROD) {
        // =============================== PRODUCER ===============================
        if (lane == 0) {
            const uint64_t policy = createpolicy_evict_first();
            const uint32_t ring0 = smem_u32(sm.ring());
            FT_BEGIN(blockIdx.x == 0);
            for (uint32_t k = 0;; ++k) {
                const int d = (int)(k % F_ND);
                FP(0, k, 1);
                mbar_wait_sleep<200>(sm.bar(FB_DESC + d), (k / F_ND) & 1u);
                FP(0, k, 2);
                FT_MARK(0);   // producer: waiting for a descriptor
                const FlowSlot* sl = sm.slot(d);
                if (sl->sid < 0) break;
                const int n = sl->nstages;
                const int ngroups = (n + F_GS - 1) / F_GS;
                const uint32_t grp0 = sl->g0 / F_GS;
                const uint32_t total = sl->total_bytes;
                const unsigned char* src = reinterpret_cast<const unsigned char*>(sl->src);
                for (int gi = 1; gi < F_PF && gi < ngroups; ++gi) {
                    const uint32_t off = (uint32_t)gi * F_GROUP_BYTES;
                    const uint32_t pb = total > off ? min(total - off, (uint32_t)F_GROUP_BYTES) : 0u;
                    if (pb) bulk_prefetch_l2(src + off, pb);
                }
                for (int gi = 0; gi < ngroups; ++gi) {
                    const uint32_t grp = grp0 + (uint32_t)gi;
                    uint32_t s, u;
                    ring_slot(grp, single, s, u);
                    FT_MARK(2);   // producer: issuing
                    FP(0, k, 16 + gi);
                    if (u > 0) mbar_wait(sm.bar(FB_EMPTY + s), (u - 1u) & 1u);
                    FT_MARK(1);   // producer: waiting for a free group
                    const uint32_t off = (uint32_t)gi * F_GROUP_BYTES;
                    const uint32_t bytes = total > off ? min(total - off, (uint32_t)F_GROUP_BYTES) : 0u;
                    const uint32_t dst = ring0 + s * (uint32_t)F_GROUP_BYTES;
                    if (gi + F_PF < ngroups) {
                        const uint32_t poff = (uint32_t)(gi + F_PF) * F_GROUP_BYTES;
                        const uint32_t pb = total > poff ? min(total - poff, (uint32_t)F_GROUP_BYTES) : 0u;
                        if (pb) bulk_prefetch_l2(src + poff, pb);
                    }
                    if (gi == ngroups - 1 && sl->tail_fix) {   // the array's very last element, read on its own
                        const int e = sl->L + sl->lead - 1 - gi * (F_GS * F_STAGE);
                        const double x = a.values[sl->start + sl->L - 1];
                        sts_f64(dst + (uint32_t)e * 8u, x);
                    }
                    *sm.issued() = grp + 1u;   // (after the wait for the slot: its previous use is over)
                    if (bytes) {
                        mbar_expect_tx(sm.bar(FB_FULL + s), bytes);
                        tma_bulk_g2s_hint(dst, src + off, bytes, smem_u32(sm.bar(FB_FULL + s)), policy);
                    } else {
                        mbar_arrive(sm.bar(FB_FULL + s));
                    }
                }
            }
        }
    } else if (warp >= w_plan) {
        // =============================== PLANNER ===============================
        const int ptid = tid - w_plan * 32;
        PlanScratch* ps = sm.ps();
        uint32_t g = 0;
        const long long arr_end = a.offsets[a.n_series];
        int prev_depth1 = 0;
        FT_BEGIN(blockIdx.x == 0 && ptid == 0);
        for (uint32_t k = 0;; ++k) {
            const int d = (int)(k % F_ND);
            FT_MARK(6);   // planner: planning
            // slot reuse: the finish of series k - F_ND; long series: stay one series ahead of the stream warps only
            if (ptid == 0) FP(1, k, 1);
            if (k >= (uint32_t)F_ND) mbar_wait_sleep<400>(sm.bar(FB_FREE + d), ((k / F_ND) - 1u) & 1u);
            if (ptid == 0) FP(1, k, 2);
            if (prev_depth1 && k >= 1u) {   // the stream warps have begun series k - 1
                const uint32_t kk = k - 1u;
                mbar_wait_sleep<400>(sm.bar(FB_START + (int)(kk % F_ND)), (kk / F_ND) & 1u);
            }
            if (ptid == 0) FP(1, k, 3);
            FT_MARK(5);   // planner: waiting for the finish warps (slot / depth)
            FlowSlot* sl = sm.slot(d);
            if (ptid == 0) {
                int w, item = 0, code = 7;
                for (;;) {
                    w = atomicAdd(&a.counters[a.work_cur], 1);
                    if (w >= n_work) break;
                    item = a.work_list[w];
                    // second attempt: a work item is (code, series) when series numbers fit 27 bits: code 0..3 = that
                    // quantile alone, 7 = all quantiles, 6 = nothing to gain here, straight on to the slow exact path
                    code = (a.wide && a.pack) ? (int)((unsigned)item >> 27) : 7;
                    if (code != 6) break;
                    const int pos = atomicAdd(&a.counters[a.out_cnt], 1);
                    a.out_list[pos] = item & 0x7FFFFFF;
                }
                if (w >= n_work) {
                    sl->sid = -1;
                } else {
                    const int sid = (a.wide && a.pack) ? (item & 0x7FFFFFF) : item;
                    sl->qsel = code < NI_MAX ? code : -1;
                    const long long off0 = a.offsets[sid];
                    const int L = (int)(a.offsets[sid + 1] - off0);
                    const double* v = a.values + off0;
                    const int lead = (int)((reinterpret_cast<uintptr_t>(v) >> 3) & 1u);
                    const int tot = L + lead;
                    const int n = (tot + F_STAGE - 1) / F_STAGE;
                    const int rem = tot - (n - 1) * F_STAGE;
                    const bool at_end = (off0 + L == arr_end);
                    sl->sid = sid;
                    sl->L = L;
                    sl->lead = lead;
                    sl->nstages = n;
                    sl->g0 = g;
                    sl->tail_fix = ((rem & 1) && at_end) ? 1 : 0;
                    sl->total_bytes = (uint32_t)(n - 1) * F_STAGE_BYTES +
                                      (uint32_t)(sl->tail_fix ? (rem - 1) : ((rem + 1) & ~1)) * 8u;
                    sl->depth1 = L > 16384 ? 1 : 0;
                    sl->src = v - lead;
                    sl->start = off0;
                }
                ps->work = sl->sid;
                mbar_arrive(sm.bar(FB_DESC + d));
            }
            plan_sync(np * 32);
            if (ps->work < 0) {
                if (ptid == 0) mbar_arrive(sm.bar(FB_PLAN + d));
                break;
            }
            g += (uint32_t)((sl->nstages + F_GS - 1) / F_GS * F_GS);   // every series starts a ring group
            prev_depth1 = sl->depth1;
            if (ptid == 0) FP(1, k, 4);
            planner_plan(a, sm, sl, ptid, single, k, np * 32);
            if (ptid == 0) FP(1, k, 5);
            if (ptid == 0) mbar_arrive(sm.bar(FB_PLAN + d));
        }
    } else if (warp >= w_fin) {
        // =============================== FINISH ===============================
        const int ftid = tid - w_fin * 32;
        FT_BEGIN(blockIdx.x == 0 && ftid == 0);
        for (uint32_t k = 0;; ++k) {
            const int d = (int)(k % F_ND);
            const uint32_t par = (k / F_ND) & 1u;
            FT_MARK(22);   // finish: working
            if (ftid == 0) FP(4, k, 1);
            mbar_wait_sleep<400>(sm.bar(FB_PLAN + d), par);
            if (ftid == 0) FP(4, k, 2);
            FT_MARK(20);   // finish: waiting for the plan
            const FlowSlot* sl = sm.slot(d);
            if (sl->sid < 0) break;
            mbar_wait_sleep<400>(sm.bar(FB_SDONE + d), par);
            if (ftid == 0) FP(4, k, 3);
            FT_MARK(21);   // finish: waiting for the stream warps
            const PoolWin pw = pool_window(single, k, a.spill_cap);
            finish_series(a, sm, sl, d, ftid, pw, spill, ns);
            if (ftid == 0) FP(4, k, 5);
            if (ftid == 0) {   // (finish_series ends with a barrier of the finish warps: all reads of the pool are done)
                __threadfence_block();
                atomicAdd(&sm.rel()[pw.win], 1u);
                mbar_arrive(sm.bar(FB_FREE + d));
            }
        }
    } else {
        // =============================== STREAM ===============================
        const int ws = warp;
        FT_BEGIN(blockIdx.x == 0 && ws == 0 && lane == 0);
        for (uint32_t k = 0;; ++k) {
            const int d = (int)(k % F_ND);
            const uint32_t par = (k / F_ND) & 1u;
            FT_MARK(31);   // stream warp 0: working (scan + epilogue)
            if (lane == 0 && ws < 3) FP(5 + ws, k, 1);
            mbar_wait_sleep<100>(sm.bar(FB_PLAN + d), par);
            if (lane == 0 && ws < 3) FP(5 + ws, k, 2);
            FT_MARK(30);   // stream warp 0: waiting for the plan
            const FlowSlot* sl = sm.slot(d);
            if (sl->sid < 0) break;
            if (ws == 0 && lane == 0) mbar_arrive(sm.bar(FB_START + d));
            StreamAcc acc;
            acc.s0 = 0.0;
            acc.s1 = 0.0;
            acc.mn = sl->mn_seed;
            acc.mx = sl->mx_seed;
            acc.nan = 0;
#pragma unroll
            for (int i = 0; i < NI_MAX; ++i) acc.c[i] = 0;
            const int ni = sl->ni;
            // the shared-memory window was last used by series k - 1 (one pool) or k - 2 (two halves); until the
            // finish warps have released it this series' candidates go to the spill window
            uint32_t flags = (pool_window(single, k, a.spill_cap).use == 0u ? 1u : 0u) | (sl->depth1 == 0 ? 8u : 0u);
            switch (ni) {
                case 0: stream_series<0>(a, sm, sl, acc, ws, ns, lane, single, k, flags); break;
                case 1: stream_series<1>(a, sm, sl, acc, ws, ns, lane, single, k, flags); break;
                case 2: stream_series<2>(a, sm, sl, acc, ws, ns, lane, single, k, flags); break;
                case 3: stream_series<3>(a, sm, sl, acc, ws, ns, lane, single, k, flags); break;
                default: stream_series<4>(a, sm, sl, acc, ws, ns, lane, single, k, flags); break;
            }
            if (__any_sync(0xffffffffu, (flags & 4u) != 0u)) __threadfence();   // spilled keys: in L2 before the finish warps read them
            const bool over = __any_sync(0xffffffffu, (flags & 2u) != 0u);
            if (lane == 0 && over) sm.ovf()[d] = 1u;
            // warp totals -> partial of (slot, warp)
            double s = acc.s0 + acc.s1, mn = acc.mn, mx = acc.mx;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                s += shfl_xor_f64(s, o);
                mn = fmin(mn, shfl_xor_f64(mn, o));
                mx = fmax(mx, shfl_xor_f64(mx, o));
            }
            const unsigned nn = __reduce_add_sync(0xffffffffu, acc.nan);
            unsigned c[NI_MAX];
#pragma unroll
            for (int i = 0; i < NI_MAX; ++i) c[i] = __reduce_add_sync(0xffffffffu, acc.c[i]);
            __syncwarp();
            if (lane == 0) {
                FlowPartial* p = sm.part(d, ws);
                p->sum = s;
                p->mn = mn;
                p->mx = mx;
                p->nan = nn;
#pragma unroll
                for (int i = 0; i < NI_MAX; ++i) p->c[i] = c[i];
                mbar_arrive(sm.bar(FB_SDONE + d));
            }
        }
    }
}

}  // namespace

// ------------------------------------------------------------------------------------------------ host
size_t ko_flow_smem_bytes() { return FLOW_SMEM_BYTES; }

cudaError_t ko_flow_set_attributes() {
    return cudaFuncSetAttribute(series_flow, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)FLOW_SMEM_BYTES);
}

#ifdef KO_FLOW_PROGRESS
extern "C" int ko_debug_flow_progress(unsigned* out2048) {
    cudaStream_t st;
    KO_CUDA(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
    void* p = nullptr;
    KO_CUDA(cudaGetSymbolAddress(&p, g_flow_prog));
    KO_CUDA(cudaMemcpyAsync(out2048, p, 2032 * sizeof(unsigned), cudaMemcpyDeviceToHost, st));
    KO_CUDA(cudaGetSymbolAddress(&p, g_flow_trap));
    KO_CUDA(cudaMemcpyAsync(out2048 + 2032, p, 16 * sizeof(unsigned), cudaMemcpyDeviceToHost, st));
    KO_CUDA(cudaStreamSynchronize(st));
    cudaStreamDestroy(st);
    return KO_OK;
}
#endif

#ifdef KO_FLOW_TIMING
extern "C" int ko_debug_flow_phases(unsigned long long* out64, int reset) {
    KO_CUDA(cudaMemcpyFromSymbol(out64, g_flow_phase, sizeof(unsigned long long) * 64));
    if (reset) {
        unsigned long long z[64] = {0};
        KO_CUDA(cudaMemcpyToSymbol(g_flow_phase, z, sizeof(z)));
    }
    return KO_OK;
}
#endif

// Enqueue one series_flow launch over the work list selected in `a` (grid <= one CTA per SM).
cudaError_t ko_flow_launch(const void* args, int grid, cudaStream_t st) {
    const SeriesArgs& a = *reinterpret_cast<const SeriesArgs*>(args);
    series_flow<<<grid, F_THREADS, FLOW_SMEM_BYTES, st>>>(a);
    return cudaGetLastError();
}
