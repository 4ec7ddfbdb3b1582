// Pod rollup (cluster_monitor.py:93-117) and the stable descending restart sort that reproduces
// quick_sort_pods / quick_sort_error_loki_container (cluster_monitor.py:450-479) as an LSD radix sort.
//
// The reference's quick sort is a STABLE DESCENDING sort of the member list.  Here every pod gets a 33-bit
// key  (non-member << 32) | ~orderable(restart_count)  and all pods are sorted ascending and stably by it:
// members come first, in descending restart order, ties in input order — exactly the reference's order —
// and no compaction pass is needed.  One 8-bit digit per pass: tile histogram -> exclusive scan ->
// stable scatter (ranks from warp match_any, warps own consecutive runs of the tile).
#include "ko_common.cuh"

namespace {

constexpr int SORT_THREADS = 256;
constexpr int SORT_WARPS = SORT_THREADS / 32;
constexpr int SORT_ITEMS = 8;                             // elements per thread
constexpr int SORT_TILE = SORT_THREADS * SORT_ITEMS;      // 2048 elements per CTA
constexpr int WARP_RUN = SORT_TILE / SORT_WARPS;          // 256 consecutive elements per warp

__device__ __forceinline__ unsigned orderable_desc(int32_t r) {
    // ascending order of the result == descending order of r (signed)
    return ~((unsigned)r ^ 0x80000000u);
}

// restart_count = sum of container restart counts; warn flag per container; sort keys of both lists
__global__ void pod_kernel(const int32_t* __restrict__ c_restart, const uint8_t* __restrict__ c_ready,
                           const int64_t* __restrict__ pod_off, const uint8_t* __restrict__ phase_ok,
                           int64_t n_pods, int32_t* o_restart, uint8_t* o_warn, unsigned long long* key_r,
                           unsigned long long* key_e, unsigned* idx, unsigned long long* n_members) {
    int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    bool mr = false, me = false;
    if (p < n_pods) {
        int32_t rc = 0;
        for (int64_t c = pod_off[p]; c < pod_off[p + 1]; ++c) {
            rc += c_restart[c];
            if (o_warn) o_warn[c] = c_ready[c] ? 0 : 1;
        }
        o_restart[p] = rc;
        mr = rc > 0;
        me = phase_ok[p] == 0;
        const unsigned long long k = orderable_desc(rc);
        key_r[p] = k | (mr ? 0ull : (1ull << 32));
        key_e[p] = k | (me ? 0ull : (1ull << 32));
        idx[p] = (unsigned)p;
    }
    const unsigned br = __ballot_sync(0xffffffffu, mr), be = __ballot_sync(0xffffffffu, me);
    if ((threadIdx.x & 31) == 0) {
        if (br) atomicAdd(&n_members[0], (unsigned long long)__popc(br));
        if (be) atomicAdd(&n_members[1], (unsigned long long)__popc(be));
    }
}

__global__ void keys_from_int32(const int32_t* __restrict__ keys, int64_t n, unsigned long long* out,
                                unsigned* idx) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) {
        out[i] = orderable_desc(keys[i]);
        idx[i] = (unsigned)i;
    }
}

// per-tile digit histogram, stored digit-major: hist[d * ntiles + tile]
__global__ void __launch_bounds__(SORT_THREADS) sort_hist(const unsigned long long* __restrict__ keys, int64_t n,
                                                          int shift, unsigned* hist, int ntiles) {
    __shared__ unsigned h[256];
    h[threadIdx.x] = 0;
    __syncthreads();
    const int64_t base = (int64_t)blockIdx.x * SORT_TILE;
    for (int k = 0; k < SORT_ITEMS; ++k) {
        const int64_t i = base + k * SORT_THREADS + threadIdx.x;
        if (i < n) atomicAdd(&h[(unsigned)(keys[i] >> shift) & 255u], 1u);
    }
    __syncthreads();
    hist[(size_t)threadIdx.x * ntiles + blockIdx.x] = h[threadIdx.x];
}

// exclusive scan of m = 256*ntiles counters by one CTA (chunks of 1024 with a running carry)
__global__ void __launch_bounds__(1024) sort_scan(unsigned* hist, int64_t m) {
    __shared__ unsigned wsum[32];
    __shared__ unsigned carry_s;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) carry_s = 0;
    __syncthreads();
    for (int64_t base = 0; base < m; base += 1024) {
        const int64_t i = base + threadIdx.x;
        const unsigned x = i < m ? hist[i] : 0u;
        unsigned incl = x;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            unsigned t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        if (lane == 31) wsum[warp] = incl;
        __syncthreads();
        if (warp == 0) {
            unsigned w = wsum[lane], wi = w;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                unsigned t = __shfl_up_sync(0xffffffffu, wi, o);
                if (lane >= o) wi += t;
            }
            wsum[lane] = wi - w;  // exclusive warp offsets
        }
        __syncthreads();
        const unsigned carry = carry_s;
        if (i < m) hist[i] = carry + wsum[warp] + incl - x;
        __syncthreads();
        if (threadIdx.x == 1023) carry_s = carry + wsum[warp] + incl;
        __syncthreads();
    }
}

// stable scatter of one tile
__global__ void __launch_bounds__(SORT_THREADS) sort_scatter(const unsigned long long* __restrict__ keys_in,
                                                             const unsigned* __restrict__ idx_in, int64_t n,
                                                             int shift, const unsigned* __restrict__ offs,
                                                             int ntiles, unsigned long long* keys_out,
                                                             unsigned* idx_out) {
    __shared__ unsigned cnt[SORT_WARPS][256];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < SORT_WARPS * 256; i += SORT_THREADS) (&cnt[0][0])[i] = 0;
    __syncthreads();
    const int64_t wbase = (int64_t)blockIdx.x * SORT_TILE + (int64_t)warp * WARP_RUN;
    unsigned long long k[WARP_RUN / 32];
    unsigned v[WARP_RUN / 32];
    // pass A: per-warp digit counts (each warp owns a consecutive run, visited in order)
#pragma unroll
    for (int s = 0; s < WARP_RUN / 32; ++s) {
        const int64_t i = wbase + s * 32 + lane;
        const bool ok = i < n;
        k[s] = ok ? keys_in[i] : 0ull;
        v[s] = ok ? idx_in[i] : 0u;
        const unsigned d = ok ? ((unsigned)(k[s] >> shift) & 255u) : 256u;
        const unsigned peers = __match_any_sync(0xffffffffu, d);
        if (ok && lane == (__ffs(peers) - 1)) cnt[warp][d] += __popc(peers);
        __syncwarp();
    }
    __syncthreads();
    // base[w][d] = global offset of (d, tile) + counts of earlier warps
    {
        const int d = threadIdx.x;
        unsigned run = offs[(size_t)d * ntiles + blockIdx.x];
        for (int w = 0; w < SORT_WARPS; ++w) {
            const unsigned c = cnt[w][d];
            cnt[w][d] = run;
            run += c;
        }
    }
    __syncthreads();
    // pass B: ranks inside the warp's run, in order
#pragma unroll
    for (int s = 0; s < WARP_RUN / 32; ++s) {
        const int64_t i = wbase + s * 32 + lane;
        const bool ok = i < n;
        const unsigned d = ok ? ((unsigned)(k[s] >> shift) & 255u) : 256u;
        const unsigned peers = __match_any_sync(0xffffffffu, d);
        if (ok) {
            const unsigned rank = __popc(peers & ((1u << lane) - 1u));
            const unsigned pos = cnt[warp][d] + rank;
            keys_out[pos] = k[s];
            idx_out[pos] = v[s];
        }
        __syncwarp();
        if (ok && lane == (__ffs(peers) - 1)) cnt[warp][d] += __popc(peers);
        __syncwarp();
    }
}

__global__ void widen_idx(const unsigned* __restrict__ idx, int64_t n, long long* out) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = (long long)idx[i];
}

struct SortBufs {
    unsigned long long* key[2];
    unsigned* idx[2];
    unsigned* hist;
};

size_t sort_scratch_bytes(int64_t n) {
    const size_t ntiles = (size_t)((n + SORT_TILE - 1) / SORT_TILE);
    return 2 * ko_align_up((size_t)n * 8, 256) + 2 * ko_align_up((size_t)n * 4, 256) +
           ko_align_up(ntiles * 256 * 4, 256) + 256;
}

void carve_sort(unsigned char* p, int64_t n, SortBufs* b) {
    const size_t ntiles = (size_t)((n + SORT_TILE - 1) / SORT_TILE);
    b->key[0] = (unsigned long long*)p; p += ko_align_up((size_t)n * 8, 256);
    b->key[1] = (unsigned long long*)p; p += ko_align_up((size_t)n * 8, 256);
    b->idx[0] = (unsigned*)p;           p += ko_align_up((size_t)n * 4, 256);
    b->idx[1] = (unsigned*)p;           p += ko_align_up((size_t)n * 4, 256);
    b->hist = (unsigned*)p;
    (void)ntiles;
}

// Sorts key[0]/idx[0] by the low `bits` bits; returns which buffer holds the result.
int radix_sort(ko_ctx* ctx, SortBufs* b, int64_t n, int bits, cudaStream_t st, int* result_buf) {
    const int ntiles = (int)((n + SORT_TILE - 1) / SORT_TILE);
    int cur = 0;
    for (int shift = 0; shift < bits; shift += 8) {
        sort_hist<<<ntiles, SORT_THREADS, 0, st>>>(b->key[cur], n, shift, b->hist, ntiles);
        sort_scan<<<1, 1024, 0, st>>>(b->hist, (int64_t)ntiles * 256);
        sort_scatter<<<ntiles, SORT_THREADS, 0, st>>>(b->key[cur], b->idx[cur], n, shift, b->hist, ntiles,
                                                      b->key[cur ^ 1], b->idx[cur ^ 1]);
        ctx->counters.kernel_launches += 3;
        cur ^= 1;
    }
    KO_CUDA(cudaGetLastError());
    *result_buf = cur;
    return KO_OK;
}

}  // namespace

extern "C" {

int ko_stable_desc_order(ko_ctx* ctx, const int32_t* keys, int64_t n, int64_t* out_order) {
    KO_REQUIRE(ctx && ctx->alive, "ko_stable_desc_order: ctx not initialised");
    KO_REQUIRE(n >= 0 && n < 4294967295LL, "ko_stable_desc_order: bad size");
    if (n == 0) return KO_OK;
    KO_REQUIRE(keys && out_order, "ko_stable_desc_order: NULL argument");
    std::lock_guard<std::mutex> lk(ctx->mu);
    KO_CUDA(cudaSetDevice(ctx->device));
    int rc;
    if ((rc = ko_buf_reserve(&ctx->scratch, sort_scratch_bytes(n)))) return rc;
    if ((rc = ko_buf_reserve(&ctx->in_small, (size_t)n * 4 + 256))) return rc;
    if ((rc = ko_buf_reserve(&ctx->out_small, (size_t)n * 8 + 256))) return rc;
    SortBufs b;
    carve_sort((unsigned char*)ctx->scratch.p, n, &b);
    cudaStream_t st = ctx->stream;
    KO_CUDA(cudaMemcpyAsync(ctx->in_small.p, keys, (size_t)n * 4, cudaMemcpyHostToDevice, st));
    keys_from_int32<<<(unsigned)((n + 255) / 256), 256, 0, st>>>((const int32_t*)ctx->in_small.p, n, b.key[0], b.idx[0]);
    ctx->counters.kernel_launches += 1;
    int res;
    if ((rc = radix_sort(ctx, &b, n, 32, st, &res))) return rc;
    widen_idx<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(b.idx[res], n, (long long*)ctx->out_small.p);
    ctx->counters.kernel_launches += 1;
    KO_CUDA(cudaGetLastError());
    KO_CUDA(cudaMemcpyAsync(out_order, ctx->out_small.p, (size_t)n * 8, cudaMemcpyDeviceToHost, st));
    KO_CUDA(cudaStreamSynchronize(st));
    ctx->counters.h2d_bytes += n * 4;
    ctx->counters.d2h_bytes += n * 8;
    return KO_OK;
}

int ko_pod_rollup(ko_ctx* ctx, const int32_t* container_restart, const uint8_t* container_ready,
                  const int64_t* pod_offsets, const uint8_t* pod_phase_ok, int64_t n_pods, int32_t* out_restart,
                  uint8_t* out_warn, int64_t* out_order_restart_desc, int64_t* out_n_restart,
                  int64_t* out_order_error_desc, int64_t* out_n_error) {
    KO_REQUIRE(ctx && ctx->alive, "ko_pod_rollup: ctx not initialised");
    KO_REQUIRE(n_pods >= 0 && n_pods < 4294967295LL, "ko_pod_rollup: bad size");
    KO_REQUIRE(out_n_restart && out_n_error, "ko_pod_rollup: NULL count output");
    *out_n_restart = 0;
    *out_n_error = 0;
    if (n_pods == 0) return KO_OK;
    KO_REQUIRE(pod_offsets && pod_phase_ok && out_restart && out_order_restart_desc && out_order_error_desc,
               "ko_pod_rollup: NULL argument");
    KO_REQUIRE(pod_offsets[0] == 0, "ko_pod_rollup: pod_offsets must start at 0");
    for (int64_t p = 0; p < n_pods; ++p)
        KO_REQUIRE(pod_offsets[p + 1] >= pod_offsets[p], "ko_pod_rollup: pod_offsets must be non-decreasing");
    const size_t P = (size_t)n_pods, C = (size_t)pod_offsets[n_pods];
    KO_REQUIRE(C == 0 || (container_restart && container_ready), "ko_pod_rollup: NULL container arrays");
    std::lock_guard<std::mutex> lk(ctx->mu);
    KO_CUDA(cudaSetDevice(ctx->device));
    int rc;
    const size_t sortb = sort_scratch_bytes(n_pods);
    // scratch: sort buffers + second key column + member counters
    if ((rc = ko_buf_reserve(&ctx->scratch, sortb + ko_align_up(P * 8, 256) + ko_align_up(P * 4, 256) + 512))) return rc;
    if ((rc = ko_buf_reserve(&ctx->in_small, ko_align_up(C * 4, 256) + ko_align_up(C, 256) +
                                                 ko_align_up((P + 1) * 8, 256) + ko_align_up(P, 256) + 256)))
        return rc;
    if ((rc = ko_buf_reserve(&ctx->out_small, ko_align_up(P * 4, 256) + ko_align_up(C, 256) +
                                                  2 * ko_align_up(P * 8, 256) + 256)))
        return rc;
    unsigned char* sp = (unsigned char*)ctx->scratch.p;
    SortBufs b;
    carve_sort(sp, n_pods, &b);
    unsigned long long* key_e = (unsigned long long*)(sp + sortb);
    unsigned* idx0 = (unsigned*)(sp + sortb + ko_align_up(P * 8, 256));
    unsigned long long* d_nm = (unsigned long long*)(sp + sortb + ko_align_up(P * 8, 256) + ko_align_up(P * 4, 256));
    unsigned char* ip = (unsigned char*)ctx->in_small.p;
    int32_t* d_cr = (int32_t*)ip;            ip += ko_align_up(C * 4, 256);
    uint8_t* d_rd = (uint8_t*)ip;            ip += ko_align_up(C, 256);
    int64_t* d_po = (int64_t*)ip;            ip += ko_align_up((P + 1) * 8, 256);
    uint8_t* d_ok = (uint8_t*)ip;
    unsigned char* op = (unsigned char*)ctx->out_small.p;
    int32_t* o_restart = (int32_t*)op;       op += ko_align_up(P * 4, 256);
    uint8_t* o_warn = (uint8_t*)op;          op += ko_align_up(C, 256);
    long long* o_ord_r = (long long*)op;     op += ko_align_up(P * 8, 256);
    long long* o_ord_e = (long long*)op;
    cudaStream_t st = ctx->stream;
    if (C) {
        KO_CUDA(cudaMemcpyAsync(d_cr, container_restart, C * 4, cudaMemcpyHostToDevice, st));
        KO_CUDA(cudaMemcpyAsync(d_rd, container_ready, C, cudaMemcpyHostToDevice, st));
    }
    KO_CUDA(cudaMemcpyAsync(d_po, pod_offsets, (P + 1) * 8, cudaMemcpyHostToDevice, st));
    KO_CUDA(cudaMemcpyAsync(d_ok, pod_phase_ok, P, cudaMemcpyHostToDevice, st));
    KO_CUDA(cudaMemsetAsync(d_nm, 0, 16, st));
    const unsigned pb = (unsigned)((P + 255) / 256);
    pod_kernel<<<pb, 256, 0, st>>>(d_cr, d_rd, d_po, d_ok, n_pods, o_restart, out_warn ? o_warn : nullptr, b.key[0],
                                   key_e, idx0, d_nm);
    ctx->counters.kernel_launches += 1;
    // restart list
    KO_CUDA(cudaMemcpyAsync(b.idx[0], idx0, P * 4, cudaMemcpyDeviceToDevice, st));
    int res;
    if ((rc = radix_sort(ctx, &b, n_pods, 40, st, &res))) return rc;
    widen_idx<<<pb, 256, 0, st>>>(b.idx[res], n_pods, o_ord_r);
    // error list
    KO_CUDA(cudaMemcpyAsync(b.key[0], key_e, P * 8, cudaMemcpyDeviceToDevice, st));
    KO_CUDA(cudaMemcpyAsync(b.idx[0], idx0, P * 4, cudaMemcpyDeviceToDevice, st));
    if ((rc = radix_sort(ctx, &b, n_pods, 40, st, &res))) return rc;
    widen_idx<<<pb, 256, 0, st>>>(b.idx[res], n_pods, o_ord_e);
    ctx->counters.kernel_launches += 2;
    KO_CUDA(cudaGetLastError());
    unsigned long long nm[2] = {0, 0};
    KO_CUDA(cudaMemcpyAsync(nm, d_nm, 16, cudaMemcpyDeviceToHost, st));
    KO_CUDA(cudaMemcpyAsync(out_restart, o_restart, P * 4, cudaMemcpyDeviceToHost, st));
    if (out_warn && C) KO_CUDA(cudaMemcpyAsync(out_warn, o_warn, C, cudaMemcpyDeviceToHost, st));
    KO_CUDA(cudaMemcpyAsync(out_order_restart_desc, o_ord_r, P * 8, cudaMemcpyDeviceToHost, st));
    KO_CUDA(cudaMemcpyAsync(out_order_error_desc, o_ord_e, P * 8, cudaMemcpyDeviceToHost, st));
    KO_CUDA(cudaStreamSynchronize(st));
    *out_n_restart = (int64_t)nm[0];
    *out_n_error = (int64_t)nm[1];
    ctx->counters.h2d_bytes += (int64_t)(C * 5 + (P + 1) * 8 + P);
    ctx->counters.d2h_bytes += (int64_t)(P * 20 + (out_warn ? C : 0) + 16);
    return KO_OK;
}

}  // extern "C"
