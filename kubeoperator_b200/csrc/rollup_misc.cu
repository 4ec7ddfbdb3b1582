// Node / fleet / availability / Loki / counter kernels of libkorollup.so and their host wrappers.
// Reference lines replaced are cited per entry point in include/korollup.h.
#include "ko_common.cuh"

namespace {

// Bump allocator over a ctx-owned device buffer (256-byte aligned pieces).
struct Bump {
    unsigned char* base;
    size_t off;
    template <typename T>
    T* take(size_t n) {
        T* p = reinterpret_cast<T*>(base + off);
        off += ko_align_up(n * sizeof(T), 256);
        return p;
    }
};
inline size_t piece(size_t n, size_t elem) { return ko_align_up(n * elem, 256); }

__device__ __forceinline__ double wsum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ long long wsum_ll(long long v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// progress-color.pipe.ts:10-20 on a percent value
__device__ __forceinline__ uint8_t band_of(double pct) {
    if (!(pct == pct) || pct == 0.0) return KO_BAND_NONE;
    if (pct <= 60.0) return KO_BAND_GREEN;
    if (pct <= 80.0) return KO_BAND_YELLOW;
    return KO_BAND_RED;
}

// cluster_monitor.py:174-176 per node (+ colour bands, + validation of the grouping)
__global__ void node_flags(const double* __restrict__ cu, const double* __restrict__ mu,
                           const int32_t* __restrict__ cl, int64_t n, int32_t n_clusters, double thr,
                           uint8_t* flag, uint8_t* band_c, uint8_t* band_m, int* err) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double c = cu[i], m = mu[i];
    const bool zero = (c == 0.0) && (m == 0.0);
    if (flag) flag[i] = (!zero && (c > thr || m > thr)) ? 1 : 0;
    if (band_c) band_c[i] = band_of(c * 100.0);
    if (band_m) band_m[i] = band_of(m * 100.0);
    const int32_t k = cl[i];
    if (k < 0 || k >= n_clusters || (i > 0 && cl[i - 1] > k)) atomicExch(err, 1);
}

// cluster_monitor.py:163-173,180-182: one warp per cluster over its contiguous node range.  The four sums are taken in
// NODE ORDER, like the reference's loop (x = x + float(...), :170-173), so that the floats of the blob are the
// reference's to the last bit: the warp loads 32 nodes at a time (coalesced) into shared memory and lane 0 adds them
// one after the other - four independent chains, about ten cycles per node.
__global__ void cluster_reduce(const double* __restrict__ cpu, const double* __restrict__ mem,
                               const double* __restrict__ cu, const double* __restrict__ mu,
                               const int32_t* __restrict__ cl, int64_t n, int32_t n_clusters,
                               double* o_cpu, double* o_mem, double* o_cu, double* o_mu, int32_t* o_cnt) {
    const int lane = threadIdx.x & 31;
    const int64_t c = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (c >= n_clusters) return;
    // lower_bound(c) and lower_bound(c+1)
    int64_t lo = 0, hi = n;
    while (lo < hi) {
        int64_t mid = (lo + hi) >> 1;
        if (cl[mid] < (int32_t)c) lo = mid + 1; else hi = mid;
    }
    const int64_t b = lo;
    hi = n;
    while (lo < hi) {
        int64_t mid = (lo + hi) >> 1;
        if (cl[mid] <= (int32_t)c) lo = mid + 1; else hi = mid;
    }
    const int64_t e = lo;
    __shared__ double stage[8][4][32];   // per warp of the block: 32 nodes x 4 columns
    double (*st)[32] = stage[(threadIdx.x >> 5) & 7];
    double s_cpu = 0, s_mem = 0, s_cu = 0, s_mu = 0;
    long long zeros = 0;
    for (int64_t i0 = b; i0 < e; i0 += 32) {
        const int64_t i = i0 + lane;
        if (i < e) {
            const double a = cu[i], m = mu[i];
            st[0][lane] = cpu[i];
            st[1][lane] = mem[i];
            st[2][lane] = a;
            st[3][lane] = m;
            zeros += (a == 0.0 && m == 0.0) ? 1 : 0;
        }
        __syncwarp();
        if (lane == 0) {
            const int cnt = (int)((e - i0) < 32 ? (e - i0) : 32);
            for (int k = 0; k < cnt; ++k) {
                s_cpu += st[0][k];
                s_mem += st[1][k];
                s_cu += st[2][k];
                s_mu += st[3][k];
            }
        }
        __syncwarp();
    }
    zeros = wsum_ll(zeros);
    if (lane == 0) {
        const long long count = (long long)(e - b) - zeros;
        if (count > 0) {
            s_cu = s_cu / (double)count;
            s_mu = s_mu / (double)count;
        }
        o_cpu[c] = s_cpu;
        o_mem[c] = s_mem;
        o_cu[c] = s_cu;
        o_mu[c] = s_mu;
        o_cnt[c] = (int32_t)count;
    }
}

// dashboard.component.ts:117-153 with one warp
__global__ void fleet_rollup_kernel(const int64_t* __restrict__ counts, const double* __restrict__ cpu_total,
                                    const double* __restrict__ mem_total, const double* __restrict__ cpu_usage,
                                    const double* __restrict__ mem_usage, const int32_t* __restrict__ max_pods,
                                    int64_t n, double* out) {
    const int lane = threadIdx.x;
    long long pods = 0, ns = 0, dep = 0, cont = 0, nodes = 0, maxpod = 0, zero = 0;
    double ct = 0, mt = 0, cu = 0, mu = 0;
    // the four float sums in CLUSTER ORDER (the UI's forEach, dashboard.component.ts:121-133): lane 0 adds what the
    // warp staged in shared memory, 32 clusters at a time
    __shared__ double st[4][32];
    for (int64_t c0 = 0; c0 < n; c0 += 32) {
        const int64_t c = c0 + lane;
        if (c < n) {
            st[0][lane] = cpu_total[c];
            st[1][lane] = mem_total[c];
            st[2][lane] = cpu_usage[c];
            st[3][lane] = mem_usage[c];
        }
        __syncwarp();
        if (lane == 0) {
            const int cnt = (int)((n - c0) < 32 ? (n - c0) : 32);
            for (int k = 0; k < cnt; ++k) {
                ct += st[0][k];
                mt += st[1][k];
                cu += st[2][k];
                mu += st[3][k];
            }
        }
        __syncwarp();
    }
    for (int64_t c = lane; c < n; c += 32) {
        pods += counts[c * 6 + 0];
        ns += counts[c * 6 + 1];
        dep += counts[c * 6 + 2];
        cont += counts[c * 6 + 3];
        nodes += counts[c * 6 + 4];
        zero += (cpu_total[c] == 0.0 && mem_total[c] == 0.0) ? 1 : 0;
        // :136-143: only a cluster the UI's own list knows adds capacity (ui_nodes >= 0 = cluster.nodes.length);
        // MAX_PODS === undefined (passed as a negative value) means 110, a configured 0 stays 0
        const long long ui_nodes = counts[c * 6 + 5];
        const long long mp = max_pods[c] >= 0 ? max_pods[c] : 110;
        if (ui_nodes >= 0) maxpod += mp * ui_nodes;
    }
    pods = wsum_ll(pods); ns = wsum_ll(ns); dep = wsum_ll(dep); cont = wsum_ll(cont);
    nodes = wsum_ll(nodes); maxpod = wsum_ll(maxpod); zero = wsum_ll(zero);
    if (lane == 0) {
        const long long count = (long long)n - zero;
        if (count > 0) {
            cu = cu / (double)count * 100;
            mu = mu / (double)count * 100;
        }
        double cusage = 0;
        if (maxpod > 0) cusage = (double)pods / (double)maxpod * 100;
        out[0] = (double)pods; out[1] = (double)ns; out[2] = (double)dep; out[3] = (double)cont;
        out[4] = (double)nodes; out[5] = ct; out[6] = mt; out[7] = cu; out[8] = mu;
        out[9] = (double)maxpod; out[10] = cusage; out[11] = (double)count;
    }
}

// prometheus_client.py:75-85, one thread per cluster
__global__ void available_rate_kernel(const int32_t* __restrict__ up, const int32_t* __restrict__ len,
                                      const int64_t* __restrict__ joff, int64_t n_clusters, double* job_rate,
                                      double* cluster_rate) {
    int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= n_clusters) return;
    int service_up = 0;
    const int64_t b = joff[c], e = joff[c + 1];
    for (int64_t j = b; j < e; ++j) {
        double r = len[j] > 0 ? (double)up[j] / (double)len[j] * 100 : 0;
        if (r == 100) service_up++;
        job_rate[j] = r;
    }
    cluster_rate[c] = (e > b) ? (double)service_up / (double)(e - b) * 100 : 0;
}

// one refinement round of the fleet quantile
__global__ void key_hist_kernel(const double* __restrict__ v, int64_t n, int shift, unsigned long long pmask,
                                unsigned long long prefix, unsigned long long* hist) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const double x = v[i];
        if (x == x) {
            const unsigned long long k = ko_okey(x + 0.0);
            if ((k & pmask) == prefix) atomicAdd(&hist[(k >> shift) & 0xFFFFull], 1ull);
        }
    }
}

// The same with the refinement state (prefix, mask, wanted rank, result) in device memory, so that the rounds of a
// fleet quantile follow each other on the stream without a trip to the host: histogram -> all-reduce -> select.
struct KoSelState {
    unsigned long long prefix, mask;
    long long rank;      // 1-based rank still wanted inside the prefix; < 1: no valid value in the fleet
    double result;       // set by the round with shift == 0 (NaN when there is no value)
};
__global__ void key_hist_state_kernel(const double* __restrict__ v, int64_t n, int shift, const KoSelState* st,
                                      unsigned long long* hist) {
    const unsigned long long pmask = st->mask, prefix = st->prefix;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const double x = v[i];
        if (x == x) {
            const unsigned long long k = ko_okey(x + 0.0);
            if ((k & pmask) == prefix) atomicAdd(&hist[(k >> shift) & 0xFFFFull], 1ull);
        }
    }
}
// One CTA of 1024 threads: the digit whose cumulative count first reaches the wanted rank (64 bins per thread).
__global__ void __launch_bounds__(1024) key_select_kernel(const unsigned long long* __restrict__ hist, int shift,
                                                          KoSelState* st) {
    __shared__ unsigned long long wsum[32];
    __shared__ int found;
    const int t = threadIdx.x, lane = t & 31, w = t >> 5;
    const long long rank = st->rank;
    unsigned long long s = 0;
    for (int i = 0; i < 64; ++i) s += hist[t * 64 + i];
    unsigned long long incl = s;
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned long long u = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += u;
    }
    if (lane == 31) wsum[w] = incl;
    if (t == 0) found = 0;
    __syncthreads();
    unsigned long long before_w = 0;
    for (int i = 0; i < w; ++i) before_w += wsum[i];
    incl += before_w;
    const unsigned long long excl = incl - s;
    if (rank >= 1 && excl < (unsigned long long)rank && (unsigned long long)rank <= incl) {
        unsigned long long run = excl;
        int d = 63;
        for (int i = 0; i < 64; ++i) {
            const unsigned long long h = hist[t * 64 + i];
            if (run + h >= (unsigned long long)rank) {
                d = i;
                break;
            }
            run += h;
        }
        const unsigned long long digit = (unsigned long long)(t * 64 + d);
        st->prefix |= digit << shift;
        st->mask |= 0xFFFFull << shift;
        st->rank = rank - (long long)run;
        if (shift == 0) st->result = ko_key2d(st->prefix) + 0.0;
        found = 1;
    }
    __syncthreads();
    if (t == 0 && !found) {   // no valid value anywhere (or an inconsistent rank): NaN, and stay that way
        st->rank = 0;
        st->result = __longlong_as_double(0x7FF8000000000000LL);
    }
}

// prometheus_client.py:142-145: one warp per line looks for "level=error"
__global__ void loki_count_kernel(const uint8_t* __restrict__ buf, const int64_t* __restrict__ loff,
                                  int64_t n_lines, const int64_t* __restrict__ goff, int64_t n_groups,
                                  unsigned long long* out) {
    const int lane = threadIdx.x & 31;
    const int64_t l = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (l >= n_lines) return;
    const int64_t a = loff[l], b = loff[l + 1];
    const char needle[11] = {'l', 'e', 'v', 'e', 'l', '=', 'e', 'r', 'r', 'o', 'r'};
    bool hit = false;
    for (int64_t base = a; base + 11 <= b && !hit; base += 32) {
        const int64_t p = base + lane;
        bool h = false;
        if (p + 11 <= b && buf[p] == 'l') {
            h = true;
#pragma unroll
            for (int k = 1; k < 11; ++k) h = h && (buf[p + k] == (uint8_t)needle[k]);
        }
        hit = __any_sync(0xffffffffu, h);
    }
    if (lane == 0 && hit) {
        int64_t lo = 0, hi = n_groups;  // group g with goff[g] <= l < goff[g+1]
        while (lo < hi) {
            int64_t mid = (lo + hi) >> 1;
            if (goff[mid + 1] <= l) lo = mid + 1; else hi = mid;
        }
        if (lo < n_groups) atomicAdd(&out[lo], 1ull);
    }
}

// increase()/rate() with counter resets; one CTA per series
__global__ void __launch_bounds__(256) counter_increase_kernel(const double* __restrict__ values,
                                                               const int64_t* __restrict__ offsets,
                                                               double step, double* o_inc, double* o_rate) {
    __shared__ double s_sum[8];
    __shared__ long long s_first[8], s_last[8];
    const int64_t s = blockIdx.x;
    const double* v = values + offsets[s];
    const int64_t L = offsets[s + 1] - offsets[s];
    double acc = 0.0;
    long long first = 0x7FFFFFFFFFFFFFFFLL, last = -1;
    for (int64_t i = threadIdx.x; i < L; i += blockDim.x) {
        const double x = v[i];
        if (x == x) {
            if (i < first) first = i;
            if (i > last) last = i;
            int64_t j = i - 1;
            while (j >= 0 && !(v[j] == v[j])) --j;
            if (j >= 0) {
                const double p = v[j];
                acc += (x >= p) ? x - p : x;
            }
        }
    }
    acc = wsum(acc);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        first = min(first, __shfl_xor_sync(0xffffffffu, first, o));
        last = max(last, __shfl_xor_sync(0xffffffffu, last, o));
    }
    const int w = threadIdx.x >> 5;
    if ((threadIdx.x & 31) == 0) {
        s_sum[w] = acc;
        s_first[w] = first;
        s_last[w] = last;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0;
        long long f = 0x7FFFFFFFFFFFFFFFLL, l = -1;
        for (int k = 0; k < 8; ++k) {
            t += s_sum[k];
            f = min(f, s_first[k]);
            l = max(l, s_last[k]);
        }
        const bool ok = (l > f);
        if (!ok) t = 0.0;
        if (o_inc) o_inc[s] = t;
        if (o_rate) o_rate[s] = ok ? t / ((double)(l - f) * step) : 0.0;
    }
}

}  // namespace

extern "C" {

int ko_cluster_rollup(ko_ctx* ctx, const double* cpu, const double* mem, const double* cpu_usage,
                      const double* mem_usage, const int32_t* cluster_of_node, int64_t n_nodes,
                      int32_t n_clusters, double threshold, double* out_cpu_total, double* out_mem_total,
                      double* out_cpu_usage, double* out_mem_usage, int32_t* out_count, uint8_t* out_alert_flag,
                      uint8_t* out_band_cpu, uint8_t* out_band_mem) {
    KO_REQUIRE(ctx && ctx->alive, "ko_cluster_rollup: ctx not initialised");
    KO_REQUIRE(n_nodes >= 0 && n_clusters >= 0, "ko_cluster_rollup: negative size");
    KO_REQUIRE(n_nodes == 0 || (cpu && mem && cpu_usage && mem_usage && cluster_of_node),
               "ko_cluster_rollup: NULL input");
    KO_REQUIRE(n_clusters == 0 || (out_cpu_total && out_mem_total && out_cpu_usage && out_mem_usage && out_count),
               "ko_cluster_rollup: NULL output");
    if (n_clusters == 0) {
        KO_REQUIRE(n_nodes == 0, "ko_cluster_rollup: nodes without clusters");
        return KO_OK;
    }
    std::lock_guard<std::mutex> lk(ctx->mu);
    KO_CUDA(cudaSetDevice(ctx->device));
    const size_t N = (size_t)n_nodes, K = (size_t)n_clusters;
    int rc;
    if ((rc = ko_buf_reserve(&ctx->in_small, 4 * piece(N, 8) + piece(N, 4) + 256))) return rc;
    if ((rc = ko_buf_reserve(&ctx->out_small, 4 * piece(K, 8) + piece(K, 4) + 3 * piece(N, 1) + 512))) return rc;
    Bump in{(unsigned char*)ctx->in_small.p, 0}, out{(unsigned char*)ctx->out_small.p, 0};
    double* d_cpu = in.take<double>(N);
    double* d_mem = in.take<double>(N);
    double* d_cu = in.take<double>(N);
    double* d_mu = in.take<double>(N);
    int32_t* d_cl = in.take<int32_t>(N);
    double* o_cpu = out.take<double>(K);
    double* o_mem = out.take<double>(K);
    double* o_cu = out.take<double>(K);
    double* o_mu = out.take<double>(K);
    int32_t* o_cnt = out.take<int32_t>(K);
    uint8_t* o_flag = out.take<uint8_t>(N);
    uint8_t* o_bc = out.take<uint8_t>(N);
    uint8_t* o_bm = out.take<uint8_t>(N);
    int* d_err = out.take<int>(1);
    cudaStream_t st = ctx->stream;
    if (N) {
        KO_CUDA(cudaMemcpyAsync(d_cpu, cpu, N * 8, cudaMemcpyHostToDevice, st));
        KO_CUDA(cudaMemcpyAsync(d_mem, mem, N * 8, cudaMemcpyHostToDevice, st));
        KO_CUDA(cudaMemcpyAsync(d_cu, cpu_usage, N * 8, cudaMemcpyHostToDevice, st));
        KO_CUDA(cudaMemcpyAsync(d_mu, mem_usage, N * 8, cudaMemcpyHostToDevice, st));
        KO_CUDA(cudaMemcpyAsync(d_cl, cluster_of_node, N * 4, cudaMemcpyHostToDevice, st));
    }
    ctx->counters.h2d_bytes += (int64_t)(N * 36);
    KO_CUDA(cudaMemsetAsync(d_err, 0, sizeof(int), st));
    if (N) {
        node_flags<<<(unsigned)((N + 255) / 256), 256, 0, st>>>(d_cu, d_mu, d_cl, n_nodes, n_clusters, threshold,
                                                                o_flag, out_band_cpu ? o_bc : nullptr,
                                                                out_band_mem ? o_bm : nullptr, d_err);
        ctx->counters.kernel_launches += 1;
    }
    cluster_reduce<<<(unsigned)((K * 32 + 255) / 256), 256, 0, st>>>(d_cpu, d_mem, d_cu, d_mu, d_cl, n_nodes,
                                                                     n_clusters, o_cpu, o_mem, o_cu, o_mu, o_cnt);
    ctx->counters.kernel_launches += 1;
    KO_CUDA(cudaGetLastError());
    int err = 0;
    KO_CUDA(cudaMemcpyAsync(&err, d_err, sizeof(int), cudaMemcpyDeviceToHost, st));
    KO_CUDA(cudaMemcpyAsync(out_cpu_total, o_cpu, K * 8, cudaMemcpyDeviceToHost, st));
    KO_CUDA(cudaMemcpyAsync(out_mem_total, o_mem, K * 8, cudaMemcpyDeviceToHost, st));
    KO_CUDA(cudaMemcpyAsync(out_cpu_usage, o_cu, K * 8, cudaMemcpyDeviceToHost, st));
    KO_CUDA(cudaMemcpyAsync(out_mem_usage, o_mu, K * 8, cudaMemcpyDeviceToHost, st));
    KO_CUDA(cudaMemcpyAsync(out_count, o_cnt, K * 4, cudaMemcpyDeviceToHost, st));
    if (N && out_alert_flag) KO_CUDA(cudaMemcpyAsync(out_alert_flag, o_flag, N, cudaMemcpyDeviceToHost, st));
    if (N && out_band_cpu) KO_CUDA(cudaMemcpyAsync(out_band_cpu, o_bc, N, cudaMemcpyDeviceToHost, st));
    if (N && out_band_mem) KO_CUDA(cudaMemcpyAsync(out_band_mem, o_bm, N, cudaMemcpyDeviceToHost, st));
    KO_CUDA(cudaStreamSynchronize(st));
    ctx->counters.d2h_bytes += (int64_t)(K * 36 + N * ((out_alert_flag != 0) + (out_band_cpu != 0) + (out_band_mem != 0)));
    if (err) {
        ko_set_error("ko_cluster_rollup: cluster_of_node must be non-decreasing and within [0, n_clusters)");
        return KO_EINVAL;
    }
    return KO_OK;
}

int ko_fleet_rollup(ko_ctx* ctx, const int64_t* counts, const double* cpu_total, const double* mem_total,
                    const double* cpu_usage, const double* mem_usage, const int32_t* max_pods, int64_t n_clusters,
                    double* out) {
    KO_REQUIRE(ctx && ctx->alive, "ko_fleet_rollup: ctx not initialised");
    KO_REQUIRE(n_clusters >= 0 && out, "ko_fleet_rollup: bad argument");
    KO_REQUIRE(n_clusters == 0 || (counts && cpu_total && mem_total && cpu_usage && mem_usage && max_pods),
               "ko_fleet_rollup: NULL input");
    std::lock_guard<std::mutex> lk(ctx->mu);
    KO_CUDA(cudaSetDevice(ctx->device));
    const size_t K = (size_t)n_clusters;
    int rc;
    if ((rc = ko_buf_reserve(&ctx->in_small, piece(K * 6, 8) + 4 * piece(K, 8) + piece(K, 4) + 256))) return rc;
    if ((rc = ko_buf_reserve(&ctx->out_small, 256))) return rc;
    Bump in{(unsigned char*)ctx->in_small.p, 0};
    int64_t* d_counts = in.take<int64_t>(K * 6);
    double* d_ct = in.take<double>(K);
    double* d_mt = in.take<double>(K);
    double* d_cu = in.take<double>(K);
    double* d_mu = in.take<double>(K);
    int32_t* d_mp = in.take<int32_t>(K);
    double* d_out = (double*)ctx->out_small.p;
    cudaStream_t st = ctx->stream;
    if (K) {
        KO_CUDA(cudaMemcpyAsync(d_counts, counts, K * 48, cudaMemcpyHostToDevice, st));
        KO_CUDA(cudaMemcpyAsync(d_ct, cpu_total, K * 8, cudaMemcpyHostToDevice, st));
        KO_CUDA(cudaMemcpyAsync(d_mt, mem_total, K * 8, cudaMemcpyHostToDevice, st));
        KO_CUDA(cudaMemcpyAsync(d_cu, cpu_usage, K * 8, cudaMemcpyHostToDevice, st));
        KO_CUDA(cudaMemcpyAsync(d_mu, mem_usage, K * 8, cudaMemcpyHostToDevice, st));
        KO_CUDA(cudaMemcpyAsync(d_mp, max_pods, K * 4, cudaMemcpyHostToDevice, st));
    }
    fleet_rollup_kernel<<<1, 32, 0, st>>>(d_counts, d_ct, d_mt, d_cu, d_mu, d_mp, n_clusters, d_out);
    KO_CUDA(cudaGetLastError());
    ctx->counters.kernel_launches += 1;
    KO_CUDA(cudaMemcpyAsync(out, d_out, 12 * 8, cudaMemcpyDeviceToHost, st));
    KO_CUDA(cudaStreamSynchronize(st));
    ctx->counters.h2d_bytes += (int64_t)(K * 84);
    ctx->counters.d2h_bytes += 96;
    return KO_OK;
}

int ko_available_rate(ko_ctx* ctx, const int32_t* job_up, const int32_t* job_len, const int64_t* job_offsets,
                      int64_t n_clusters, double* out_job_rate, double* out_cluster_rate) {
    KO_REQUIRE(ctx && ctx->alive, "ko_available_rate: ctx not initialised");
    KO_REQUIRE(n_clusters >= 0, "ko_available_rate: negative size");
    if (n_clusters == 0) return KO_OK;
    KO_REQUIRE(job_offsets && out_cluster_rate, "ko_available_rate: NULL argument");
    const int64_t nj = job_offsets[n_clusters] - job_offsets[0];
    KO_REQUIRE(job_offsets[0] == 0 && nj >= 0, "ko_available_rate: job_offsets must start at 0");
    for (int64_t c = 0; c < n_clusters; ++c)
        KO_REQUIRE(job_offsets[c + 1] >= job_offsets[c], "ko_available_rate: job_offsets must be non-decreasing");
    KO_REQUIRE(nj == 0 || (job_up && job_len && out_job_rate), "ko_available_rate: NULL job arrays");
    std::lock_guard<std::mutex> lk(ctx->mu);
    KO_CUDA(cudaSetDevice(ctx->device));
    const size_t K = (size_t)n_clusters, J = (size_t)nj;
    int rc;
    if ((rc = ko_buf_reserve(&ctx->in_small, 2 * piece(J, 4) + piece(K + 1, 8) + 256))) return rc;
    if ((rc = ko_buf_reserve(&ctx->out_small, piece(J, 8) + piece(K, 8) + 256))) return rc;
    Bump in{(unsigned char*)ctx->in_small.p, 0}, out{(unsigned char*)ctx->out_small.p, 0};
    int32_t* d_up = in.take<int32_t>(J);
    int32_t* d_len = in.take<int32_t>(J);
    int64_t* d_off = in.take<int64_t>(K + 1);
    double* o_jr = out.take<double>(J);
    double* o_cr = out.take<double>(K);
    cudaStream_t st = ctx->stream;
    if (J) {
        KO_CUDA(cudaMemcpyAsync(d_up, job_up, J * 4, cudaMemcpyHostToDevice, st));
        KO_CUDA(cudaMemcpyAsync(d_len, job_len, J * 4, cudaMemcpyHostToDevice, st));
    }
    KO_CUDA(cudaMemcpyAsync(d_off, job_offsets, (K + 1) * 8, cudaMemcpyHostToDevice, st));
    available_rate_kernel<<<(unsigned)((K + 127) / 128), 128, 0, st>>>(d_up, d_len, d_off, n_clusters, o_jr, o_cr);
    KO_CUDA(cudaGetLastError());
    ctx->counters.kernel_launches += 1;
    if (J) KO_CUDA(cudaMemcpyAsync(out_job_rate, o_jr, J * 8, cudaMemcpyDeviceToHost, st));
    KO_CUDA(cudaMemcpyAsync(out_cluster_rate, o_cr, K * 8, cudaMemcpyDeviceToHost, st));
    KO_CUDA(cudaStreamSynchronize(st));
    ctx->counters.h2d_bytes += (int64_t)(J * 8 + (K + 1) * 8);
    ctx->counters.d2h_bytes += (int64_t)(J * 8 + K * 8);
    return KO_OK;
}

// (unlocked part of ko_key_histogram_dev: the host-pointer call below holds the mutex across staging, launch and
// read-back, as the header promises)
static int key_histogram_launch(ko_ctx* ctx, const double* d_values, int64_t n, int shift, uint64_t prefix_mask,
                                uint64_t prefix, int64_t* d_out_hist, cudaStream_t st) {
    KO_CUDA(cudaMemsetAsync(d_out_hist, 0, (size_t)KO_HIST_BINS * 8, st));
    if (n) {
        int64_t blocks = (n + 255) / 256;
        if (blocks > (int64_t)ctx->num_sms * 8) blocks = (int64_t)ctx->num_sms * 8;
        key_hist_kernel<<<(unsigned)blocks, 256, 0, st>>>(d_values, n, shift, prefix_mask, prefix,
                                                         (unsigned long long*)d_out_hist);
        KO_CUDA(cudaGetLastError());
        ctx->counters.kernel_launches += 1;
    }
    return KO_OK;
}

int ko_key_histogram_dev(ko_ctx* ctx, const double* d_values, int64_t n, int shift, uint64_t prefix_mask,
                         uint64_t prefix, int64_t* d_out_hist, void* stream) {
    KO_REQUIRE(ctx && ctx->alive, "ko_key_histogram_dev: ctx not initialised");
    KO_REQUIRE(n >= 0 && d_out_hist && (n == 0 || d_values), "ko_key_histogram_dev: bad argument");
    KO_REQUIRE(shift >= 0 && shift <= 48, "ko_key_histogram_dev: shift must be 0..48");
    std::lock_guard<std::mutex> lk(ctx->mu);
    KO_CUDA(cudaSetDevice(ctx->device));
    return key_histogram_launch(ctx, d_values, n, shift, prefix_mask, prefix, d_out_hist, (cudaStream_t)stream);
}

int ko_key_histogram_state_dev(ko_ctx* ctx, const double* d_values, int64_t n, int shift, const void* d_state,
                               int64_t* d_out_hist, void* stream) {
    KO_REQUIRE(ctx && ctx->alive, "ko_key_histogram_state_dev: ctx not initialised");
    KO_REQUIRE(n >= 0 && d_out_hist && d_state && (n == 0 || d_values), "ko_key_histogram_state_dev: bad argument");
    KO_REQUIRE(shift >= 0 && shift <= 48, "ko_key_histogram_state_dev: shift must be 0..48");
    std::lock_guard<std::mutex> lk(ctx->mu);
    KO_CUDA(cudaSetDevice(ctx->device));
    cudaStream_t st = (cudaStream_t)stream;
    KO_CUDA(cudaMemsetAsync(d_out_hist, 0, (size_t)KO_HIST_BINS * 8, st));
    if (n) {
        int64_t blocks = (n + 255) / 256;
        if (blocks > (int64_t)ctx->num_sms * 8) blocks = (int64_t)ctx->num_sms * 8;
        key_hist_state_kernel<<<(unsigned)blocks, 256, 0, st>>>(d_values, n, shift, (const KoSelState*)d_state,
                                                               (unsigned long long*)d_out_hist);
        KO_CUDA(cudaGetLastError());
        ctx->counters.kernel_launches += 1;
    }
    return KO_OK;
}

int ko_key_select_dev(ko_ctx* ctx, const int64_t* d_hist, int shift, void* d_state, void* stream) {
    KO_REQUIRE(ctx && ctx->alive, "ko_key_select_dev: ctx not initialised");
    KO_REQUIRE(d_hist && d_state, "ko_key_select_dev: bad argument");
    KO_REQUIRE(shift >= 0 && shift <= 48, "ko_key_select_dev: shift must be 0..48");
    std::lock_guard<std::mutex> lk(ctx->mu);
    KO_CUDA(cudaSetDevice(ctx->device));
    key_select_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>((const unsigned long long*)d_hist, shift,
                                                            (KoSelState*)d_state);
    KO_CUDA(cudaGetLastError());
    ctx->counters.kernel_launches += 1;
    return KO_OK;
}

int ko_key_histogram(ko_ctx* ctx, const double* values, int64_t n, int shift, uint64_t prefix_mask,
                     uint64_t prefix, int64_t* out_hist) {
    KO_REQUIRE(ctx && ctx->alive, "ko_key_histogram: ctx not initialised");
    KO_REQUIRE(n >= 0 && out_hist && (n == 0 || values), "ko_key_histogram: bad argument");
    KO_REQUIRE(shift >= 0 && shift <= 48, "ko_key_histogram: shift must be 0..48");
    std::lock_guard<std::mutex> lk(ctx->mu);   // one lock for staging, launch and read-back
    KO_CUDA(cudaSetDevice(ctx->device));
    int rc;
    if ((rc = ko_buf_reserve(&ctx->in_small, (size_t)n * 8 + 256))) return rc;
    if ((rc = ko_buf_reserve(&ctx->out_small, (size_t)KO_HIST_BINS * 8))) return rc;
    double* d_v = (double*)ctx->in_small.p;
    int64_t* d_h = (int64_t*)ctx->out_small.p;
    if (n) KO_CUDA(cudaMemcpyAsync(d_v, values, (size_t)n * 8, cudaMemcpyHostToDevice, ctx->stream));
    ctx->counters.h2d_bytes += n * 8;
    if ((rc = key_histogram_launch(ctx, d_v, n, shift, prefix_mask, prefix, d_h, ctx->stream))) return rc;
    KO_CUDA(cudaMemcpyAsync(out_hist, d_h, (size_t)KO_HIST_BINS * 8, cudaMemcpyDeviceToHost, ctx->stream));
    KO_CUDA(cudaStreamSynchronize(ctx->stream));
    ctx->counters.d2h_bytes += (int64_t)KO_HIST_BINS * 8;
    return KO_OK;
}

int ko_loki_count(ko_ctx* ctx, const uint8_t* buf, const int64_t* line_offsets, int64_t n_lines,
                  const int64_t* group_offsets, int64_t n_groups, int64_t* out_count) {
    KO_REQUIRE(ctx && ctx->alive, "ko_loki_count: ctx not initialised");
    KO_REQUIRE(n_lines >= 0 && n_groups >= 0, "ko_loki_count: negative size");
    if (n_groups == 0) return KO_OK;
    KO_REQUIRE(line_offsets && group_offsets && out_count, "ko_loki_count: NULL argument");
    KO_REQUIRE(group_offsets[0] == 0 && group_offsets[n_groups] == n_lines,
               "ko_loki_count: group_offsets must cover [0, n_lines]");
    for (int64_t g = 0; g < n_groups; ++g)
        KO_REQUIRE(group_offsets[g + 1] >= group_offsets[g], "ko_loki_count: group_offsets must be non-decreasing");
    for (int64_t l = 0; l < n_lines; ++l)
        KO_REQUIRE(line_offsets[l + 1] >= line_offsets[l], "ko_loki_count: line_offsets must be non-decreasing");
    const int64_t nbytes = line_offsets[n_lines];
    KO_REQUIRE(line_offsets[0] == 0 && (nbytes == 0 || buf), "ko_loki_count: bad buffer");
    std::lock_guard<std::mutex> lk(ctx->mu);
    KO_CUDA(cudaSetDevice(ctx->device));
    const size_t B = (size_t)nbytes, NL = (size_t)n_lines, G = (size_t)n_groups;
    int rc;
    if ((rc = ko_buf_reserve(&ctx->in_small, piece(B, 1) + piece(NL + 1, 8) + piece(G + 1, 8) + 256))) return rc;
    if ((rc = ko_buf_reserve(&ctx->out_small, piece(G, 8) + 256))) return rc;
    Bump in{(unsigned char*)ctx->in_small.p, 0};
    uint8_t* d_buf = in.take<uint8_t>(B);
    int64_t* d_lo = in.take<int64_t>(NL + 1);
    int64_t* d_go = in.take<int64_t>(G + 1);
    unsigned long long* d_out = (unsigned long long*)ctx->out_small.p;
    cudaStream_t st = ctx->stream;
    if (B) KO_CUDA(cudaMemcpyAsync(d_buf, buf, B, cudaMemcpyHostToDevice, st));
    KO_CUDA(cudaMemcpyAsync(d_lo, line_offsets, (NL + 1) * 8, cudaMemcpyHostToDevice, st));
    KO_CUDA(cudaMemcpyAsync(d_go, group_offsets, (G + 1) * 8, cudaMemcpyHostToDevice, st));
    KO_CUDA(cudaMemsetAsync(d_out, 0, G * 8, st));
    if (NL) {
        loki_count_kernel<<<(unsigned)((NL * 32 + 255) / 256), 256, 0, st>>>(d_buf, d_lo, n_lines, d_go, n_groups, d_out);
        KO_CUDA(cudaGetLastError());
        ctx->counters.kernel_launches += 1;
    }
    KO_CUDA(cudaMemcpyAsync(out_count, d_out, G * 8, cudaMemcpyDeviceToHost, st));
    KO_CUDA(cudaStreamSynchronize(st));
    ctx->counters.h2d_bytes += (int64_t)(B + (NL + 1) * 8 + (G + 1) * 8);
    ctx->counters.d2h_bytes += (int64_t)(G * 8);
    return KO_OK;
}

int ko_counter_increase(ko_ctx* ctx, const double* values, const int64_t* offsets, int64_t n_series,
                        double step_seconds, double* out_increase, double* out_rate) {
    KO_REQUIRE(ctx && ctx->alive, "ko_counter_increase: ctx not initialised");
    KO_REQUIRE(n_series >= 0, "ko_counter_increase: negative size");
    if (n_series == 0) return KO_OK;
    KO_REQUIRE(offsets && offsets[0] == 0, "ko_counter_increase: offsets must start at 0");
    for (int64_t i = 0; i < n_series; ++i)
        KO_REQUIRE(offsets[i + 1] >= offsets[i], "ko_counter_increase: offsets must be non-decreasing");
    const size_t T = (size_t)offsets[n_series], S = (size_t)n_series;
    KO_REQUIRE(T == 0 || values, "ko_counter_increase: values is NULL");
    KO_REQUIRE(step_seconds > 0, "ko_counter_increase: step_seconds must be positive");
    std::lock_guard<std::mutex> lk(ctx->mu);
    KO_CUDA(cudaSetDevice(ctx->device));
    int rc;
    if ((rc = ko_buf_reserve(&ctx->stage[0], T * 8 + 256))) return rc;
    if ((rc = ko_buf_reserve(&ctx->in_small, (S + 1) * 8 + 256))) return rc;
    if ((rc = ko_buf_reserve(&ctx->out_small, 2 * piece(S, 8) + 256))) return rc;
    double* d_v = (double*)ctx->stage[0].p;
    int64_t* d_off = (int64_t*)ctx->in_small.p;
    Bump out{(unsigned char*)ctx->out_small.p, 0};
    double* o_inc = out.take<double>(S);
    double* o_rate = out.take<double>(S);
    cudaStream_t st = ctx->stream;
    if (T) KO_CUDA(cudaMemcpyAsync(d_v, values, T * 8, cudaMemcpyHostToDevice, st));
    KO_CUDA(cudaMemcpyAsync(d_off, offsets, (S + 1) * 8, cudaMemcpyHostToDevice, st));
    counter_increase_kernel<<<(unsigned)S, 256, 0, st>>>(d_v, d_off, step_seconds, o_inc, o_rate);
    KO_CUDA(cudaGetLastError());
    ctx->counters.kernel_launches += 1;
    if (out_increase) KO_CUDA(cudaMemcpyAsync(out_increase, o_inc, S * 8, cudaMemcpyDeviceToHost, st));
    if (out_rate) KO_CUDA(cudaMemcpyAsync(out_rate, o_rate, S * 8, cudaMemcpyDeviceToHost, st));
    KO_CUDA(cudaStreamSynchronize(st));
    ctx->counters.h2d_bytes += (int64_t)(T * 8 + (S + 1) * 8);
    ctx->counters.d2h_bytes += (int64_t)(S * 8 * ((out_increase != 0) + (out_rate != 0)));
    return KO_OK;
}

}  // extern "C"
