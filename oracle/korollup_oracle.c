/* TEST INFRASTRUCTURE ONLY — plain-C restatement (oracle) of the KubeOperator monitoring rollup.
 *
 * Same specification as oracle/rollup_oracle.py, for sizes the Python form cannot finish in seconds and
 * as the timed CPU baseline of bench.py (cpu_baseline kind "port", and `--impl reference`).  Only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load this library;
 * nothing under kubeoperator_b200/ does.
 *
 * Reference lines restated (paths relative to /root/reference):
 *   koo_cluster_rollup   core/apps/kubeops_api/cluster_monitor.py:163-182   (left-to-right float sums,
 *                        zero-node exclusion :174-175, alert condition :176, count guard :180-182)
 *   koo_pod_rollup       core/apps/kubeops_api/cluster_monitor.py:93-117    (restart sum, ready / phase)
 *   koo_stable_desc_order core/apps/kubeops_api/cluster_monitor.py:450-463  (stable descending order)
 *   koo_available_rate   core/apps/kubeops_api/prometheus_client.py:75-85
 *   koo_loki_count       core/apps/kubeops_api/prometheus_client.py:140-146 ('level=error' in line)
 *   koo_series_rollup    BUILD-DEFINED (no reference code) — parity unpinned; follows
 *                        the reference style: strict left-to-right double accumulation, count guard.
 * Pinned by tests/test_oracle_golden.py against tests/golden/ (reference outputs) through the Python
 * form, and tests/test_oracle_c.py checks this C form against the Python form on seeded inputs.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

static int64_t nearest_rank(int64_t n, double q) {
    double r = ceil(q * (double)n);
    int64_t k = (int64_t)r;
    if (k < 1) k = 1;
    if (k > n) k = n;
    return k;
}

/* k-th smallest (0-based) of a[0..n) by value; partitions a in place (Hoare, median of three). */
static double select_kth(double* a, int64_t n, int64_t k) {
    int64_t lo = 0, hi = n - 1;
    while (lo < hi) {
        int64_t mid = lo + (hi - lo) / 2;
        double x = a[lo], y = a[mid], z = a[hi];
        double p = (x < y) ? ((y < z) ? y : (x < z ? z : x)) : ((x < z) ? x : (y < z ? z : y));
        int64_t i = lo, j = hi;
        while (i <= j) {
            while (a[i] < p) i++;
            while (a[j] > p) j--;
            if (i <= j) { double t = a[i]; a[i] = a[j]; a[j] = t; i++; j--; }
        }
        if (k <= j) hi = j;
        else if (k >= i) lo = i;
        else return a[k];
    }
    return a[k];
}

int koo_series_rollup(const double* values, const int64_t* offsets, int64_t n_series,
                      const double* q, int nq, double* out_sum, double* out_mean, double* out_min,
                      double* out_max, double* out_last, int64_t* out_count, double* out_quant,
                      int nthreads) {
    int err = 0;
#ifdef _OPENMP
    if (nthreads > 0) omp_set_num_threads(nthreads);
#endif
#pragma omp parallel
    {
        double* tmp = NULL;
        int64_t cap = 0;
#pragma omp for schedule(dynamic, 4)
        for (int64_t s = 0; s < n_series; ++s) {
            const double* v = values + offsets[s];
            int64_t L = offsets[s + 1] - offsets[s];
            double total = 0.0, mn = NAN, mx = NAN, last = NAN;
            int64_t cnt = 0;
            if (nq > 0 && L > cap) {
                free(tmp);
                tmp = (double*)malloc(sizeof(double) * (size_t)L);
                cap = L;
                if (!tmp) { err = 1; cap = 0; }
            }
            for (int64_t i = 0; i < L; ++i) {
                double x = v[i];
                if (x != x) continue;
                total = total + x;
                if (cnt == 0 || x < mn) mn = x;
                if (cnt == 0 || x > mx) mx = x;
                last = x;
                if (nq > 0 && tmp) tmp[cnt] = x;
                cnt++;
            }
            if (out_sum) out_sum[s] = total;
            if (out_mean) out_mean[s] = cnt > 0 ? total / (double)cnt : 0.0;
            if (out_min) out_min[s] = cnt > 0 ? mn + 0.0 : NAN;
            if (out_max) out_max[s] = cnt > 0 ? mx + 0.0 : NAN;
            if (out_last) out_last[s] = last;
            if (out_count) out_count[s] = cnt;
            for (int j = 0; j < nq; ++j) {
                double r = NAN;
                if (cnt > 0 && tmp) r = select_kth(tmp, cnt, nearest_rank(cnt, q[j]) - 1) + 0.0;
                if (out_quant) out_quant[s * nq + j] = r;
            }
        }
        free(tmp);
    }
    return err;
}

/* cluster_monitor.py:163-182 per cluster; nodes of a cluster are visited in input order. */
int koo_cluster_rollup(const double* cpu, const double* mem, const double* cpu_usage,
                       const double* mem_usage, const int32_t* cluster_of_node, int64_t n_nodes,
                       int32_t n_clusters, double threshold, double* out_cpu_total,
                       double* out_mem_total, double* out_cpu_usage, double* out_mem_usage,
                       int32_t* out_count, uint8_t* out_flag) {
    for (int32_t c = 0; c < n_clusters; ++c) {
        out_cpu_total[c] = out_mem_total[c] = out_cpu_usage[c] = out_mem_usage[c] = 0.0;
        out_count[c] = 0;
    }
    for (int64_t i = 0; i < n_nodes; ++i) {
        int32_t c = cluster_of_node[i];
        if (c < 0 || c >= n_clusters) return -1;
        out_count[c] += 1;                                  /* :167 count = len(nodes) */
        out_cpu_total[c] = out_cpu_total[c] + cpu[i];       /* :170 */
        out_cpu_usage[c] = out_cpu_usage[c] + cpu_usage[i]; /* :171 */
        out_mem_total[c] = out_mem_total[c] + mem[i];       /* :172 */
        out_mem_usage[c] = out_mem_usage[c] + mem_usage[i]; /* :173 */
        uint8_t f = 0;
        if (cpu_usage[i] == 0 && mem_usage[i] == 0) out_count[c] -= 1;            /* :174-175 */
        else if (cpu_usage[i] > threshold || mem_usage[i] > threshold) f = 1;     /* :176 */
        if (out_flag) out_flag[i] = f;
    }
    for (int32_t c = 0; c < n_clusters; ++c)
        if (out_count[c] > 0) {                             /* :180 */
            out_cpu_usage[c] = out_cpu_usage[c] / out_count[c];
            out_mem_usage[c] = out_mem_usage[c] / out_count[c];
        }
    return 0;
}

/* stable descending order of int32 keys: bottom-up merge sort on indices (equivalent to :450-463). */
int koo_stable_desc_order(const int32_t* keys, int64_t n, int64_t* order) {
    int64_t* tmp = (int64_t*)malloc(sizeof(int64_t) * (size_t)(n > 0 ? n : 1));
    if (!tmp) return 1;
    for (int64_t i = 0; i < n; ++i) order[i] = i;
    int64_t *src = order, *dst = tmp;
    for (int64_t w = 1; w < n; w *= 2) {
        for (int64_t lo = 0; lo < n; lo += 2 * w) {
            int64_t mid = lo + w < n ? lo + w : n, hi = lo + 2 * w < n ? lo + 2 * w : n;
            int64_t i = lo, j = mid, k = lo;
            while (i < mid && j < hi) dst[k++] = (keys[src[j]] > keys[src[i]]) ? src[j++] : src[i++];
            while (i < mid) dst[k++] = src[i++];
            while (j < hi) dst[k++] = src[j++];
        }
        int64_t* t = src; src = dst; dst = t;
    }
    if (src != order) memcpy(order, src, sizeof(int64_t) * (size_t)n);
    free(tmp);
    return 0;
}

/* cluster_monitor.py:93-117 + the two sorts of :183-184. */
int koo_pod_rollup(const int32_t* container_restart, const uint8_t* container_ready,
                   const int64_t* pod_offsets, const uint8_t* pod_phase_ok, int64_t n_pods,
                   int32_t* out_restart, uint8_t* out_warn, int64_t* out_order_restart,
                   int64_t* out_n_restart, int64_t* out_order_error, int64_t* out_n_error) {
    int64_t nr = 0, ne = 0;
    int64_t* mem_r = (int64_t*)malloc(sizeof(int64_t) * (size_t)(n_pods + 1));
    int64_t* mem_e = (int64_t*)malloc(sizeof(int64_t) * (size_t)(n_pods + 1));
    int32_t* key = (int32_t*)malloc(sizeof(int32_t) * (size_t)(n_pods + 1));
    int64_t* ord = (int64_t*)malloc(sizeof(int64_t) * (size_t)(n_pods + 1));
    if (!mem_r || !mem_e || !key || !ord) return 1;
    for (int64_t p = 0; p < n_pods; ++p) {
        int32_t rc = 0;                                                   /* :93 */
        for (int64_t c = pod_offsets[p]; c < pod_offsets[p + 1]; ++c) {
            rc = rc + container_restart[c];                               /* :98 */
            if (out_warn) out_warn[c] = container_ready[c] ? 0 : 1;       /* :101 */
        }
        out_restart[p] = rc;
        if (rc > 0) mem_r[nr++] = p;                                      /* :113 */
        if (!pod_phase_ok[p]) mem_e[ne++] = p;                            /* :115 */
    }
    for (int64_t i = 0; i < nr; ++i) key[i] = out_restart[mem_r[i]];
    koo_stable_desc_order(key, nr, ord);
    for (int64_t i = 0; i < nr; ++i) out_order_restart[i] = mem_r[ord[i]];
    for (int64_t i = 0; i < ne; ++i) key[i] = out_restart[mem_e[i]];
    koo_stable_desc_order(key, ne, ord);
    for (int64_t i = 0; i < ne; ++i) out_order_error[i] = mem_e[ord[i]];
    *out_n_restart = nr;
    *out_n_error = ne;
    free(mem_r); free(mem_e); free(key); free(ord);
    return 0;
}

/* prometheus_client.py:75-85 */
int koo_available_rate(const int32_t* job_up, const int32_t* job_len, int32_t n_jobs,
                       double* out_job_rate, double* out_rate) {
    int32_t service_up = 0;
    for (int32_t j = 0; j < n_jobs; ++j) {
        double r = job_len[j] > 0 ? (double)job_up[j] / (double)job_len[j] * 100 : 0;   /* :82 */
        if (r == 100) service_up += 1;                                                  /* :83 */
        out_job_rate[j] = r;
    }
    *out_rate = n_jobs > 0 ? (double)service_up / (double)n_jobs * 100 : 0;             /* :85 */
    return 0;
}

/* prometheus_client.py:142-145 per group of lines: count lines containing "level=error". */
int koo_loki_count(const uint8_t* buf, const int64_t* line_offsets, const int64_t* group_offsets,
                   int64_t n_groups, int64_t* out_count) {
    static const char needle[] = "level=error";
    const int64_t m = 11;
    for (int64_t g = 0; g < n_groups; ++g) {
        int64_t cnt = 0;
        for (int64_t l = group_offsets[g]; l < group_offsets[g + 1]; ++l) {
            int64_t a = line_offsets[l], b = line_offsets[l + 1];
            int hit = 0;
            for (int64_t i = a; i + m <= b && !hit; ++i)
                if (memcmp(buf + i, needle, (size_t)m) == 0) hit = 1;
            cnt += hit;
        }
        out_count[g] = cnt;
    }
    return 0;
}

/* BUILD-DEFINED (§8f rank 2): increase() with counter-reset handling, NaN skipped. */
int koo_counter_increase(const double* values, const int64_t* offsets, int64_t n_series,
                         double step_seconds, double* out_increase, double* out_rate) {
    for (int64_t s = 0; s < n_series; ++s) {
        double inc = 0.0, prev = 0.0;
        int64_t first = -1, last = -1;
        for (int64_t i = offsets[s]; i < offsets[s + 1]; ++i) {
            double v = values[i];
            if (v != v) continue;
            if (first < 0) first = i;
            else inc = inc + (v >= prev ? v - prev : v);
            prev = v;
            last = i;
        }
        if (first == last) { inc = 0.0; first = last = -1; }
        if (out_increase) out_increase[s] = inc;
        if (out_rate) out_rate[s] = first < 0 ? 0.0 : inc / ((double)(last - first) * step_seconds);
    }
    return 0;
}
