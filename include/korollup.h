/* korollup.h — C ABI of libkorollup.so, the B200-native (sm_100a) monitoring rollup for KubeOperator.
 *
 * This library replaces ONLY the arithmetic of KubeOperator's Day-2 monitoring rollup; the Django/Celery
 * code keeps its REST / Redis / JSON surface and calls these entry points through ctypes
 * (kubeoperator_b200/korollup.py; reference-side patch in INTEGRATION.md).  The reference has no
 * FFI of its own (it is pure Python), so each entry point cites the reference lines whose loop it
 * replaces.  Paths are relative to /root/reference.
 *
 * Conventions
 *   - All buffers are caller-allocated, C-contiguous, native endian.  `*_dev` entry points take DEVICE
 *     pointers and a CUDA stream (as void*) and only enqueue work (ko_series_status is the call that waits for
 *     the stream and reads the flags back); the others take HOST pointers, do their own H2D/D2H and return
 *     when the outputs are valid.
 *   - Return 0 (KO_OK) or a negative KO_E* code; ko_last_error() gives the message (thread-local,
 *     library-owned).  There is NO CPU fallback: without a CUDA device every compute call fails.
 *   - CUDA is initialised lazily inside ko_init(), never at load time, so the library may be dlopen'ed
 *     before a Celery prefork (core/kubeops.py:154-160); call ko_init() in the child.
 *   - One ko_ctx per process and GPU; calls on one ctx are serialised by an internal mutex.
 *   - NaN in a series is the Prometheus staleness marker and is skipped.
 */
#ifndef KOROLLUP_H
#define KOROLLUP_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define KO_OK 0
#define KO_EINVAL (-1) /* bad argument (NULL, negative size, unsorted cluster ids, too many quantiles) */
#define KO_ECUDA (-2)  /* CUDA runtime / driver error, message has the CUDA string                    */
#define KO_ENOMEM (-3) /* device or pinned allocation failed                                          */
#define KO_ESTATE (-4) /* ctx not initialised / destroyed                                              */

#define KO_MAX_QUANTILES 4     /* quantiles per call (the dashboard asks for p50/p95/p99)              */
#define KO_BAND_NONE 255       /* ProgressColorPipe returns null                                       */
#define KO_BAND_GREEN 0        /* <= 60                                                                */
#define KO_BAND_YELLOW 1       /* (60, 80]                                                             */
#define KO_BAND_RED 2          /* > 80                                                                 */
#define KO_HIST_BINS 65536     /* bins of one fleet-quantile refinement round                          */
#define KO_STATIC_MAX_SERIES 65536 /* largest batch of ko_series_rollup_static                         */

typedef struct ko_ctx ko_ctx;

typedef struct ko_counters {
    int64_t kernel_launches; /* kernels of this library launched through the ctx since ko_init       */
    int64_t h2d_bytes;       /* bytes copied host->device by host-pointer entry points               */
    int64_t d2h_bytes;       /* bytes copied device->host by host-pointer entry points               */
    int64_t series_fallback; /* long series the first attempt of the one-pass kernel could not finish (last call)  */
} ko_counters;

/* ---- life cycle (no reference counterpart: the reference is pure Python) ------------------------- */
int ko_version(void);
const char* ko_last_error(void);
/* devices/n_devices: one process drives one GPU, so n_devices must be 1 (devices[0] = CUDA ordinal). */
int ko_init(const int* devices, int n_devices, ko_ctx** out);
void ko_destroy(ko_ctx* ctx);
int ko_get_counters(ko_ctx* ctx, ko_counters* out);
/* Device time of the long-series kernels (series_flow, its second attempt, series_long) of the most recent series
 * rollup on this ctx, measured with CUDA events on the launching stream; waits for them to finish.
 * ko_last_series_pipe_ms: the dominant kernel alone - the first launch of the one-pass kernel of that rollup
 * (series_flow today; the entry point keeps the name it had when that kernel was called series_pipe).
 * Used for roofline reports; no counterpart in the reference. */
int ko_last_series_kernel_ms(ko_ctx* ctx, float* out_ms);
int ko_last_series_pipe_ms(ko_ctx* ctx, float* out_ms);
/* Page-lock / unlock a caller buffer so host-pointer entry points copy at full PCIe speed. */
int ko_pin(void* ptr, int64_t bytes);
int ko_unpin(void* ptr);

/* ---- a2: per-cluster node rollup --------------------------------------------------------------------
 * Replaces the loop of ClusterMonitor.set_cluster_data, core/apps/kubeops_api/cluster_monitor.py:163-182
 * for many clusters at once: cpu_total = sum cpu, mem_total = sum mem, cpu_usage/mem_usage = sum/count
 * with count = nodes - #{cpu_usage==0 and mem_usage==0} (:174-175; sums are left undivided when
 * count==0, :180); out_alert_flag[i]=1 where the reference enters the alert branch :176 (usage >
 * threshold, evaluated only for nodes that are not all-zero).  cluster_of_node must be non-decreasing
 * (nodes grouped by cluster, as the reference visits them).  The sums are taken node by node in input order,
 * so the four floats equal the reference's bit for bit.  out_band_* (may be NULL) give the UI colour
 * band of usage*100 per node: ui/src/app/shared/pipe/progress-color.pipe.ts:8-23. */
int ko_cluster_rollup(ko_ctx* ctx, const double* cpu, const double* mem, const double* cpu_usage,
                      const double* mem_usage, const int32_t* cluster_of_node, int64_t n_nodes,
                      int32_t n_clusters, double threshold, double* out_cpu_total, double* out_mem_total,
                      double* out_cpu_usage, double* out_mem_usage, int32_t* out_count,
                      uint8_t* out_alert_flag, uint8_t* out_band_cpu, uint8_t* out_band_mem);

/* ---- a4 + a5: pod rollup and the two sorted lists --------------------------------------------------
 * Replaces the inner loop of ClusterMonitor.list_pods, cluster_monitor.py:93-117 (restart_count = sum of
 * container restart counts :98; warn container = ready==False :101; restart member = restart_count>0
 * :113; error member = phase not Running/Succeeded :115) and the two quick_sort_pods calls of :183-184
 * (cluster_monitor.py:450-463: a STABLE DESCENDING sort by restart_count).
 * pod_offsets[n_pods+1] index the container arrays.  out_order_restart_desc / out_order_error_desc have
 * room for n_pods indices; the first *out_n_* entries are the members in the reference's sorted order. */
int ko_pod_rollup(ko_ctx* ctx, const int32_t* container_restart, const uint8_t* container_ready,
                  const int64_t* pod_offsets, const uint8_t* pod_phase_ok, int64_t n_pods,
                  int32_t* out_restart, uint8_t* out_warn, int64_t* out_order_restart_desc,
                  int64_t* out_n_restart, int64_t* out_order_error_desc, int64_t* out_n_error);

/* ---- a5 alone: stable descending order of int32 keys ------------------------------------------------
 * Replaces quick_sort_pods / quick_sort_error_loki_container (cluster_monitor.py:450-479) where they are
 * applied to an existing list: the fleet re-sort of DashBoardView.get, core/apps/kubeops_api/api.py:500-501,
 * and set_loki_data_to_cluster, cluster_monitor.py:235.  out_order[i] = index of the i-th item. */
int ko_stable_desc_order(ko_ctx* ctx, const int32_t* keys, int64_t n, int64_t* out_order);

/* ---- window statistics over columnar series (BUILD-DEFINED) -------------------------
 * No reference code computes range-window statistics (Prometheus evaluates rate(...[5m]) server side,
 * core/apps/kubeops_api/prometheus_client.py:90); the specification is oracle/rollup_oracle.py
 * series_stats(): per series sum, count (non-NaN), mean (=sum/count, 0 when count==0 like the guard at
 * cluster_monitor.py:180), min, max, last valid sample and nearest-rank quantiles
 * sorted(valid)[ceil(q*n)-1]; zero results are +0.0; min/max/last/quantiles are NaN for an empty series.
 * values: f64 samples of all series back to back; offsets[n_series+1] (int64, non-decreasing).
 * q[nq] in [0,1], nq <= KO_MAX_QUANTILES; out_quantiles is [n_series][nq].  Any out_* may be NULL. */
int ko_series_rollup(ko_ctx* ctx, const double* values, const int64_t* offsets, int64_t n_series,
                     const double* q, int nq, double* out_sum, double* out_mean, double* out_min,
                     double* out_max, double* out_last, int64_t* out_count, double* out_quantiles);
/* Same with DEVICE pointers (q stays a host array); enqueues on `stream` and returns.  The offsets are then
 * checked on the GPU only: a series whose offsets decrease is skipped (its outputs are left untouched) and the
 * device error flag is raised: read it with ko_series_status() after the call.  Scratch of the
 * context is reused by every call: calls on one context must not overlap on the GPU. */
int ko_series_rollup_dev(ko_ctx* ctx, const double* d_values, const int64_t* d_offsets,
                         int64_t n_series, const double* q, int nq, double* d_out_sum,
                         double* d_out_mean, double* d_out_min, double* d_out_max, double* d_out_last,
                         int64_t* d_out_count, double* d_out_quantiles, void* stream);

/* Outcome of the most recent ko_series_rollup_dev on this ctx as the GPU saw it (waits for that call's kernels):
 * out[0] = long series (> 1024 samples) of the call, out[1] = series whose first one-pass attempt had to be repeated
 * with wider intervals, out[2] = series that went on to the exact multi-pass slow path, out[3] = device error flag
 * (non-zero: at least one series had decreasing offsets or a length >= 2^31 and was skipped).  The host-pointer
 * ko_series_rollup reports the same through its return code and ko_counters.  No counterpart in the reference. */
int ko_series_status(ko_ctx* ctx, int32_t* out4);

/* Same kernels and results as ko_series_rollup_dev, for callers that must not allocate device memory when
 * they ask for a rollup: a worker whose GPU memory belongs entirely to another component's pool, a request
 * thread that cannot afford the device-wide synchronisation hidden in cudaMalloc / cudaFree, or any section of
 * a program in which the CUDA allocator is off limits.  This entry point has NO COUNTERPART IN THE REFERENCE
 * (and needs no ko_ctx): the work lists, counters and candidate spill space that ko_series_rollup_dev keeps in
 * memory it cudaMallocs on demand live here in a fixed array inside the library image - 6.6 MiB: counters and two
 * lists of 4 * KO_STATIC_MAX_SERIES ints (2 MiB) plus 4.6 MiB of spill space - which every GPU that loads the
 * library holds until the process exits.  It calls no cudaMalloc, cudaFree or other allocating routine, and the
 * free device memory reported by cudaMemGetInfo is the same before and after a call.  Limits and rules:
 *   - n_series > KO_STATIC_MAX_SERIES is refused with KO_EINVAL and a message naming the limit (split the batch);
 *   - the first call on a device looks up the array's device address (cudaGetSymbolAddress, once per device) and
 *     sets the kernels' shared-memory attributes; later calls only clear the counters and launch kernels;
 *   - all calls on a device share the array, which keeps the previous call's contents (its counters are
 *     cleared on `stream` at the start of each call): calls whose kernels could overlap on the GPU race on it,
 *     so issue them on one stream or otherwise serialise them;
 *   - the spill space is small: series of more than about 100 000 samples are computed by the exact multi-pass
 *     kernel instead of the one-pass kernel (same results, about a third of the speed).
 * `device` is the CUDA ordinal; it is made current.  Pointers are device pointers as in ko_series_rollup_dev. */
int ko_series_rollup_static(int device, const double* d_values, const int64_t* d_offsets, int64_t n_series,
                            const double* q, int nq, double* d_out_sum, double* d_out_mean, double* d_out_min,
                            double* d_out_max, double* d_out_last, int64_t* d_out_count, double* d_out_quantiles,
                            void* stream);

/* ---- a7: scrape-target availability rate --------------------------------------------------------------
 * Replaces PrometheusClient.calculate_available_rate, core/apps/kubeops_api/prometheus_client.py:75-85, for
 * many clusters: jobs of cluster c are job_offsets[c]..job_offsets[c+1]; per job rate = up/len*100 (0 when
 * len==0, :82); cluster rate = #{rate==100}/n_jobs*100 (:83-85). */
int ko_available_rate(ko_ctx* ctx, const int32_t* job_up, const int32_t* job_len,
                      const int64_t* job_offsets, int64_t n_clusters, double* out_job_rate,
                      double* out_cluster_rate);

/* ---- a11: fleet roll-up the browser performs ------------------------------------------------------------
 * Replaces DashboardComponent.getClusterData, ui/src/app/dashboard/dashboard.component.ts:117-153.
 * counts is [n_clusters][6] = pods, namespaces, deployments, containers, nodes (all of the cluster's blob), and
 * ui_nodes = cluster.nodes.length of the UI's own cluster list, or -1 when that list does not hold the cluster (it then
 * adds no pod capacity, :136-143).  max_pods[c] < 0 means MAX_PODS === undefined (-> 110, :138-141); 0 is a configured
 * zero.  out[12] = pod_count, namespace_count, deployment_count, container_count, node_count, cpu_total, mem_total,
 * cpu_usage(%), mem_usage(%), max_pod_count, container_usage(%), count. */
int ko_fleet_rollup(ko_ctx* ctx, const int64_t* counts, const double* cpu_total,
                    const double* mem_total, const double* cpu_usage, const double* mem_usage,
                    const int32_t* max_pods, int64_t n_clusters, double* out);

/* ---- e: one round of the fleet-wide quantile merge (BUILD-DEFINED) ----------------------
 * Histogram of digit ((key >> shift) & 0xFFFF) over the valid values whose (key & prefix_mask)==prefix,
 * key = order-preserving u64 image of the f64 (zero canonicalised to +0.0).  Ranks all-gather the 65536
 * bins (NCCL via torch.distributed in kubeoperator_b200/fleet.py), pick the digit holding the target
 * rank and repeat for shift = 48, 32, 16, 0.  out_hist has KO_HIST_BINS int64 entries. */
int ko_key_histogram(ko_ctx* ctx, const double* values, int64_t n, int shift, uint64_t prefix_mask,
                     uint64_t prefix, int64_t* out_hist);
int ko_key_histogram_dev(ko_ctx* ctx, const double* d_values, int64_t n, int shift,
                         uint64_t prefix_mask, uint64_t prefix, int64_t* d_out_hist, void* stream);

/* The same round with the refinement state in DEVICE memory, so that the four rounds of a fleet quantile follow each
 * other on the stream without a trip to the host: ko_key_histogram_state_dev (local histogram under the state's
 * prefix), an all-reduce of the 65536 bins across the ranks (NCCL, by the caller), ko_key_select_dev (one CTA finds the
 * digit holding the wanted rank and updates the state).  d_state points to 32 bytes: {uint64 prefix, uint64 mask,
 * int64 rank (1-based; < 1: no valid value), double result}; start with prefix = mask = 0 and
 * rank = clamp(ceil(q * n_valid_total), 1, n_valid_total); after the round with shift 0 `result` holds the exact
 * nearest-rank quantile (NaN when the fleet has no valid value).  BUILD-DEFINED, no counterpart in the reference. */
int ko_key_histogram_state_dev(ko_ctx* ctx, const double* d_values, int64_t n, int shift, const void* d_state,
                               int64_t* d_out_hist, void* stream);
int ko_key_select_dev(ko_ctx* ctx, const int64_t* d_hist, int shift, void* d_state, void* stream);

/* ---- a6: Loki error-line count ------------------------------------------------------
 * Replaces the scan of PrometheusClient.get_msg_from_loki, prometheus_client.py:140-146: per container
 * (group) the number of lines containing the bytes "level=error".  buf holds all lines back to back,
 * line_offsets[n_lines+1] delimit them, group_offsets[n_groups+1] index line_offsets. */
int ko_loki_count(ko_ctx* ctx, const uint8_t* buf, const int64_t* line_offsets, int64_t n_lines,
                  const int64_t* group_offsets, int64_t n_groups, int64_t* out_count);

/* ---- §8f rank 2: increase()/rate() with counter-reset handling (BUILD-DEFINED) -------------------------
 * The [5m] window the reference delegates to Prometheus (prometheus_client.py:90).  Specification:
 * oracle/rollup_oracle.py counter_increase(): sum of deltas between consecutive valid samples, a sample
 * lower than its predecessor counts as a reset and contributes its own value; rate = increase / seconds
 * between first and last valid sample (0 with fewer than two valid samples). */
int ko_counter_increase(ko_ctx* ctx, const double* values, const int64_t* offsets, int64_t n_series,
                        double step_seconds, double* out_increase, double* out_rate);

/* ---- §8f rank 1: Prometheus HTTP API response -> columnar samples (HOST code, no GPU needed) -------------
 * The reference reads Prometheus with requests.json() and picks ['data']['result'][0]['value'][1], a string,
 * one node and one query at a time (prometheus_client.py:91-116; scrape targets :67-72).  These calls turn a
 * whole response body -- a query_range "matrix" or an instant "vector" -- into the columnar layout the rollup
 * entry points take: ko_prom_open scans the structure (fails with KO_ESTATE when "status" is not "success",
 * the reference's guard at :94, and with KO_EINVAL on malformed JSON), ko_prom_dims reports sizes,
 * ko_prom_fill converts the sample strings (correctly rounded, as Python's float(); "NaN", "+Inf", "-Inf")
 * into caller buffers -- values[n_samples], optional timestamps[n_samples], offsets[n_series+1], and
 * optionally the byte range of every series' "metric" object for the caller's label handling -- on
 * n_threads host threads.  The body must stay valid until ko_prom_close.  Fill straight into a ko_pin()ned
 * buffer to hand the result to ko_series_rollup without another copy. */
typedef struct ko_prom_doc ko_prom_doc;
int ko_prom_open(const char* json, int64_t len, ko_prom_doc** out);
int ko_prom_dims(const ko_prom_doc* doc, int64_t* n_series, int64_t* n_samples);
int ko_prom_fill(const ko_prom_doc* doc, double* values, double* timestamps, int64_t* offsets,
                 int64_t* metric_begin, int64_t* metric_end, int n_threads);
void ko_prom_close(ko_prom_doc* doc);

#ifdef __cplusplus
}
#endif
#endif /* KOROLLUP_H */
