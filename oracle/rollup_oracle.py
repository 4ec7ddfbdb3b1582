"""TEST INFRASTRUCTURE ONLY — CPU restatement (oracle) of KubeOperator's monitoring rollup.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / ``--impl reference`` legs
may import this module, and only as the checker.  Nothing under ``kubeoperator_b200/`` imports it.

Every function cites the reference lines it restates (paths relative to /root/reference).  Two kinds:

* REFERENCE-DEFINED (pinned against the reference's own functions executed under stubs by
  ``oracle/gen_golden.py`` -> ``tests/golden/reference_*.json``): node rollup, pod rollup, the two quick
  sorts, available rate, Loki error count, health day mean.
* BUILD-DEFINED (the reference has no implementation: SURVEY.md §0.3): range-window statistics, nearest
  rank quantiles, fleet quantile / top-k merge, rate()/increase().  They follow the reference's *style*
  (left-to-right ``float`` accumulation as in cluster_monitor.py:170-173, ``if count > 0`` guard as in
  :180).  **Parity unpinned** for these — the oracle is the specification.
* The fleet roll-up of dashboard.component.ts:117-153 is TypeScript; ``node`` is absent, so it is
  restated line by line and pinned only by hand-computed vectors (**parity unpinned** by execution).

NumPy twins (``*_np``) compute bit-identical results for large inputs (``np.cumsum`` is a strict
left-to-right accumulation) and are checked against the pure-Python forms in tests.
"""
import math
import struct

import numpy as np

NAN = float("nan")


# --------------------------------------------------------------------------------------------------
# a2 — ClusterMonitor.set_cluster_data aggregation loop   core/apps/kubeops_api/cluster_monitor.py:163-182
# --------------------------------------------------------------------------------------------------
def node_rollup(nodes, threshold=80):
    """Restates cluster_monitor.py:163-182.

    ``nodes``: list of dicts with keys cpu, mem, cpu_usage, mem_usage (numbers — the reference raises on
    the strings Prometheus returns; the adapter converts with float() as :170-173 do).
    Returns dict(cpu_total, mem_total, cpu_usage, mem_usage, count, flags) where ``flags[i]`` is 1 when
    the reference would enter the alert branch :176 for node i (the branch body raises in the reference,
    SURVEY §0.5b, so only the condition is restated).
    """
    cpu_usage = 0
    cpu_total = 0
    mem_total = 0
    mem_usage = 0
    count = len(nodes)                                            # :167
    flags = []
    for n in nodes:                                               # :168
        cpu_total = cpu_total + float(n["cpu"])                   # :170
        cpu_usage = cpu_usage + float(n["cpu_usage"])             # :171
        mem_total = mem_total + float(n["mem"])                   # :172
        mem_usage = mem_usage + float(n["mem_usage"])             # :173
        flag = 0
        if n["cpu_usage"] == 0 and n["mem_usage"] == 0:           # :174
            count = count - 1                                     # :175
        elif n["cpu_usage"] > threshold or n["mem_usage"] > threshold:   # :176
            flag = 1
        flags.append(flag)
    if count > 0:                                                 # :180
        cpu_usage = cpu_usage / count                             # :181
        mem_usage = mem_usage / count                             # :182
    return dict(cpu_total=cpu_total, mem_total=mem_total, cpu_usage=cpu_usage, mem_usage=mem_usage,
                count=count, flags=flags)


def node_rollup_np(cpu, mem, cpu_usage, mem_usage, threshold=80.0):
    """Vector twin of :func:`node_rollup` for one cluster (strict left-to-right sums via cumsum)."""
    cpu, mem = np.asarray(cpu, np.float64), np.asarray(mem, np.float64)
    cu, mu = np.asarray(cpu_usage, np.float64), np.asarray(mem_usage, np.float64)
    n = len(cpu)
    seq = lambda a: float(np.cumsum(a)[-1]) if n else 0.0
    zero = (cu == 0) & (mu == 0)
    with np.errstate(invalid="ignore"):
        flags = (~zero) & ((cu > threshold) | (mu > threshold))
    count = int(n - zero.sum())
    cpu_total, mem_total, cus, mus = seq(cpu), seq(mem), seq(cu), seq(mu)
    if count > 0:
        cus, mus = cus / count, mus / count
    return dict(cpu_total=cpu_total, mem_total=mem_total, cpu_usage=cus, mem_usage=mus, count=count,
                flags=flags.astype(np.uint8))


# --------------------------------------------------------------------------------------------------
# a11 — UI colour band   ui/src/app/shared/pipe/progress-color.pipe.ts:8-23
# --------------------------------------------------------------------------------------------------
BAND_NONE, BAND_GREEN, BAND_YELLOW, BAND_RED = 255, 0, 1, 2


def progress_band(value):
    """Restates progress-color.pipe.ts:10-20.  ``value`` is a percent.  JS ``if (value)`` is false for
    0, NaN, null/undefined -> null (BAND_NONE)."""
    if value is None or value == 0 or value != value:            # :10  falsy
        return BAND_NONE
    if value <= 60:                                               # :11
        return BAND_GREEN
    elif value > 60 and value <= 80:                              # :14
        return BAND_YELLOW
    elif value > 80:                                              # :17
        return BAND_RED
    return BAND_NONE


# --------------------------------------------------------------------------------------------------
# a4 — ClusterMonitor.list_pods inner loop   core/apps/kubeops_api/cluster_monitor.py:90-117
# --------------------------------------------------------------------------------------------------
def pod_rollup(pods):
    """Restates cluster_monitor.py:90-117 on the columnar view of a pod list.

    ``pods``: list of dicts {phase_ok: bool (phase in Running/Succeeded, :115),
    containers: None | list of (restart_count:int, ready:bool)}.
    Returns dict(restart=[..], warn=[flat per-container 0/1 in order, :101], restart_members=[pod idx,
    :113], error_members=[pod idx, :115]).
    """
    restart, warn, restart_members, error_members = [], [], [], []
    for i, p in enumerate(pods):
        restart_count = 0                                         # :93
        if p["containers"] is not None:                           # :96
            for rc, ready in p["containers"]:                     # :97
                restart_count = restart_count + rc                # :98
                warn.append(1 if ready == False else 0)           # :101  noqa: E712 (reference idiom)
        restart.append(restart_count)
        if restart_count > 0:                                     # :113
            restart_members.append(i)
        if not p["phase_ok"]:                                     # :115
            error_members.append(i)
    return dict(restart=restart, warn=warn, restart_members=restart_members,
                error_members=error_members)


# --------------------------------------------------------------------------------------------------
# a5 — quick_sort_pods / quick_sort_error_loki_container   cluster_monitor.py:450-479
# --------------------------------------------------------------------------------------------------
def quick_sort_desc(items, key):
    """Line-for-line restatement of cluster_monitor.py:450-463 (and :466-479 with key='error_count'),
    made iterative-safe only by raising the recursion limit at the call site.  First element is the
    pivot; ``<=`` go right in encounter order -> a STABLE DESCENDING sort.  Unlike the reference this
    does not mutate its input (the reference ``remove``s the pivot from the caller's list, :456)."""
    if len(items) < 2:                                            # :451
        return list(items)
    mid = items[0]                                                # :453
    left, right = [], []
    for item in items[1:]:                                        # :456-458
        if item[key] <= mid[key]:                                 # :459
            right.append(item)
        else:
            left.append(item)
    return quick_sort_desc(left, key) + [mid] + quick_sort_desc(right, key)   # :463


def stable_desc_order(keys):
    """Index order equivalent to :func:`quick_sort_desc` (property-tested): stable sort by -key."""
    return sorted(range(len(keys)), key=lambda i: -keys[i])


def stable_desc_order_np(keys):
    keys = np.asarray(keys, dtype=np.int64)
    return np.argsort(-keys, kind="stable")


# --------------------------------------------------------------------------------------------------
# a7 — available rate   core/apps/kubeops_api/prometheus_client.py:75-85
# --------------------------------------------------------------------------------------------------
def available_rate(job_up, job_len):
    """Restates prometheus_client.py:75-85 on per-job (#Ready, #targets) counts.
    Returns (per-job rates list, cluster rate)."""
    service_up = 0                                                # :76
    rates = []
    for up, ln in zip(job_up, job_len):
        rate = up / ln * 100 if ln > 0 else 0                     # :82
        if rate == 100:                                           # :83
            service_up = service_up + 1
        rates.append(rate)
    total = service_up / len(job_len) * 100 if len(job_len) > 0 else 0   # :85
    return rates, total


def health_day_mean(hour_rates):
    """Restates cluster_health_utils.py:30-33 (mean over all HOUR rows; raises on empty like :33)."""
    cluster_rate = 0
    for r in hour_rates:
        cluster_rate = cluster_rate + r
    return cluster_rate / len(hour_rates)


# --------------------------------------------------------------------------------------------------
# a6 — Loki error-line count   core/apps/kubeops_api/prometheus_client.py:140-146
# --------------------------------------------------------------------------------------------------
def loki_error_count(lines):
    """Restates prometheus_client.py:142-145: number of non-None lines containing 'level=error'."""
    error_count = 0
    for line in lines:
        if line is not None and "level=error" in line:
            error_count = error_count + 1
    return error_count


def loki_error_count_bytes(buf, line_offsets):
    """Columnar twin: ``buf`` bytes, ``line_offsets`` int64[n+1].  A match must lie inside one line."""
    needle = b"level=error"
    out = 0
    for i in range(len(line_offsets) - 1):
        if needle in bytes(buf[line_offsets[i]:line_offsets[i + 1]]):
            out += 1
    return out


# --------------------------------------------------------------------------------------------------
# a11 — fleet roll-up in the browser   ui/src/app/dashboard/dashboard.component.ts:117-153
# --------------------------------------------------------------------------------------------------
def fleet_rollup(clusters):
    """Line-by-line restatement of dashboard.component.ts:117-153 (TypeScript; JS numbers = f64).

    ``clusters``: list of dicts {pods, namespaces, deployments, containers, nodes (ints),
    cpu_total, mem_total, cpu_usage, mem_usage (floats), max_pods (None = undefined -> 110, :138-141; 0 stays 0)}
    and optionally ``ui_nodes``: cluster.nodes.length of the UI's own cluster list (:142), None when that list
    does not hold the cluster (the loop at :136 then finds nothing and no capacity is added); default: ``nodes``.
    """
    pod = ns = dep = cont = node = 0
    cpu_total = mem_total = cpu_usage = mem_usage = 0
    max_pod_count = 0
    container_usage = 0
    count = len(clusters)                                         # :117
    for d in clusters:                                            # :118
        pod = pod + d["pods"]                                     # :120
        ns = ns + d["namespaces"]                                 # :121
        dep = dep + d["deployments"]                              # :122
        cont = cont + d["containers"]                             # :123-125
        node = node + d["nodes"]                                  # :126
        cpu_total = cpu_total + d["cpu_total"]                    # :127
        mem_total = mem_total + d["mem_total"]                    # :128
        cpu_usage = cpu_usage + d["cpu_usage"]                    # :129
        mem_usage = mem_usage + d["mem_usage"]                    # :130
        if d["cpu_total"] == 0 and d["mem_total"] == 0:           # :131
            count -= 1
        ui_nodes = d.get("ui_nodes", d["nodes"])
        if ui_nodes is not None:                                  # :136-137 the UI's cluster list holds it
            max_pod = d.get("max_pods")                           # :138
            if max_pod is None:                                   # :139 (=== undefined; 0 is a value)
                max_pod = 110                                     # :140
            max_pod_count = max_pod_count + max_pod * ui_nodes    # :142-143
    if count > 0:                                                 # :147
        cpu_usage = cpu_usage / count * 100                       # :148
        mem_usage = mem_usage / count * 100                       # :149
    if max_pod_count > 0:                                         # :151
        container_usage = pod / max_pod_count * 100               # :152
    return dict(pod_count=pod, namespace_count=ns, deployment_count=dep, container_count=cont,
                node_count=node, cpu_total=cpu_total, mem_total=mem_total, cpu_usage=cpu_usage,
                mem_usage=mem_usage, max_pod_count=max_pod_count, container_usage=container_usage,
                count=count)


# --------------------------------------------------------------------------------------------------
# BUILD-DEFINED — range-window statistics over one series (no reference implementation)
# --------------------------------------------------------------------------------------------------
def nearest_rank(n, q):
    """1-based nearest rank: clamp(ceil(q*n), 1, n) with q*n evaluated in IEEE double."""
    r = int(math.ceil(q * n))
    return max(1, min(n, r))


def series_stats(values, qs=()):
    """Window statistics of one series in the reference's style (left-to-right float accumulation as
    cluster_monitor.py:170-173; ``if count > 0`` guard as :180).  NaN = Prometheus staleness marker,
    skipped.  min/max/last/quantiles are NaN when no valid sample exists.  Zero results are
    canonicalised to +0.0 (``x + 0.0``) so that -0.0/+0.0 ties are bit-defined.
    Quantile = nearest-rank order statistic ``sorted(valid)[ceil(q*n)-1]`` (an element of the data)."""
    total = 0.0
    count = 0
    mn = mx = last = NAN
    valid = []
    for v in values:
        v = float(v)
        if v != v:
            continue
        total = total + v
        count = count + 1
        if count == 1 or v < mn:
            mn = v
        if count == 1 or v > mx:
            mx = v
        last = v
        valid.append(v)
    mean = total / count if count > 0 else 0.0
    quant = []
    if qs is not None and len(qs):
        srt = sorted(valid)
        for q in qs:
            quant.append(srt[nearest_rank(count, q) - 1] + 0.0 if count > 0 else NAN)
    if count > 0:
        mn, mx = mn + 0.0, mx + 0.0
    return dict(sum=total, count=count, mean=mean, min=mn, max=mx, last=last, quantiles=quant)


def series_rollup_np(values, offsets, qs=()):
    """NumPy twin of :func:`series_stats` over a columnar batch (values f64[], offsets i64[n+1]).
    Bit-identical to the pure-Python form (cumsum is strictly sequential)."""
    values = np.asarray(values, np.float64)
    offsets = np.asarray(offsets, np.int64)
    n = len(offsets) - 1
    nq = len(qs)
    out = dict(sum=np.zeros(n), mean=np.zeros(n), min=np.full(n, np.nan), max=np.full(n, np.nan),
               last=np.full(n, np.nan), count=np.zeros(n, np.int64),
               quantiles=np.full((n, nq), np.nan))
    for i in range(n):
        v = values[offsets[i]:offsets[i + 1]]
        v = v[v == v]
        c = len(v)
        out["count"][i] = c
        if c == 0:
            continue
        s = float(np.cumsum(v)[-1])
        out["sum"][i] = s
        out["mean"][i] = s / c
        out["min"][i] = v.min() + 0.0
        out["max"][i] = v.max() + 0.0
        out["last"][i] = v[-1]
        if nq:
            srt = np.sort(v)
            for j, q in enumerate(qs):
                out["quantiles"][i, j] = srt[nearest_rank(c, q) - 1] + 0.0
    return out


# --------------------------------------------------------------------------------------------------
# BUILD-DEFINED — fleet-wide merge across shards
# --------------------------------------------------------------------------------------------------
def ordered_key(x):
    """Order-preserving u64 image of an f64 (the key the histogram merge refines on)."""
    (u,) = struct.unpack("<Q", struct.pack("<d", x))
    return (u ^ 0xFFFFFFFFFFFFFFFF) if (u >> 63) else (u | 0x8000000000000000)


def ordered_key_np(x):
    u = np.asarray(x, np.float64).view(np.uint64)
    neg = (u >> np.uint64(63)).astype(bool)
    return np.where(neg, ~u, u | np.uint64(0x8000000000000000))


def key_to_double_np(k):
    k = np.asarray(k, np.uint64)
    pos = (k >> np.uint64(63)).astype(bool)
    u = np.where(pos, k & np.uint64(0x7FFFFFFFFFFFFFFF), ~k)
    return u.view(np.float64)


def fleet_quantile(shards, q):
    """Nearest-rank quantile of the pooled valid values of all shards (zero canonicalised)."""
    pooled = sorted(float(v) for s in shards for v in s if v == v)
    if not pooled:
        return NAN
    return pooled[nearest_rank(len(pooled), q) - 1] + 0.0


def key_histogram_np(values, shift, prefix_mask, prefix, bits=16):
    """Histogram of ((key >> shift) & (2^bits-1)) over valid values whose (key & prefix_mask)==prefix.
    One round of the allgather-and-refine fleet quantile."""
    v = np.asarray(values, np.float64)
    v = v[v == v]
    v = np.where(v == 0, 0.0, v)                 # canonical +0.0
    k = ordered_key_np(v)
    sel = (k & np.uint64(prefix_mask)) == np.uint64(prefix)
    d = ((k[sel] >> np.uint64(shift)) & np.uint64((1 << bits) - 1)).astype(np.int64)
    return np.bincount(d, minlength=1 << bits).astype(np.int64)


def fleet_topk(shard_keys, shard_global_idx, k):
    """Fleet top-k of (key desc, global index asc) — the order quick_sort_pods gives on the
    concatenated fleet list (api.py:495-500) truncated to k."""
    allp = [(-int(key), int(g)) for keys, gi in zip(shard_keys, shard_global_idx)
            for key, g in zip(keys, gi)]
    allp.sort()
    return [g for _, g in allp[:k]]


# --------------------------------------------------------------------------------------------------
# BUILD-DEFINED (§8f rank 2) — rate()/increase() with counter-reset handling over a window
# --------------------------------------------------------------------------------------------------
def counter_increase(values):
    """Sum of positive deltas with reset handling: a sample lower than its predecessor is a counter
    reset and contributes its own value (Prometheus ``increase`` without extrapolation).  NaN skipped.
    Returns (increase, first_valid_index, last_valid_index) (indices -1 when < 2 valid samples)."""
    inc = 0.0
    prev = None
    first = last = -1
    for i, v in enumerate(values):
        v = float(v)
        if v != v:
            continue
        if prev is None:
            first = i
        else:
            inc = inc + (v - prev if v >= prev else v)
        prev = v
        last = i
    if first == last:
        return 0.0, -1, -1
    return inc, first, last


def counter_rate(values, step_seconds):
    """increase / elapsed seconds between first and last valid sample (no extrapolation)."""
    inc, first, last = counter_increase(values)
    if first < 0:
        return 0.0
    return inc / ((last - first) * step_seconds)


# ---------------------------------------------------------------------------------------------------------
# §8f rank 1: Prometheus response -> columnar.  Restates what the reference does with a response
# (core/apps/kubeops_api/prometheus_client.py:91-116: req.json(), status == 'success', result[i]['value'][1]
# is a STRING that the caller later float()s, cluster_monitor.py:171-182) for whole matrices / vectors.
def prom_columnar(body):
    """(values, offsets, timestamps, metrics) of a query / query_range response body; raises ValueError
    when status is not 'success' (the reference skips the assignment in that case)."""
    import json as _json
    doc = _json.loads(body)
    if doc.get("status") != "success":
        raise ValueError("status is not success: %s" % doc.get("error"))
    data = doc.get("data", {})
    if data.get("resultType") not in ("matrix", "vector"):
        raise ValueError("resultType %r carries no series" % data.get("resultType"))
    values, stamps, offsets, metrics = [], [], [0], []
    for s in data.get("result", []):
        pairs = s["values"] if "values" in s else ([s["value"]] if "value" in s else [])
        for t, v in pairs:
            stamps.append(float(t))
            values.append(float(v))          # float("NaN"), float("+Inf"), float("-Inf") as Prometheus prints them
        offsets.append(len(values))
        metrics.append(s.get("metric", {}))
    return values, offsets, stamps, metrics
