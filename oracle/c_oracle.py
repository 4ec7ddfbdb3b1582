"""TEST INFRASTRUCTURE ONLY — ctypes loader for the plain-C oracle (oracle/korollup_oracle.c).

Allowed importers: tests/, __graft_entry__.smoke(), bench.py (cpu_baseline / --impl reference)."""
import ctypes
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "_build", "libkooracle.so")
_lib = None


def build(force=False):
    src = os.path.join(HERE, "korollup_oracle.c")
    if force or not os.path.exists(LIB_PATH) or os.path.getmtime(LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(["make", "-s", "-C", HERE])
    return LIB_PATH


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            build()
        _lib = ctypes.CDLL(LIB_PATH)
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(ctypes.c_void_p)


def series_rollup(values, offsets, qs=(), nthreads=1):
    values = np.ascontiguousarray(values, np.float64)
    offsets = np.ascontiguousarray(offsets, np.int64)
    n = len(offsets) - 1
    q = np.ascontiguousarray(qs, np.float64)
    out = dict(sum=np.empty(n), mean=np.empty(n), min=np.empty(n), max=np.empty(n), last=np.empty(n),
               count=np.empty(n, np.int64), quantiles=np.empty((n, len(q))))
    rc = lib().koo_series_rollup(_p(values), _p(offsets), ctypes.c_int64(n), _p(q), ctypes.c_int(len(q)),
                                 _p(out["sum"]), _p(out["mean"]), _p(out["min"]), _p(out["max"]),
                                 _p(out["last"]), _p(out["count"]), _p(out["quantiles"]),
                                 ctypes.c_int(nthreads))
    assert rc == 0
    return out


def cluster_rollup(cpu, mem, cpu_usage, mem_usage, cluster_of_node, n_clusters, threshold=80.0):
    a = [np.ascontiguousarray(x, np.float64) for x in (cpu, mem, cpu_usage, mem_usage)]
    cl = np.ascontiguousarray(cluster_of_node, np.int32)
    n = len(cl)
    out = dict(cpu_total=np.empty(n_clusters), mem_total=np.empty(n_clusters),
               cpu_usage=np.empty(n_clusters), mem_usage=np.empty(n_clusters),
               count=np.empty(n_clusters, np.int32), flags=np.empty(n, np.uint8))
    rc = lib().koo_cluster_rollup(_p(a[0]), _p(a[1]), _p(a[2]), _p(a[3]), _p(cl), ctypes.c_int64(n),
                                  ctypes.c_int32(n_clusters), ctypes.c_double(threshold),
                                  _p(out["cpu_total"]), _p(out["mem_total"]), _p(out["cpu_usage"]),
                                  _p(out["mem_usage"]), _p(out["count"]), _p(out["flags"]))
    assert rc == 0
    return out


def pod_rollup(container_restart, container_ready, pod_offsets, pod_phase_ok):
    cr = np.ascontiguousarray(container_restart, np.int32)
    rd = np.ascontiguousarray(container_ready, np.uint8)
    po = np.ascontiguousarray(pod_offsets, np.int64)
    ok = np.ascontiguousarray(pod_phase_ok, np.uint8)
    n = len(po) - 1
    restart = np.empty(n, np.int32)
    warn = np.empty(len(cr), np.uint8)
    o_r, o_e = np.empty(n, np.int64), np.empty(n, np.int64)
    nr, ne = ctypes.c_int64(0), ctypes.c_int64(0)
    rc = lib().koo_pod_rollup(_p(cr), _p(rd), _p(po), _p(ok), ctypes.c_int64(n), _p(restart), _p(warn),
                              _p(o_r), ctypes.byref(nr), _p(o_e), ctypes.byref(ne))
    assert rc == 0
    return dict(restart=restart, warn=warn, order_restart=o_r[:nr.value].copy(),
                order_error=o_e[:ne.value].copy())


def loki_count(buf, line_offsets, group_offsets):
    buf = np.ascontiguousarray(buf, np.uint8)
    lo = np.ascontiguousarray(line_offsets, np.int64)
    go = np.ascontiguousarray(group_offsets, np.int64)
    out = np.empty(len(go) - 1, np.int64)
    rc = lib().koo_loki_count(_p(buf), _p(lo), _p(go), ctypes.c_int64(len(go) - 1), _p(out))
    assert rc == 0
    return out


def counter_increase(values, offsets, step_seconds=15.0):
    values = np.ascontiguousarray(values, np.float64)
    offsets = np.ascontiguousarray(offsets, np.int64)
    n = len(offsets) - 1
    inc, rate = np.empty(n), np.empty(n)
    rc = lib().koo_counter_increase(_p(values), _p(offsets), ctypes.c_int64(n),
                                    ctypes.c_double(step_seconds), _p(inc), _p(rate))
    assert rc == 0
    return inc, rate


def available_rate(job_up, job_len):
    up = np.ascontiguousarray(job_up, np.int32)
    ln = np.ascontiguousarray(job_len, np.int32)
    rates = np.empty(len(up))
    total = ctypes.c_double(0)
    rc = lib().koo_available_rate(_p(up), _p(ln), ctypes.c_int32(len(up)), _p(rates), ctypes.byref(total))
    assert rc == 0
    return rates, total.value
