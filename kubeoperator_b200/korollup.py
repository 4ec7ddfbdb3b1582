"""ctypes shim over libkorollup.so (C ABI: include/korollup.h).

The shim adds nothing but argument marshalling: NumPy arrays in, NumPy arrays out, non-zero return codes
raised as :class:`KoRollupError` (a RuntimeError, which the reference's own handlers catch and log:
core/apps/kubeops_api/tasks.py:52-53, core/apps/kubeops_api/api.py:450-452).  There is no CPU fallback:
if the library is missing, or no B200 is present, calls fail loudly.

CUDA is initialised lazily by ``Rollup()`` (ko_init), never at import, so the module can be imported
before a Celery prefork (core/kubeops.py:154-160).
"""
import ctypes
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_ENV = "KOROLLUP_LIB"          # config key mirroring the reference's env/config.yml style
DEFAULT_LIB = os.path.join(HERE, "libkorollup.so")
MAX_QUANTILES = 4
HIST_BINS = 65536
BAND_NONE, BAND_GREEN, BAND_YELLOW, BAND_RED = 255, 0, 1, 2

_c_i64, _c_i32, _c_int, _c_dbl, _c_vp = (ctypes.c_int64, ctypes.c_int32, ctypes.c_int, ctypes.c_double,
                                         ctypes.c_void_p)

# name -> argtypes (restype is int unless noted); must list EVERY symbol include/korollup.h declares
ABI = {
    "ko_version": [],
    "ko_last_error": [],
    "ko_init": [ctypes.POINTER(_c_int), _c_int, ctypes.POINTER(_c_vp)],
    "ko_destroy": [_c_vp],
    "ko_get_counters": [_c_vp, _c_vp],
    "ko_last_series_kernel_ms": [_c_vp, _c_vp],
    "ko_last_series_pipe_ms": [_c_vp, _c_vp],
    "ko_pin": [_c_vp, _c_i64],
    "ko_unpin": [_c_vp],
    "ko_cluster_rollup": [_c_vp] + [_c_vp] * 5 + [_c_i64, _c_i32, _c_dbl] + [_c_vp] * 8,
    "ko_pod_rollup": [_c_vp] + [_c_vp] * 4 + [_c_i64] + [_c_vp] * 6,
    "ko_stable_desc_order": [_c_vp, _c_vp, _c_i64, _c_vp],
    "ko_series_rollup": [_c_vp, _c_vp, _c_vp, _c_i64, _c_vp, _c_int] + [_c_vp] * 7,
    "ko_series_rollup_dev": [_c_vp, _c_vp, _c_vp, _c_i64, _c_vp, _c_int] + [_c_vp] * 7 + [_c_vp],
    "ko_series_status": [_c_vp, _c_vp],
    "ko_series_rollup_static": [_c_int, _c_vp, _c_vp, _c_i64, _c_vp, _c_int] + [_c_vp] * 7 + [_c_vp],
    "ko_available_rate": [_c_vp, _c_vp, _c_vp, _c_vp, _c_i64, _c_vp, _c_vp],
    "ko_fleet_rollup": [_c_vp] + [_c_vp] * 6 + [_c_i64, _c_vp],
    "ko_key_histogram": [_c_vp, _c_vp, _c_i64, _c_int, ctypes.c_uint64, ctypes.c_uint64, _c_vp],
    "ko_key_histogram_dev": [_c_vp, _c_vp, _c_i64, _c_int, ctypes.c_uint64, ctypes.c_uint64, _c_vp, _c_vp],
    "ko_key_histogram_state_dev": [_c_vp, _c_vp, _c_i64, _c_int, _c_vp, _c_vp, _c_vp],
    "ko_key_select_dev": [_c_vp, _c_vp, _c_int, _c_vp, _c_vp],
    "ko_loki_count": [_c_vp, _c_vp, _c_vp, _c_i64, _c_vp, _c_i64, _c_vp],
    "ko_counter_increase": [_c_vp, _c_vp, _c_vp, _c_i64, _c_dbl, _c_vp, _c_vp],
    "ko_prom_open": [ctypes.c_char_p, _c_i64, ctypes.POINTER(_c_vp)],
    "ko_prom_dims": [_c_vp, ctypes.POINTER(_c_i64), ctypes.POINTER(_c_i64)],
    "ko_prom_fill": [_c_vp, _c_vp, _c_vp, _c_vp, _c_vp, _c_vp, _c_int],
    "ko_prom_close": [_c_vp],
}


class KoRollupError(RuntimeError):
    def __init__(self, code, message):
        super().__init__("libkorollup error %d: %s" % (code, message))
        self.code = code


class _Counters(ctypes.Structure):
    _fields_ = [("kernel_launches", _c_i64), ("h2d_bytes", _c_i64), ("d2h_bytes", _c_i64),
                ("series_fallback", _c_i64)]


_lib = None


def lib_path():
    return os.environ.get(LIB_ENV, DEFAULT_LIB)


def load_library():
    """dlopen the library and bind the ABI.  Does NOT touch CUDA."""
    global _lib
    if _lib is not None:
        return _lib
    path = lib_path()
    if not os.path.exists(path):
        raise KoRollupError(-4, "%s not found: build it with `python -m kubeoperator_b200.build` "
                                "(there is no CPU fallback)" % path)
    lib = ctypes.CDLL(path)
    for name, argtypes in ABI.items():
        fn = getattr(lib, name)
        fn.argtypes = argtypes
        fn.restype = _c_int
    lib.ko_last_error.restype = ctypes.c_char_p
    lib.ko_destroy.restype = None
    lib.ko_prom_close.restype = None
    _lib = lib
    return lib


def _ptr(a):
    return None if a is None else a.ctypes.data


def _arr(a, dtype):
    return np.ascontiguousarray(a, dtype=dtype)


STATIC_MAX_SERIES = 65536


def series_rollup_static(device, d_values, d_offsets, n_series, qs, d_sum=0, d_mean=0, d_min=0, d_max=0, d_last=0,
                         d_count=0, d_quant=0, stream=0):
    """``ko_series_rollup_static``: the device-pointer series rollup for callers that must not allocate GPU
    memory (no ``Rollup`` / ko_ctx needed, nothing is cudaMalloc'ed; scratch is a fixed 6.6 MiB array inside the
    library).  At most ``STATIC_MAX_SERIES`` series per call (more raise ``KoRollupError``); calls on one device
    must not overlap on the GPU.  Arguments as ``Rollup.series_rollup_dev``."""
    lib = load_library()
    q = _arr(qs, np.float64)
    rc = lib.ko_series_rollup_static(int(device), d_values, d_offsets, int(n_series), _ptr(q) if len(q) else None,
                                     len(q), d_sum or None, d_mean or None, d_min or None, d_max or None,
                                     d_last or None, d_count or None, d_quant or None, stream or None)
    if rc != 0:
        raise KoRollupError(rc, lib.ko_last_error().decode())


class Rollup:
    """One ``ko_ctx`` (one process, one GPU).  Create it in the worker process, after the fork."""

    def __init__(self, device=0):
        self._lib = load_library()
        self._ctx = _c_vp()
        dev = (_c_int * 1)(device)
        rc = self._lib.ko_init(dev, 1, ctypes.byref(self._ctx))
        if rc != 0:
            raise KoRollupError(rc, self._lib.ko_last_error().decode())
        self.device = device

    # ------------------------------------------------------------------ plumbing
    def close(self):
        if getattr(self, "_ctx", None) is not None and self._ctx.value:
            self._lib.ko_destroy(self._ctx)
            self._ctx = _c_vp()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def _check(self, rc):
        if rc != 0:
            raise KoRollupError(rc, self._lib.ko_last_error().decode())

    def counters(self):
        c = _Counters()
        self._check(self._lib.ko_get_counters(self._ctx, ctypes.byref(c)))
        return dict(kernel_launches=c.kernel_launches, h2d_bytes=c.h2d_bytes, d2h_bytes=c.d2h_bytes,
                    series_fallback=c.series_fallback)

    def series_status(self):
        """Device-side outcome of the last ``series_rollup_dev`` (waits for it): long series, series repeated with
        wide intervals, series that took the slow exact path, device error flag."""
        out = (ctypes.c_int32 * 4)()
        self._check(self._lib.ko_series_status(self._ctx, out))
        return dict(long_series=out[0], second_attempt=out[1], slow_path=out[2], error_flag=out[3])

    def last_series_kernel_ms(self):
        ms = ctypes.c_float(0)
        self._check(self._lib.ko_last_series_kernel_ms(self._ctx, ctypes.byref(ms)))
        return float(ms.value)

    def last_series_pipe_ms(self):
        """Device time of the dominant kernel alone (first series_flow launch of the last series rollup)."""
        ms = ctypes.c_float(0)
        self._check(self._lib.ko_last_series_pipe_ms(self._ctx, ctypes.byref(ms)))
        return float(ms.value)

    def pin(self, array):
        self._check(self._lib.ko_pin(array.ctypes.data, array.nbytes))

    def unpin(self, array):
        self._check(self._lib.ko_unpin(array.ctypes.data))

    # ------------------------------------------------------------------ a2
    def cluster_rollup(self, cpu, mem, cpu_usage, mem_usage, cluster_of_node, n_clusters, threshold=80.0,
                       bands=False):
        """Per-cluster node rollup — cluster_monitor.py:163-182 for many clusters (see korollup.h)."""
        cpu, mem = _arr(cpu, np.float64), _arr(mem, np.float64)
        cu, mu = _arr(cpu_usage, np.float64), _arr(mem_usage, np.float64)
        cl = _arr(cluster_of_node, np.int32)
        n, k = len(cl), int(n_clusters)
        if not (len(cpu) == len(mem) == len(cu) == len(mu) == n):
            raise ValueError("node columns must have equal length")
        out = dict(cpu_total=np.zeros(k), mem_total=np.zeros(k), cpu_usage=np.zeros(k),
                   mem_usage=np.zeros(k), count=np.zeros(k, np.int32), flags=np.zeros(n, np.uint8))
        bc = np.zeros(n, np.uint8) if bands else None
        bm = np.zeros(n, np.uint8) if bands else None
        self._check(self._lib.ko_cluster_rollup(
            self._ctx, _ptr(cpu), _ptr(mem), _ptr(cu), _ptr(mu), _ptr(cl), n, k, float(threshold),
            _ptr(out["cpu_total"]), _ptr(out["mem_total"]), _ptr(out["cpu_usage"]), _ptr(out["mem_usage"]),
            _ptr(out["count"]), _ptr(out["flags"]), _ptr(bc), _ptr(bm)))
        if bands:
            out["band_cpu"], out["band_mem"] = bc, bm
        return out

    # ------------------------------------------------------------------ a4 + a5
    def pod_rollup(self, container_restart, container_ready, pod_offsets, pod_phase_ok):
        """Pod rollup + the two sorted member lists — cluster_monitor.py:93-117 and :183-184."""
        cr, rd = _arr(container_restart, np.int32), _arr(container_ready, np.uint8)
        po, ok = _arr(pod_offsets, np.int64), _arr(pod_phase_ok, np.uint8)
        n = len(po) - 1
        if n < 0 or len(ok) != n or len(cr) != len(rd) or (n >= 0 and len(cr) != (po[-1] if len(po) else 0)):
            raise ValueError("inconsistent pod columns")
        restart, warn = np.zeros(n, np.int32), np.zeros(len(cr), np.uint8)
        o_r, o_e = np.zeros(max(n, 1), np.int64), np.zeros(max(n, 1), np.int64)
        nr, ne = _c_i64(0), _c_i64(0)
        self._check(self._lib.ko_pod_rollup(self._ctx, _ptr(cr), _ptr(rd), _ptr(po), _ptr(ok), n, _ptr(restart),
                                            _ptr(warn), _ptr(o_r), ctypes.addressof(nr), _ptr(o_e),
                                            ctypes.addressof(ne)))
        return dict(restart=restart, warn=warn, order_restart=o_r[:nr.value].copy(),
                    order_error=o_e[:ne.value].copy())

    def stable_desc_order(self, keys):
        """Index order of quick_sort_pods / quick_sort_error_loki_container — cluster_monitor.py:450-479."""
        keys = _arr(keys, np.int32)
        out = np.zeros(len(keys), np.int64)
        self._check(self._lib.ko_stable_desc_order(self._ctx, _ptr(keys), len(keys), _ptr(out)))
        return out

    # ------------------------------------------------------------------ window statistics
    def series_rollup(self, values, offsets, qs=(), want=("sum", "mean", "min", "max", "last", "count")):
        """Window statistics + nearest-rank quantiles over columnar series (build-defined; korollup.h)."""
        values, offsets = _arr(values, np.float64), _arr(offsets, np.int64)
        q = _arr(qs, np.float64)
        n = len(offsets) - 1
        if n < 0:
            raise ValueError("offsets must have n_series+1 entries")
        out = {}
        for k in ("sum", "mean", "min", "max", "last"):
            out[k] = np.zeros(n) if k in want else None
        out["count"] = np.zeros(n, np.int64) if "count" in want else None
        quant = np.zeros((n, len(q))) if len(q) else None
        self._check(self._lib.ko_series_rollup(
            self._ctx, _ptr(values), _ptr(offsets), n, _ptr(q) if len(q) else None, len(q), _ptr(out["sum"]),
            _ptr(out["mean"]), _ptr(out["min"]), _ptr(out["max"]), _ptr(out["last"]), _ptr(out["count"]),
            _ptr(quant)))
        out = {k: v for k, v in out.items() if v is not None}
        if quant is not None:
            out["quantiles"] = quant
        return out

    def series_rollup_dev(self, d_values, d_offsets, n_series, qs, d_sum=0, d_mean=0, d_min=0, d_max=0,
                          d_last=0, d_count=0, d_quant=0, stream=0):
        """Device-pointer form: integers are raw device addresses (e.g. torch ``tensor.data_ptr()``) and
        ``stream`` a cudaStream_t handle (``torch.cuda.current_stream().cuda_stream``).  Asynchronous."""
        q = _arr(qs, np.float64)
        self._check(self._lib.ko_series_rollup_dev(
            self._ctx, d_values, d_offsets, int(n_series), _ptr(q) if len(q) else None, len(q), d_sum or None,
            d_mean or None, d_min or None, d_max or None, d_last or None, d_count or None, d_quant or None,
            stream or None))

    # ------------------------------------------------------------------ a7, a11
    def available_rate(self, job_up, job_len, job_offsets):
        """calculate_available_rate for many clusters — prometheus_client.py:75-85."""
        up, ln, off = _arr(job_up, np.int32), _arr(job_len, np.int32), _arr(job_offsets, np.int64)
        k = len(off) - 1
        jr, cr = np.zeros(len(up)), np.zeros(max(k, 0))
        self._check(self._lib.ko_available_rate(self._ctx, _ptr(up), _ptr(ln), _ptr(off), k, _ptr(jr), _ptr(cr)))
        return jr, cr

    FLEET_KEYS = ("pod_count", "namespace_count", "deployment_count", "container_count", "node_count",
                  "cpu_total", "mem_total", "cpu_usage", "mem_usage", "max_pod_count", "container_usage",
                  "count")

    def fleet_rollup(self, counts, cpu_total, mem_total, cpu_usage, mem_usage, max_pods):
        """The browser's fleet roll-up — dashboard.component.ts:117-153.  ``counts`` rows: pods, namespaces,
        deployments, containers, nodes [, ui_nodes]: ``ui_nodes`` = cluster.nodes.length of the UI's own cluster list,
        -1 = the list does not hold the cluster (no pod capacity); without the sixth column the blob's node count is
        used.  ``max_pods`` entries: None or a negative value = MAX_PODS undefined (-> 110); 0 is a configured zero."""
        counts = np.asarray(counts, np.int64)
        counts = counts.reshape(-1, counts.shape[-1] if counts.ndim == 2 else 5)
        if counts.shape[1] == 5:
            counts = np.concatenate([counts, counts[:, 4:5]], axis=1)
        counts = np.ascontiguousarray(counts.reshape(-1, 6))
        k = counts.shape[0]
        cols = [_arr(x, np.float64) for x in (cpu_total, mem_total, cpu_usage, mem_usage)]
        mp = _arr([-1 if m is None else int(m) for m in max_pods], np.int32)
        out = np.zeros(12)
        self._check(self._lib.ko_fleet_rollup(self._ctx, _ptr(counts), _ptr(cols[0]), _ptr(cols[1]),
                                              _ptr(cols[2]), _ptr(cols[3]), _ptr(mp), k, _ptr(out)))
        res = dict(zip(self.FLEET_KEYS, out.tolist()))
        for key in ("pod_count", "namespace_count", "deployment_count", "container_count", "node_count",
                    "max_pod_count", "count"):
            res[key] = int(res[key])
        return res

    # ------------------------------------------------------------------ e
    def key_histogram(self, values, shift, prefix_mask, prefix):
        values = _arr(values, np.float64)
        out = np.zeros(HIST_BINS, np.int64)
        self._check(self._lib.ko_key_histogram(self._ctx, _ptr(values), len(values), int(shift),
                                               int(prefix_mask), int(prefix), _ptr(out)))
        return out

    def key_histogram_dev(self, d_values, n, shift, prefix_mask, prefix, d_hist, stream=0):
        self._check(self._lib.ko_key_histogram_dev(self._ctx, d_values, int(n), int(shift), int(prefix_mask),
                                                   int(prefix), d_hist, stream or None))

    # ------------------------------------------------------------------ next rows (§8f)
    def key_histogram_state_dev(self, d_values, n, shift, d_state, d_hist, stream=0):
        self._check(self._lib.ko_key_histogram_state_dev(self._ctx, d_values, int(n), int(shift), d_state, d_hist,
                                                         stream or None))

    def key_select_dev(self, d_hist, shift, d_state, stream=0):
        self._check(self._lib.ko_key_select_dev(self._ctx, d_hist, int(shift), d_state, stream or None))

    def loki_count(self, buf, line_offsets, group_offsets):
        """Lines containing b'level=error' per container — prometheus_client.py:140-146."""
        buf = _arr(buf, np.uint8)
        lo, go = _arr(line_offsets, np.int64), _arr(group_offsets, np.int64)
        out = np.zeros(len(go) - 1, np.int64)
        self._check(self._lib.ko_loki_count(self._ctx, _ptr(buf), _ptr(lo), len(lo) - 1, _ptr(go), len(go) - 1,
                                            _ptr(out)))
        return out

    def counter_increase(self, values, offsets, step_seconds=15.0):
        values, offsets = _arr(values, np.float64), _arr(offsets, np.int64)
        n = len(offsets) - 1
        inc, rate = np.zeros(n), np.zeros(n)
        self._check(self._lib.ko_counter_increase(self._ctx, _ptr(values), _ptr(offsets), n, float(step_seconds),
                                                  _ptr(inc), _ptr(rate)))
        return inc, rate
