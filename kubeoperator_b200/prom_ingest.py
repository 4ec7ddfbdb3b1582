"""Prometheus HTTP API response -> the columnar layout the rollup takes.

The reference goes through ``requests.json()`` and reads one string per node and query
(``core/apps/kubeops_api/prometheus_client.py:91-116``: ``json['data']['result'][0]['value'][1]``).  A
``query_range`` over a fleet is a body of hundreds of megabytes; ``parse_response`` hands it to the native
parser of libkorollup (``ko_prom_open`` / ``ko_prom_fill``, host C++ on several threads) and returns

    values   float64[n_samples]   all samples of all series back to back (NaN = staleness marker)
    offsets  int64[n_series + 1]
    timestamps float64[n_samples] (optional)
    metrics  list of dicts        the label sets, decoded lazily from the byte ranges the parser reports

``values`` can be a caller-supplied (e.g. page-locked) buffer so that ``Rollup.series_rollup`` copies it at PCIe
speed.  Host code: no GPU is needed for this step.
"""
import ctypes
import json
import os

import numpy as np

from . import korollup


def parse_response(body, want_timestamps=False, want_metrics=True, values_out=None, n_threads=None):
    """``body``: bytes of an ``/api/v1/query`` or ``/api/v1/query_range`` response.  Raises KoRollupError with
    code KO_ESTATE when ``status`` is not ``"success"`` (the reference's guard) and KO_EINVAL on malformed
    input."""
    if isinstance(body, str):
        body = body.encode("utf-8")
    lib = korollup.load_library()
    doc = ctypes.c_void_p()
    rc = lib.ko_prom_open(body, len(body), ctypes.byref(doc))
    if rc != 0:
        raise korollup.KoRollupError(rc, lib.ko_last_error().decode())
    try:
        ns, nv = ctypes.c_int64(0), ctypes.c_int64(0)
        lib.ko_prom_dims(doc, ctypes.byref(ns), ctypes.byref(nv))
        n_series, n_samples = int(ns.value), int(nv.value)
        if values_out is None:
            values = np.empty(n_samples, np.float64)
        else:
            if values_out.dtype != np.float64 or values_out.size < n_samples or not values_out.flags.c_contiguous:
                raise ValueError("values_out must be a contiguous float64 array of at least n_samples")
            values = values_out[:n_samples]
        stamps = np.empty(n_samples, np.float64) if want_timestamps else None
        offsets = np.empty(n_series + 1, np.int64)
        mb = np.empty(n_series, np.int64) if want_metrics else None
        me = np.empty(n_series, np.int64) if want_metrics else None
        threads = n_threads if n_threads else min(16, os.cpu_count() or 1)
        rc = lib.ko_prom_fill(doc, values.ctypes.data, None if stamps is None else stamps.ctypes.data,
                              offsets.ctypes.data, None if mb is None else mb.ctypes.data,
                              None if me is None else me.ctypes.data, threads)
        if rc != 0:
            raise korollup.KoRollupError(rc, lib.ko_last_error().decode())
    finally:
        lib.ko_prom_close(doc)
    out = dict(values=values, offsets=offsets)
    if want_timestamps:
        out["timestamps"] = stamps
    if want_metrics:
        out["metrics"] = [json.loads(body[b:e]) if e > b else {} for b, e in zip(mb.tolist(), me.tolist())]
    return out
