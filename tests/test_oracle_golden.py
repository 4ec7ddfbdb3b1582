"""The oracle (oracle/rollup_oracle.py) against the reference's own outputs (tests/golden/reference_*.json,
made by oracle/gen_golden.py from the unmodified reference functions) and against the hand-computed /
build-defined fixtures.  CPU only."""
import json
import math
import os

import numpy as np
import pytest

from oracle import rollup_oracle as ro


def _load(golden_dir, name):
    with open(os.path.join(golden_dir, name)) as f:
        return json.load(f)


def _same(a, b):
    if isinstance(a, float) and isinstance(b, float) and a != a and b != b:
        return True
    return a == b


def test_set_cluster_data_matches_reference(golden_dir):
    for case in _load(golden_dir, "reference_set_cluster_data.json"):
        exp = case["expect"]
        got = ro.node_rollup(case["nodes"])
        # bit-exact: the restatement performs the same left-to-right float operations
        for k in ("cpu_usage", "cpu_total", "mem_usage", "mem_total"):
            assert _same(got[k], exp[k]), (k, got[k], exp[k])
        names = [n["name"] for n in case["nodes"]]
        assert [names[i] for i, f in enumerate(got["flags"]) if f] == exp["flagged"]
        # numpy twin
        if case["nodes"]:
            tw = ro.node_rollup_np([n["cpu"] for n in case["nodes"]], [n["mem"] for n in case["nodes"]],
                                   [n["cpu_usage"] for n in case["nodes"]],
                                   [n["mem_usage"] for n in case["nodes"]])
            for k in ("cpu_usage", "cpu_total", "mem_usage", "mem_total", "count"):
                assert _same(float(tw[k]), float(got[k]))
            assert list(tw["flags"]) == got["flags"]
        pods = [dict(phase_ok=p["phase"] in ("Running", "Succeeded"),
                     containers=None if p["containers"] is None else
                     [(c["restart_count"], c["ready"]) for c in p["containers"]]) for p in case["pods"]]
        pr = ro.pod_rollup(pods)
        assert pr["restart"] == exp["pod_restart"]
        pn = [p["name"] for p in case["pods"]]
        keys = [pr["restart"][i] for i in pr["restart_members"]]
        order = ro.stable_desc_order(keys)
        assert [pn[pr["restart_members"][i]] for i in order] == exp["restart_pods"]
        keys = [pr["restart"][i] for i in pr["error_members"]]
        order = ro.stable_desc_order(keys)
        assert [pn[pr["error_members"][i]] for i in order] == exp["error_pods"]
        flat = [[p["name"], c["name"]] for p in case["pods"] if p["containers"] for c in p["containers"]]
        assert [flat[i] for i, w in enumerate(pr["warn"]) if w] == exp["warn_containers"]
        assert exp["keys"] == sorted(["id", "name", "pods", "nodes", "token", "namespaces", "deployments",
                                      "cpu_usage", "cpu_total", "mem_usage", "mem_total", "restart_pods",
                                      "warn_containers", "error_loki_containers", "error_pods"])


def test_quick_sorts_match_reference(golden_dir):
    import sys
    for case in _load(golden_dir, "reference_sorts.json"):
        keys = case["keys"]
        assert ro.stable_desc_order(keys) == case["order_pods"]
        assert ro.stable_desc_order(keys) == case["order_loki"]
        assert list(ro.stable_desc_order_np(keys)) == case["order_pods"]
        old = sys.getrecursionlimit()
        sys.setrecursionlimit(20000)
        try:
            items = [dict(name=i, restart_count=k) for i, k in enumerate(keys)]
            assert [p["name"] for p in ro.quick_sort_desc(items, "restart_count")] == case["order_pods"]
        finally:
            sys.setrecursionlimit(old)


def test_available_rate_matches_reference(golden_dir):
    g = _load(golden_dir, "reference_available_rate.json")
    for case in g["cases"]:
        rates, total = ro.available_rate(case["job_up"], case["job_len"])
        assert rates == case["rates"] and total == case["rate"]
    res = g["full"]["result"]
    up = [sum(1 for t in d["data"] if t["value"] == "Ready") for d in res["data"]]
    ln = [len(d["data"]) for d in res["data"]]
    rates, total = ro.available_rate(up, ln)
    assert rates == [d["rate"] for d in res["data"]] and total == res["rate"]
    assert g["failed"]["success"] is False and g["failed"]["rate"] == 0


def test_loki_count_matches_reference(golden_dir):
    g = _load(golden_dir, "reference_loki.json")
    exp = {c["name"]: c["error_count"] for c in g["result"]}
    for nm in g["names"]:
        lines = [e.get("line") for st in g["streams"][nm] for e in st["entries"]]
        cnt = ro.loki_error_count(lines)
        assert cnt == exp.get(nm, 0)
        # columnar twin
        kept = [l for l in lines if l is not None]
        buf = "".join(kept).encode()
        offs = np.cumsum([0] + [len(l.encode()) for l in kept])
        assert ro.loki_error_count_bytes(np.frombuffer(buf, np.uint8), offs) == cnt


def test_fleet_rollup_handcomputed(golden_dir):
    g = _load(golden_dir, "handcomputed_fleet.json")
    got = ro.fleet_rollup(g["clusters"])
    for k, v in g["expect"].items():
        assert got[k] == v, k
    for value, band in g["bands"]:
        assert ro.progress_band(value) == band
    assert ro.progress_band(float("nan")) == ro.BAND_NONE


def _dec(v):
    if v is None:
        return float("nan")
    if isinstance(v, str):
        return float(v)
    return float(v)


def test_fleet_rollup_more_handcomputed_vectors(golden_dir):
    """count == 0, every cluster zero, MAX_PODS absent / present / 0, clusters the UI list does not hold, a UI node
    count that differs from the blob's: each worked out by hand from dashboard.component.ts:117-153."""
    g = _load(golden_dir, "handcomputed_fleet.json")
    assert len(g["more"]) >= 8
    for case in g["more"]:
        got = ro.fleet_rollup(case["clusters"])
        for k, v in case["expect"].items():
            assert math.isclose(got[k], v, rel_tol=1e-15, abs_tol=0), (case["name"], k, got[k], v)


def test_series_builddefined_vectors(golden_dir):
    g = _load(golden_dir, "builddefined_series.json")
    qs = g["qs"]
    vals, offs = [], [0]
    for s in g["series"]:
        v = [_dec(x) for x in s["values"]]
        st = ro.series_stats(v, qs)
        for k, x in s["expect"].items():
            assert _same(float(st[k]), _dec(x)), k
        for a, b in zip(st["quantiles"], s["quantiles"]):
            assert _same(a, _dec(b))
        assert _same(float(ro.counter_increase(v)[0]), float(s["increase"]))
        vals += v
        offs.append(len(vals))
    tw = ro.series_rollup_np(np.array(vals), np.array(offs), qs)
    for i, s in enumerate(g["series"]):
        for k in ("sum", "mean", "min", "max", "last"):
            assert _same(float(tw[k][i]), _dec(s["expect"][k])), (i, k)
            if float(tw[k][i]) == 0:
                assert math.copysign(1.0, float(tw[k][i])) == math.copysign(1.0, _dec(s["expect"][k]))
        assert int(tw["count"][i]) == s["expect"]["count"]
        for j in range(len(qs)):
            assert _same(float(tw["quantiles"][i, j]), _dec(s["quantiles"][j]))


def test_fleet_quantile_histogram_refine_equals_pooled_sort():
    rng = np.random.default_rng(7)
    shards = [rng.beta(2, 5, 1000), np.array([0.0, -0.0, np.nan, 5.0]), rng.normal(0, 3, 777)]
    for q in (0.01, 0.5, 0.95, 0.99, 1.0):
        n = sum(int((s == s).sum()) for s in shards)
        r = ro.nearest_rank(n, q)
        prefix, mask = 0, 0
        for shift in (48, 32, 16, 0):
            h = sum(ro.key_histogram_np(s, shift, mask, prefix) for s in shards)
            c = np.cumsum(h)
            d = int(np.searchsorted(c, r, side="left"))
            r -= int(c[d - 1]) if d else 0
            prefix |= d << shift
            mask |= 0xFFFF << shift
        got = float(ro.key_to_double_np(np.array([prefix], np.uint64))[0]) + 0.0
        assert got == ro.fleet_quantile(shards, q)


def test_stable_order_equals_quicksort_property():
    hyp = pytest.importorskip("hypothesis")
    from hypothesis import given, settings, strategies as st

    @settings(max_examples=200, deadline=None)
    @given(st.lists(st.integers(min_value=-3, max_value=6), max_size=60))
    def prop(keys):
        items = [dict(i=i, restart_count=k) for i, k in enumerate(keys)]
        assert [p["i"] for p in ro.quick_sort_desc(items, "restart_count")] == ro.stable_desc_order(keys)

    prop()
