"""GPU parity of the reference-defined rollup (node / pod / sort / availability / Loki / fleet) through the
C ABI and through the host-side mirror of the reference interface, against the reference's own outputs
(tests/golden/reference_*.json) and the oracle.  Integer / flag / order results bit-exact; the four float
fields cpu_usage, cpu_total, mem_usage, mem_total within 1e-12 relative (north_star)."""
import json
import os
import types

import numpy as np
import pytest

from oracle import c_oracle, rollup_oracle as ro
from kubeoperator_b200 import synth

pytestmark = pytest.mark.gpu
RTOL = 1e-12


@pytest.fixture(scope="module")
def rollup():
    from kubeoperator_b200 import Rollup
    r = Rollup(0)
    yield r
    r.close()


def _load(golden_dir, name):
    with open(os.path.join(golden_dir, name)) as f:
        return json.load(f)


def _close(a, b):
    """The float fields of the node and fleet roll-ups: EQUAL to the reference's (the kernels add in the reference's
    order, node by node / cluster by cluster), NaN matching NaN."""
    a, b = float(a), float(b)
    if a != a and b != b:
        return True
    return a == b


def _pod_objects(pods):
    items = []
    for p in pods:
        cs = None
        if p["containers"] is not None:
            cs = [types.SimpleNamespace(name=c["name"], ready=c["ready"], restart_count=c["restart_count"])
                  for c in p["containers"]]
        items.append(types.SimpleNamespace(
            status=types.SimpleNamespace(container_statuses=cs, host_ip=p.get("host_ip"), pod_ip=p.get("pod_ip"),
                                         phase=p["phase"]),
            metadata=types.SimpleNamespace(name=p["name"], namespace=p.get("namespace", "default"))))
    return items


def test_set_cluster_data_blob_matches_reference(rollup, golden_dir):
    """Same Redis blob as the unmodified reference (cluster_monitor.py:157-192) on the golden cases."""
    from kubeoperator_b200.cluster_monitor import ClusterRollup
    for case in _load(golden_dir, "reference_set_cluster_data.json"):
        cr = ClusterRollup("c", cluster_id="00000000-0000-0000-0000-000000000001", token="tok", rollup=rollup)
        pods = cr.list_pods(_pod_objects(case["pods"]), host_name_of=lambda ip: "host-" + str(ip))
        blob = cr.set_cluster_data([dict(n) for n in case["nodes"]], pods, [], [])
        ref = case["blob"]
        assert list(blob.keys()) == list(ref.keys())
        for k in ref:
            if k in ("cpu_usage", "cpu_total", "mem_usage", "mem_total"):
                assert _close(blob[k], ref[k]), (k, blob[k], ref[k])
            else:
                assert json.loads(json.dumps(blob[k])) == ref[k], k
        assert [n["name"] for n in cr.flagged_nodes] == case["expect"]["flagged"]


def test_cluster_rollup_many_clusters(rollup):
    for n, k, hot in ((1, 1, 0.0), (4, 1, 0.0), (1000, 7, 0.1), (10000, 300, 0.05), (50000, 1, 0.02)):
        d = synth.make_nodes(11 + n, n, k, zero_frac=0.05, hot_frac=hot)
        if n >= 1000:                      # one cluster made entirely of all-zero nodes (count == 0 branch)
            sel = d["cluster_of_node"] == d["cluster_of_node"][0]
            d["cpu_usage"][sel] = 0.0
            d["mem_usage"][sel] = 0.0
        got = rollup.cluster_rollup(d["cpu"], d["mem"], d["cpu_usage"], d["mem_usage"], d["cluster_of_node"], k,
                                    80.0, bands=True)
        exp = c_oracle.cluster_rollup(d["cpu"], d["mem"], d["cpu_usage"], d["mem_usage"], d["cluster_of_node"], k)
        assert np.array_equal(got["count"], exp["count"])
        assert np.array_equal(got["flags"], exp["flags"])
        for key in ("cpu_total", "mem_total", "cpu_usage", "mem_usage"):
            assert np.array_equal(got[key], exp[key]), key          # bit for bit: sums in node order
        bc = np.array([ro.progress_band(x * 100) for x in d["cpu_usage"]], np.uint8)
        bm = np.array([ro.progress_band(x * 100) for x in d["mem_usage"]], np.uint8)
        assert np.array_equal(got["band_cpu"], bc) and np.array_equal(got["band_mem"], bm)
    # NaN usage: not zero, not flagged, sums poisoned — like the reference's float('nan')
    got = rollup.cluster_rollup([4, 8], [1, 2], [np.nan, 0.5], [0.1, 0.2], [0, 0], 1)
    exp = ro.node_rollup([dict(cpu=4, mem=1, cpu_usage=float("nan"), mem_usage=0.1),
                          dict(cpu=8, mem=2, cpu_usage=0.5, mem_usage=0.2)])
    assert got["count"][0] == exp["count"] and np.isnan(got["cpu_usage"][0]) and list(got["flags"]) == exp["flags"]
    assert _close(got["mem_usage"][0], exp["mem_usage"])


def test_cluster_rollup_rejects_ungrouped_nodes(rollup):
    from kubeoperator_b200 import KoRollupError
    with pytest.raises(KoRollupError):
        rollup.cluster_rollup([1, 1, 1], [1, 1, 1], [.1, .1, .1], [.1, .1, .1], [0, 1, 0], 2)
    with pytest.raises(KoRollupError):
        rollup.cluster_rollup([1], [1], [.1], [.1], [3], 2)


def test_pod_rollup_and_sorted_lists(rollup):
    for n in (1, 2, 33, 2048, 2049, 30000, 300000):
        d = synth.make_pods(20 + n, n)
        if n == 2049:
            d["container_restart"][:] = 1            # all keys equal: the reference recursion limit case
        got = rollup.pod_rollup(d["container_restart"], d["container_ready"], d["pod_offsets"], d["pod_phase_ok"])
        exp = c_oracle.pod_rollup(d["container_restart"], d["container_ready"], d["pod_offsets"], d["pod_phase_ok"])
        assert np.array_equal(got["restart"], exp["restart"])
        assert np.array_equal(got["warn"], exp["warn"])
        assert np.array_equal(got["order_restart"], exp["order_restart"])
        assert np.array_equal(got["order_error"], exp["order_error"])
    got = rollup.pod_rollup([], [], [0, 0, 0], [1, 0])          # pods without container statuses (:96)
    assert list(got["restart"]) == [0, 0] and list(got["order_error"]) == [1] and len(got["order_restart"]) == 0


def test_stable_desc_order_matches_reference_quicksort(rollup, golden_dir):
    for case in _load(golden_dir, "reference_sorts.json"):
        got = rollup.stable_desc_order(case["keys"])
        assert got.tolist() == case["order_pods"] == case["order_loki"]
    rng = np.random.default_rng(3)
    for n, lo, hi in ((1000000, 0, 50), (4099, -5, 5), (70000, -2 ** 31, 2 ** 31 - 1)):
        keys = rng.integers(lo, hi, n, endpoint=True).astype(np.int32)
        assert np.array_equal(rollup.stable_desc_order(keys), ro.stable_desc_order_np(keys))


def test_quick_sort_wrappers_and_dashboard_merge(rollup, golden_dir):
    from kubeoperator_b200 import cluster_monitor as cm
    rng = np.random.default_rng(4)
    blobs = []
    for c in range(5):
        rp = [dict(name="p%d-%d" % (c, i), restart_count=int(k)) for i, k in enumerate(rng.integers(1, 9, 40))]
        lk = [dict(name="l%d-%d" % (c, i), error_count=int(k), cluster_name=str(c))
              for i, k in enumerate(rng.integers(1, 5, 7))]
        blobs.append(dict(name=str(c), restart_pods=rp, warn_containers=[dict(name="w%d" % c)],
                          error_loki_containers=lk, error_pods=rp[:3]))
    blobs.insert(2, {})                                                          # api.py:494 skips empty blobs
    got = cm.dashboard_merge(blobs, rollup)
    allp = [p for b in blobs for p in b.get("restart_pods", [])]
    alll = [p for b in blobs for p in b.get("error_loki_containers", [])]
    assert got["restartPods"] == ro.quick_sort_desc(allp, "restart_count")
    assert got["errorLokiContainers"] == ro.quick_sort_desc(alll, "error_count")
    assert got["errorPods"] == [p for b in blobs for p in b.get("error_pods", [])]
    assert len(got["data"]) == 5 and json.loads(got["data"][0])["name"] == "0"
    one = [dict(restart_count=3)]
    assert cm.quick_sort_pods(one, rollup) is one                                # :451-452


def test_available_rate(rollup, golden_dir):
    from kubeoperator_b200 import cluster_monitor as cm
    g = _load(golden_dir, "reference_available_rate.json")
    up, ln, off = [], [], [0]
    for case in g["cases"]:
        up += case["job_up"]
        ln += case["job_len"]
        off.append(len(up))
    jr, cr = rollup.available_rate(up, ln, off)
    pos = 0
    for i, case in enumerate(g["cases"]):
        k = len(case["job_up"])
        assert jr[pos:pos + k].tolist() == [float(x) for x in case["rates"]]      # bit-exact doubles
        assert cr[i] == float(case["rate"])
        pos += k
    res = json.loads(json.dumps(g["full"]["result"]))
    exp_rates, exp_total = [d["rate"] for d in res["data"]], res["rate"]
    for d in res["data"]:
        d["rate"] = -1
    cm.calculate_available_rate(res, rollup)
    assert [d["rate"] for d in res["data"]] == exp_rates and res["rate"] == exp_total


def test_loki_error_count(rollup, golden_dir):
    from kubeoperator_b200 import cluster_monitor as cm
    g = _load(golden_dir, "reference_loki.json")
    lines = {nm: [e.get("line") for st in g["streams"][nm] for e in st["entries"]] for nm in g["names"]}
    assert cm.count_loki_errors(g["names"], lines, "cl", rollup) == g["result"]
    rng = np.random.default_rng(6)
    words = [b"level=error", b"level=info ", b"x", b"level=erro", b"rlevel=error!", b"", b"level=", b"error"]
    n_lines = 50000
    parts = [b" ".join(words[i] for i in rng.integers(0, len(words), rng.integers(0, 6))) for _ in range(n_lines)]
    lo = np.concatenate([[0], np.cumsum([len(p) for p in parts])]).astype(np.int64)
    go = np.concatenate([[0], np.sort(rng.integers(0, n_lines, 99)), [n_lines]]).astype(np.int64)
    buf = np.frombuffer(b"".join(parts), np.uint8)
    assert np.array_equal(rollup.loki_count(buf, lo, go), c_oracle.loki_count(buf, lo, go))


def _fleet_args(cl):
    counts = [[c["pods"], c["namespaces"], c["deployments"], c["containers"], c["nodes"],
               -1 if c.get("ui_nodes", c["nodes"]) is None else c.get("ui_nodes", c["nodes"])] for c in cl]
    return (counts, [c["cpu_total"] for c in cl], [c["mem_total"] for c in cl], [c["cpu_usage"] for c in cl],
            [c["mem_usage"] for c in cl], [c["max_pods"] for c in cl])


def test_fleet_rollup(rollup, golden_dir):
    from kubeoperator_b200 import cluster_monitor as cm
    g = _load(golden_dir, "handcomputed_fleet.json")
    got = rollup.fleet_rollup(*_fleet_args(g["clusters"]))
    for k, v in g["expect"].items():
        assert _close(got[k], v), k
    for case in g["more"]:            # hand-computed: count == 0, MAX_PODS 0 / absent, clusters the UI does not list ...
        if not case["clusters"]:
            continue
        got = rollup.fleet_rollup(*_fleet_args(case["clusters"]))
        for k, v in case["expect"].items():
            assert _close(got[k], v), (case["name"], k, got[k], v)
    rng = np.random.default_rng(8)
    K = 5000
    cl = [dict(pods=int(rng.integers(0, 500)), namespaces=int(rng.integers(0, 9)), deployments=int(rng.integers(0, 50)),
               containers=int(rng.integers(0, 900)), nodes=int(rng.integers(0, 40)),
               cpu_total=float(rng.choice([0.0, 16.0, 64.0])), mem_total=float(rng.choice([0.0, 128.5])),
               cpu_usage=float(rng.random()), mem_usage=float(rng.random()),
               max_pods=None if rng.random() < 0.5 else int(rng.integers(0, 300)),
               ui_nodes=None if rng.random() < 0.1 else int(rng.integers(0, 40))) for _ in range(K)]
    exp = ro.fleet_rollup(cl)
    got = rollup.fleet_rollup(*_fleet_args(cl))
    for k, v in exp.items():
        assert _close(got[k], v), k
    blobs = [dict(name="a", pods=[dict(containers=[1, 2])] * 3, namespaces=[1], deployments=[1, 2], nodes=[1, 2],
                  cpu_total=8.0, mem_total=4.0, cpu_usage=0.5, mem_usage=0.25),
             dict(name="gone", pods=[dict(containers=[1])] * 2, namespaces=[1], deployments=[1], nodes=[1],
                  cpu_total=4.0, mem_total=4.0, cpu_usage=0.25, mem_usage=0.25)]
    out = cm.fleet_rollup(blobs, {"a": 50}, {"a": 3}, rollup)          # "gone" is not in the UI's cluster list
    assert out["max_pod_count"] == 150 and _close(out["container_usage"], 5 / 150 * 100) and out["container_count"] == 8
    out = cm.fleet_rollup(blobs, {"a": 0}, {"a": 2, "gone": 1}, rollup)  # MAX_PODS configured as 0 is a value
    assert out["max_pod_count"] == 110


def test_flagged_nodes_reach_the_alert_sink_with_the_reference_payload(rollup, golden_dir):
    """f4 end to end: set_cluster_data's threshold flags (cluster_monitor.py:174-179) -> one message per flagged node,
    equal to the reference's CLUSTER_USAGE payload (tests/golden/reference_usage_alert.json)."""
    import json
    from kubeoperator_b200 import cluster_monitor as cm
    g = _load(golden_dir, "reference_usage_alert.json")
    c = g["cluster"]
    quiet = [dict(name="quiet", status="Ready", cpu=4.0, mem=8.0, cpu_usage=10.0, mem_usage=20.0),
             dict(name="idle", status="Ready", cpu=4.0, mem=8.0, cpu_usage=0.0, mem_usage=0.0)]
    nodes = [quiet[0]] + g["nodes"][:2] + [quiet[1]] + g["nodes"][2:]
    sent = []
    cr = cm.ClusterRollup(c["name"], rollup=rollup, alert_sink=sent.append, item_id=c["item_id"],
                          item_name=c["item_name"], status=c["status"])
    cr.rollup_nodes(nodes)
    assert [n["name"] for n in cr.flagged_nodes] == [n["name"] for n in g["nodes"]]
    assert len(sent) == len(g["nodes"])
    for msg, ref_msg, ref_content in zip(sent, g["reference_get_usage_message"], g["reference_get_usage_content"]):
        want = dict(ref_msg)
        want["content"] = ref_content
        assert json.dumps(msg, ensure_ascii=False) == json.dumps(want, ensure_ascii=False)


def test_health_day_mean_through_the_library(rollup):
    """a8: handle_cluster_health_msg_day (cluster_health_utils.py:28-33) is the arithmetic mean over all HOUR rows;
    through the library that is the ``mean`` of ko_series_rollup over the rows as one series per day."""
    rng = np.random.default_rng(31)
    days = [rng.integers(0, 101, 24).astype(np.float64) for _ in range(40)] + [np.full(24, 100.0), np.zeros(24),
                                                                                 np.array([33.0])]
    offsets = np.concatenate([[0], np.cumsum([len(d) for d in days])]).astype(np.int64)
    got = rollup.series_rollup(np.concatenate(days), offsets, ())
    for i, d in enumerate(days):
        exp = ro.health_day_mean([int(x) for x in d])
        assert abs(got["mean"][i] - exp) <= 1e-12 * max(1.0, abs(exp)), (i, got["mean"][i], exp)
        assert int(got["mean"][i]) == int(exp)              # what the IntegerField keeps (models/cluster_health_history.py:17)


def test_counter_increase(rollup):
    d = synth.make_ragged(9, [0, 1, 2, 5, 20, 240, 5760, 40320])
    vals = d["values"].copy()
    inc, rate = rollup.counter_increase(vals, d["offsets"], 15.0)
    e_inc, e_rate = c_oracle.counter_increase(vals, d["offsets"], 15.0)
    for i in range(len(inc)):
        seg = vals[d["offsets"][i]:d["offsets"][i + 1]]
        tol = RTOL * (np.nansum(np.abs(seg)) * 2 + 1e-300)
        assert abs(inc[i] - e_inc[i]) <= tol
        py = ro.counter_increase(seg.tolist())[0]
        assert abs(e_inc[i] - py) <= tol
        assert abs(rate[i] - e_rate[i]) <= tol


def test_fleet_quantile_single_process(rollup):
    from kubeoperator_b200 import fleet
    rng = np.random.default_rng(10)
    v = np.concatenate([rng.beta(2, 5, 9000), [np.nan, 0.0, -0.0, np.inf, -np.inf], rng.normal(0, 100, 500)])
    for q in (0.0, 0.5, 0.95, 0.99, 1.0):
        got = fleet.fleet_quantile(lambda s, m, p: rollup.key_histogram(v, s, m, p), int((v == v).sum()), q)
        exp = ro.fleet_quantile([v], q)
        assert np.float64(got).view(np.uint64) == np.float64(exp).view(np.uint64)
    h = rollup.key_histogram(v, 48, 0, 0)
    assert np.array_equal(h, ro.key_histogram_np(v, 48, 0, 0))


def test_fleet_quantile_selected_on_the_gpu():
    """fleet.fleet_quantile_device with one rank: NaN-padded values as one series through the library's own
    exact selection; same answer as the host path and the oracle."""
    import torch
    from kubeoperator_b200 import Rollup, fleet
    from oracle import rollup_oracle as ro
    r = Rollup(0)
    try:
        rng = np.random.default_rng(41)
        for n, cap in ((100, 100), (37, 64), (1, 8), (3000, 4096)):
            v = rng.beta(2, 5, n)
            v[rng.random(n) < 0.1] = np.nan
            ws = {}
            for q in (0.5, 0.95, 1.0):
                d = fleet.fleet_quantile_device(torch.from_numpy(v).cuda(), q, cap, r, None, ws)
                got = float(d.cpu()[0])
                exp = ro.fleet_quantile([v], q)
                assert (got != got and exp != exp) or got == exp
                assert fleet.fleet_quantile_small(v, q, cap) == exp or (exp != exp)
    finally:
        r.close()
