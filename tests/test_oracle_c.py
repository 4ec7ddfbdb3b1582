"""CPU: the plain-C oracle (oracle/korollup_oracle.c) against the Python oracle on seeded inputs."""
import numpy as np

from oracle import c_oracle, rollup_oracle as ro
from kubeoperator_b200 import synth


def _eq(a, b):
    return np.array_equal(np.asarray(a), np.asarray(b), equal_nan=True)


def test_series_c_equals_python_forms():
    d = synth.make_ragged(1, [0, 1, 2, 7, 33, 240, 1025, 5760, 9000])
    qs = [0.5, 0.95, 0.99, 0.0]
    c = c_oracle.series_rollup(d["values"], d["offsets"], qs, nthreads=2)
    n = ro.series_rollup_np(d["values"], d["offsets"], qs)
    for k in c:
        assert _eq(c[k], n[k]), k
    for i in range(len(d["offsets"]) - 1):
        st = ro.series_stats(d["values"][d["offsets"][i]:d["offsets"][i + 1]].tolist(), qs)
        assert _eq([st["sum"], st["mean"], st["min"], st["max"], st["last"]],
                   [c[k][i] for k in ("sum", "mean", "min", "max", "last")])
        assert st["count"] == c["count"][i] and _eq(st["quantiles"], c["quantiles"][i])


def test_cluster_and_pod_c_equal_python():
    d = synth.make_nodes(2, 500, 9, zero_frac=0.1, hot_frac=0.1)
    c = c_oracle.cluster_rollup(d["cpu"], d["mem"], d["cpu_usage"], d["mem_usage"], d["cluster_of_node"], 9)
    for k in range(9):
        sel = d["cluster_of_node"] == k
        nodes = [dict(cpu=a, mem=b, cpu_usage=x, mem_usage=y) for a, b, x, y in
                 zip(d["cpu"][sel], d["mem"][sel], d["cpu_usage"][sel], d["mem_usage"][sel])]
        e = ro.node_rollup(nodes)
        assert [c["cpu_total"][k], c["mem_total"][k], c["cpu_usage"][k], c["mem_usage"][k], c["count"][k]] == \
            [e["cpu_total"], e["mem_total"], e["cpu_usage"], e["mem_usage"], e["count"]]
        assert list(c["flags"][sel]) == e["flags"]
    p = synth.make_pods(3, 700)
    c = c_oracle.pod_rollup(p["container_restart"], p["container_ready"], p["pod_offsets"], p["pod_phase_ok"])
    pods = [dict(phase_ok=bool(p["pod_phase_ok"][i]),
                 containers=[(int(p["container_restart"][j]), bool(p["container_ready"][j]))
                             for j in range(p["pod_offsets"][i], p["pod_offsets"][i + 1])]) for i in range(700)]
    e = ro.pod_rollup(pods)
    assert list(c["restart"]) == e["restart"] and list(c["warn"]) == e["warn"]
    order = ro.stable_desc_order([e["restart"][i] for i in e["restart_members"]])
    assert list(c["order_restart"]) == [e["restart_members"][i] for i in order]
    order = ro.stable_desc_order([e["restart"][i] for i in e["error_members"]])
    assert list(c["order_error"]) == [e["error_members"][i] for i in order]


def test_rate_and_increase_c_equal_python():
    assert c_oracle.available_rate([1, 2, 0], [2, 2, 0])[1] == ro.available_rate([1, 2, 0], [2, 2, 0])[1]
    d = synth.make_ragged(4, [0, 1, 2, 50, 3000])
    inc, rate = c_oracle.counter_increase(d["values"], d["offsets"], 15.0)
    for i in range(5):
        seg = d["values"][d["offsets"][i]:d["offsets"][i + 1]].tolist()
        assert _eq(inc[i], ro.counter_increase(seg)[0]) and _eq(rate[i], ro.counter_rate(seg, 15.0))
