"""CPU: host-side logic — sharding, the fleet merge over torch.distributed (gloo, world_size 2), the
synthetic generator, and the adapter's pure-Python parts.  The local histogram of the merge is supplied by
the oracle here (the CUDA kernel supplies it on a GPU: tests/test_gpu_rollup.py)."""
import os
import socket

import numpy as np
import pytest

from kubeoperator_b200 import fleet, synth
from oracle import rollup_oracle as ro


def test_shard_ranges_cover_and_balance():
    for n, w in ((10, 3), (0, 4), (7, 8), (10400, 8)):
        rs = [fleet.shard_range(n, r, w) for r in range(w)]
        assert rs[0][0] == 0 and rs[-1][1] == n
        assert all(a[1] == b[0] for a, b in zip(rs, rs[1:]))
        assert max(b - a for a, b in rs) - min(b - a for a, b in rs) <= 1
    lens = np.array([5, 5, 100, 5, 5, 5, 100, 5])
    rs = fleet.shard_by_weight(lens, 2)
    assert rs[0][0] == 0 and rs[-1][1] == len(lens) and rs[0][1] == rs[1][0]
    tot = [lens[a:b].sum() for a, b in rs]
    assert abs(tot[0] - tot[1]) <= 100


def test_synth_is_deterministic_and_shaped():
    a, b = synth.make_series("c1", seed=1, max_series=9), synth.make_series("c1", seed=1, max_series=9)
    assert np.array_equal(a["values"], b["values"], equal_nan=True)
    assert synth.n_series(synth.CONFIGS["c1"]) == 256 and synth.n_series(synth.CONFIGS["c2"]) == 10400
    assert synth.n_series(synth.CONFIGS["c3"]) == 64000
    assert a["offsets"][-1] == 9 * 5760
    p = synth.make_pods(1, 100)
    assert p["pod_offsets"][-1] == len(p["container_restart"])


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, q_out):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rng = np.random.default_rng(100)
    allv = np.concatenate([rng.beta(2, 5, 5001), [np.nan, np.nan, 0.0, -0.0, 7.5]])
    lo, hi = fleet.shard_range(len(allv), rank, world)
    mine = allv[lo:hi]
    res = {}
    for q in (0.5, 0.95, 0.99, 1.0):
        res[q] = fleet.fleet_quantile(lambda s, m, p: ro.key_histogram_np(mine, s, m, p),
                                      int((mine == mine).sum()), q, dist)
        res[("small", q)] = fleet.fleet_quantile_small(mine, q, 4096, dist)   # one all-gather instead of four rounds
    keys = rng.integers(0, 9, 400)
    lo, hi = fleet.shard_range(len(keys), rank, world)
    order = ro.stable_desc_order_np(keys[lo:hi])[:10]
    res["topk"] = fleet.fleet_topk(keys[lo:hi][order], order + lo, 10, dist)
    res["sum"] = fleet.fleet_scalars([float(rank + 1), 2.0], dist).tolist()
    if rank == 0:
        q_out.put(res)
    dist.barrier()
    dist.destroy_process_group()


def test_fleet_merge_gloo_world2():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    rng = np.random.default_rng(100)
    allv = np.concatenate([rng.beta(2, 5, 5001), [np.nan, np.nan, 0.0, -0.0, 7.5]])
    for qv in (0.5, 0.95, 0.99, 1.0):
        assert res[qv] == ro.fleet_quantile([allv], qv)
        assert res[("small", qv)] == res[qv]
    keys = rng.integers(0, 9, 400)
    assert res["topk"] == ro.stable_desc_order_np(keys)[:10].tolist()
    assert res["sum"] == [3.0, 4.0]


def test_fleet_quantile_empty_and_single():
    assert np.isnan(fleet.fleet_quantile_small(np.array([np.nan, np.nan]), 0.5, 8))
    assert fleet.fleet_quantile_small(np.array([3.25]), 0.99, 8) == 3.25
    assert np.isnan(fleet.fleet_quantile(lambda s, m, p: np.zeros(65536, np.int64), 0, 0.5))
    v = np.array([3.25])
    assert fleet.fleet_quantile(lambda s, m, p: ro.key_histogram_np(v, s, m, p), 1, 0.99) == 3.25


def test_cluster_data_dict_key_order():
    from kubeoperator_b200.cluster_monitor import cluster_data_dict
    d = cluster_data_dict(1, "n", "t", [], [], [], [], 0, 0, 0, 0, [], [], [], [])
    assert list(d) == ["id", "name", "pods", "nodes", "token", "namespaces", "deployments", "cpu_usage", "cpu_total",
                       "mem_usage", "mem_total", "restart_pods", "warn_containers", "error_loki_containers",
                       "error_pods"]          # cluster_data.py:9-23


def test_usage_alert_payload_equals_reference(golden_dir):
    """f4: the CLUSTER_USAGE message of cluster_monitor.py:423-442.  get_usage_content must be the reference's dict
    (same keys, order, values); get_usage_message is the reference's message with that content (the reference wires
    get_event_content there by mistake, which labels the alert CLUSTER_EVENT)."""
    import json
    import os
    from kubeoperator_b200 import cluster_monitor as cm
    g = json.load(open(os.path.join(golden_dir, "reference_usage_alert.json")))
    c = g["cluster"]
    cr = cm.ClusterRollup(c["name"], rollup=object(), item_id=c["item_id"], item_name=c["item_name"], status=c["status"])
    for node, ref_msg, ref_content in zip(g["nodes"], g["reference_get_usage_message"], g["reference_get_usage_content"]):
        content = cr.get_usage_content(node)
        assert content == ref_content and list(content) == list(ref_content)
        msg = cr.get_usage_message(node)
        want = dict(ref_msg)
        assert want["content"]["resource_type"] == "CLUSTER_EVENT"        # the reference's slip, recorded
        want["content"] = ref_content
        assert msg == want and list(msg) == list(ref_msg)
        assert json.dumps(msg, ensure_ascii=False) == json.dumps(want, ensure_ascii=False)


def test_fleet_topk_device_on_cpu_tensors():
    """fleet_topk_device is plain torch (two stable descending sorts): the same code runs on CPU tensors."""
    import torch
    rng = np.random.default_rng(9)
    v = rng.integers(0, 7, 300).astype(np.float64)
    v[rng.random(300) < 0.1] = np.nan
    v[5], v[200] = np.inf, -np.inf
    tv, ti = fleet.fleet_topk_device(torch.from_numpy(v), 40, 12)
    keys = np.where(np.isnan(v), -np.inf, v)
    order = np.argsort(-keys, kind="stable")[:12]
    assert np.array_equal(ti.numpy(), order + 40)
    assert np.array_equal(tv.numpy(), keys[order])
    tv, ti = fleet.fleet_topk_device(torch.tensor([np.nan, 2.5], dtype=torch.float64), 0, 3)
    assert ti.tolist() == [1, -1, -1] and tv.tolist()[0] == 2.5
    tv, ti = fleet.fleet_topk_device(torch.zeros(0, dtype=torch.float64), 0, 2)
    assert ti.tolist() == [-1, -1]
