"""TEST INFRASTRUCTURE ONLY — generates tests/golden/*.json.

Run in the build container (needs /root/reference):  ``python oracle/gen_golden.py``

``reference_*.json`` hold inputs and the outputs of the REFERENCE'S OWN functions executed unmodified
under stubs (oracle/ref_harness.py).  tests/test_oracle_golden.py checks oracle/rollup_oracle.py against
them on any machine (the GPU box has no /root/reference).  ``builddefined_series.json`` holds vectors
for the build-defined window statistics produced by the pure-Python oracle form (parity unpinned: the
reference has no such code) — they pin the NumPy/C twins and the CUDA path to one specification.
"""
import json
import os
import random
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)

from oracle import ref_harness as rh          # noqa: E402
from oracle import rollup_oracle as ro        # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")


def _nodes(rng, n, zero_frac=0.25, ints=False, hot=False):
    nodes = []
    for i in range(n):
        if rng.random() < zero_frac:
            cu, mu = 0, 0
        else:
            cu, mu = rng.betavariate(2, 5), rng.betavariate(2, 5)
            if hot and rng.random() < 0.3:
                cu = 80 + rng.random() * 20
        cpu = rng.choice([4, 8, 16, 64]) if ints else float(rng.choice([4, 8, 16, 64]))
        mem = int(rng.uniform(8, 512) * 2 ** 30) / 1024 / 1024 / 1024      # prometheus_client.py:109
        nodes.append(dict(name="node-%d" % i, status=None, cpu=cpu, mem=mem, cpu_usage=cu, mem_usage=mu))
    return nodes


def _pods(rng, n):
    pods = []
    for i in range(n):
        nc = rng.choice([0, 1, 1, 2, 2, 3])
        cont = None if rng.random() < 0.05 else [
            dict(name="c%d" % j, ready=rng.random() > 0.1,
                 restart_count=rng.choice([0, 0, 0, 1, 2, 5, 40])) for j in range(nc)]
        pods.append(dict(name="pod-%d" % i, namespace="ns%d" % (i % 3),
                         phase=rng.choice(["Running"] * 8 + ["Succeeded", "Pending", "Failed", "Unknown"]),
                         host_ip=None if rng.random() < 0.05 else "10.0.0.%d" % (i % 5),
                         pod_ip="1.1.%d.%d" % (i // 250, i % 250), containers=cont))
    return pods


def gen_set_cluster_data():
    cases = []
    rng = random.Random(1)
    specs = [(4, 30, 0.25, False, False), (3, 0, 1.0, False, False), (0, 5, 0.0, False, False),
             (100, 300, 0.02, False, False), (17, 64, 0.5, True, False), (1, 1, 0.0, False, False),
             (40, 120, 0.1, False, True), (7, 9, 0.0, True, True)]
    for n_nodes, n_pods, zf, ints, hot in specs:
        nodes, pods = _nodes(rng, n_nodes, zf, ints, hot), _pods(rng, n_pods)
        blob, flagged = rh.run_set_cluster_data("c", nodes, pods, flag_mode=hot)
        cases.append(dict(
            nodes=nodes, pods=pods, flag_mode=hot, blob=blob,
            expect=dict(cpu_usage=blob["cpu_usage"], cpu_total=blob["cpu_total"],
                        mem_usage=blob["mem_usage"], mem_total=blob["mem_total"],
                        flagged=flagged,
                        pod_restart=[p["restart_count"] for p in blob["pods"]],
                        restart_pods=[p["name"] for p in blob["restart_pods"]],
                        error_pods=[p["name"] for p in blob["error_pods"]],
                        warn_containers=[[c["pod_name"], c["name"]] for c in blob["warn_containers"]],
                        keys=sorted(blob.keys()))))
    return cases


def gen_sorts():
    rng = random.Random(2)
    cases = []
    for n, hi in [(0, 1), (1, 1), (2, 1), (10, 3), (100, 50), (1000, 5), (3000, 1000), (1500, 0)]:
        keys = [rng.randint(0, hi) for _ in range(n)]
        pods = [dict(name=i, restart_count=k) for i, k in enumerate(keys)]
        got = rh.run_quick_sort_pods(pods)
        lk = [dict(name=i, error_count=k, cluster_name="c") for i, k in enumerate(keys)]
        got2 = rh.run_quick_sort_loki(lk)
        cases.append(dict(keys=keys, order_pods=[p["name"] for p in got],
                          order_loki=[c["name"] for c in got2]))
    return cases


def gen_available_rate():
    rng = random.Random(3)
    cases = []
    jobs = ['kubernetes-control-manager', 'etcd', 'kubernetes-nodes', 'kubernetes-schedule',
            'kubernetes-apiservers']
    for trial in range(12):
        n_jobs = rng.choice([0, 1, 3, 5, 5, 5])
        data = []
        for j in range(n_jobs):
            ln = rng.choice([0, 1, 2, 3, 7])
            p_up = rng.choice([1.0, 1.0, 0.5, 0.9])
            data.append(dict(job=jobs[j], rate=0, data=[
                dict(key="h%d" % t, value="Ready" if rng.random() < p_up else "NotReady") for t in range(ln)]))
        res = rh.run_available_rate(dict(success=True, data=data, rate=0))
        cases.append(dict(job_up=[sum(1 for t in d["data"] if t["value"] == "Ready") for d in data],
                          job_len=[len(d["data"]) for d in data],
                          rates=[d["rate"] for d in res["data"]], rate=res["rate"]))
    # full handle_targets_message path incl. the master exclusion at prometheus_client.py:65
    targets = []
    names = {}
    for i in range(9):
        ip = "10.1.0.%d" % i
        names[ip] = ("master-%d" % i) if i < 3 else ("worker-%d" % i)
        for job in jobs + ["other-job"]:
            targets.append(dict(labels=dict(job=job), discoveredLabels={"__address__": ip + ":9100"},
                                health="up" if (i * 7 + len(job)) % 5 else "down"))
    res = rh.run_handle_targets(dict(status="success", data=dict(activeTargets=targets)), names)
    full = dict(targets=targets, names=names, result=res)
    bad = rh.run_handle_targets(dict(status="error"), names)
    return dict(cases=cases, full=full, failed=bad)


def gen_loki():
    rng = random.Random(4)
    words = ["level=error", "level=info", "level=warn", "msg=ok", "level=erro", "xlevel=errory", "err"]
    names = ["a", "b", "c", "d"]
    streams = {}
    for nm in names:
        st = []
        for _ in range(rng.randint(0, 3)):
            st.append(dict(entries=[dict(line=None if rng.random() < 0.05 else
                                         " ".join(rng.choice(words) for _ in range(rng.randint(0, 5))))
                                    for _ in range(rng.randint(0, 40))]))
        streams[nm] = st
    res = rh.run_loki_count(names, streams, "cl")
    return dict(names=names, streams=streams, result=res)


def gen_series():
    """Build-defined vectors from the pure-Python oracle form (parity unpinned)."""
    rng = random.Random(5)
    qs = [0.5, 0.95, 0.99]
    series = []
    specs = [0, 1, 2, 3, 20, 33, 240, 1000, 1025, 5760]
    for L in specs:
        kind = rng.choice(["gauge", "const", "counter", "discrete"])
        vals, acc = [], 0.0
        for t in range(L):
            if kind == "gauge":
                v = rng.betavariate(2, 5)
            elif kind == "const":
                v = 16.0
            elif kind == "counter":
                acc = 0.0 if rng.random() < 0.002 else acc + rng.gammavariate(2, 0.05)
                v = acc
            else:
                v = float(rng.randint(40, 80))
            if rng.random() < 0.01:
                v = float("nan")
            vals.append(v)
        series.append(vals)
    series.append([float("nan")] * 7)
    series.append([0.0, -0.0, 0.0, -0.0, -1.0, 1.0])
    series.append([float("inf"), 1.0, float("-inf"), 2.0, float("nan")])
    out = []
    for vals in series:
        st = ro.series_stats(vals, qs)
        inc = ro.counter_increase(vals)
        out.append(dict(values=[None if v != v else (repr(v) if v in (float("inf"), float("-inf")) else v)
                                for v in vals],
                        expect={k: (None if isinstance(x, float) and x != x else
                                    (repr(x) if isinstance(x, float) and x in (float("inf"), float("-inf")) else x))
                                for k, x in st.items() if k != "quantiles"},
                        quantiles=[None if x != x else (repr(x) if x in (float("inf"), float("-inf")) else x)
                                   for x in st["quantiles"]],
                        increase=inc[0]))
    return dict(qs=qs, series=out)


def gen_fleet():
    """Hand-computed vectors for the TypeScript fleet roll-up (dashboard.component.ts:117-153)."""
    clusters = [
        dict(pods=120, namespaces=5, deployments=20, containers=180, nodes=4, cpu_total=12.0, mem_total=46.5,
             cpu_usage=0.5, mem_usage=0.3125, max_pods=None),
        dict(pods=0, namespaces=0, deployments=0, containers=0, nodes=0, cpu_total=0, mem_total=0,
             cpu_usage=0, mem_usage=0, max_pods=None),
        dict(pods=30, namespaces=2, deployments=3, containers=45, nodes=2, cpu_total=8.0, mem_total=16.0,
             cpu_usage=0.25, mem_usage=0.5, max_pods=50),
    ]
    # by hand: count = 3-1 = 2; cpuUsage = (0.5+0+0.25)/2*100 = 37.5; memUsage = (0.3125+0+0.5)/2*100 = 40.625
    # maxPod = 110*4 + 110*0 + 50*2 = 540; containerUsage = 150/540*100
    expect = dict(pod_count=150, namespace_count=7, deployment_count=23, container_count=225, node_count=6,
                  cpu_total=20.0, mem_total=62.5, cpu_usage=37.5, mem_usage=40.625, max_pod_count=540,
                  container_usage=150 / 540 * 100, count=2)
    bands = [[0, 255], [None, 255], [0.5, 0], [60, 0], [60.0001, 1], [80, 1], [80.5, 2], [100, 2], [-5, 0]]
    # More vectors, each worked out by hand from dashboard.component.ts:117-153 (node is not available to run it).
    # ui_nodes = cluster.nodes.length of the UI's own cluster list; None = the list does not hold the cluster.
    def cl(pods, ns, dep, cont, nodes, ct, mt, cu, mu, max_pods=None, **kw):
        return dict(pods=pods, namespaces=ns, deployments=dep, containers=cont, nodes=nodes, cpu_total=ct, mem_total=mt,
                    cpu_usage=cu, mem_usage=mu, max_pods=max_pods, **kw)

    def ex(pod, ns, dep, cont, node, ct, mt, cu, mu, maxpod, cusage, count):
        return dict(pod_count=pod, namespace_count=ns, deployment_count=dep, container_count=cont, node_count=node,
                    cpu_total=ct, mem_total=mt, cpu_usage=cu, mem_usage=mu, max_pod_count=maxpod,
                    container_usage=cusage, count=count)
    more = [
        # no cluster at all: every sum 0, count 0 -> usages stay 0 (:147), maxPodCount 0 -> containerUsage 0 (:151)
        dict(name="empty fleet", clusters=[], expect=ex(0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0)),
        # every cluster all-zero: count = 2 - 2 = 0, so the usage SUMS are left undivided (0.25 + 0.5, 0.125 + 0.25)
        dict(name="count == 0", clusters=[cl(3, 1, 1, 3, 1, 0, 0, 0.25, 0.125), cl(5, 1, 2, 6, 2, 0, 0, 0.5, 0.25)],
             expect=ex(8, 2, 3, 9, 3, 0, 0, 0.75, 0.375, 110 * 1 + 110 * 2, 8 / 330 * 100, 0)),
        # cpu_total 0 but mem_total not: the cluster still counts (:131 needs both)
        dict(name="only one total zero", clusters=[cl(10, 2, 2, 12, 2, 0, 32.0, 0.5, 0.25), cl(10, 2, 2, 12, 2, 16.0, 0, 0.25, 0.5)],
             expect=ex(20, 4, 4, 24, 4, 16.0, 32.0, 0.75 / 2 * 100, 0.75 / 2 * 100, 440, 20 / 440 * 100, 2)),
        # MAX_PODS configured as 0: capacity 0 (0 !== undefined), so maxPodCount 0 and containerUsage stays 0
        dict(name="MAX_PODS 0", clusters=[cl(7, 1, 1, 7, 3, 8.0, 16.0, 0.5, 0.5, max_pods=0)],
             expect=ex(7, 1, 1, 7, 3, 8.0, 16.0, 50.0, 50.0, 0, 0, 1)),
        # MAX_PODS present / absent side by side
        dict(name="MAX_PODS mixed", clusters=[cl(100, 3, 9, 150, 5, 40.0, 160.0, 0.75, 0.5, max_pods=200),
                                               cl(20, 1, 2, 20, 1, 4.0, 8.0, 0.25, 0.25)],
             expect=ex(120, 4, 11, 170, 6, 44.0, 168.0, 50.0, 37.5, 200 * 5 + 110, 120 / 1110 * 100, 2)),
        # a cluster whose blob exists but which the UI's cluster list does not hold: it adds no capacity (:136)
        dict(name="cluster missing from the UI list", clusters=[cl(50, 2, 4, 60, 4, 16.0, 64.0, 0.5, 0.25, max_pods=100, ui_nodes=None),
                                                                 cl(10, 1, 1, 10, 2, 8.0, 16.0, 0.25, 0.75)],
             expect=ex(60, 3, 5, 70, 6, 24.0, 80.0, 37.5, 50.0, 220, 60 / 220 * 100, 2)),
        # the UI list knows a different number of nodes than the blob holds (a node was added since the last rollup)
        dict(name="ui list has more nodes", clusters=[cl(30, 1, 3, 30, 2, 8.0, 16.0, 0.5, 0.5, max_pods=50, ui_nodes=3)],
             expect=ex(30, 1, 3, 30, 2, 8.0, 16.0, 50.0, 50.0, 150, 20.0, 1)),
        # only missing clusters: maxPodCount 0 -> containerUsage 0 although there are pods
        dict(name="no capacity known", clusters=[cl(9, 1, 1, 9, 3, 12.0, 24.0, 1.0, 1.0, ui_nodes=None)],
             expect=ex(9, 1, 1, 9, 3, 12.0, 24.0, 100.0, 100.0, 0, 0, 1)),
        # pods above capacity: the UI does not clamp
        dict(name="over capacity", clusters=[cl(300, 1, 1, 300, 1, 4.0, 8.0, 0.9, 0.9, max_pods=100)],
             expect=ex(300, 1, 1, 300, 1, 4.0, 8.0, 90.0, 90.0, 100, 300.0, 1)),
    ]
    return dict(clusters=clusters, expect=expect, bands=bands, more=more)


def gen_alerts():
    """The usage alert payload of cluster_monitor.py:423-442 for nodes of every shape set_cluster_data flags."""
    cluster = dict(name="prod-a", item_id="3f2c9a10-1111-2222-3333-444455556666", item_name="项目一", status="RUNNING")
    nodes = [dict(name="worker1.prod-a", status="Ready", cpu=16.0, mem=62.5, cpu_usage=91.25, mem_usage=40.0),
             dict(name="worker2.prod-a", status="Ready", cpu=8.0, mem=31.25, cpu_usage=12.5, mem_usage=80.5),
             dict(name="master1.prod-a", status="NotReady", cpu=4.0, mem=15.5, cpu_usage=100.0, mem_usage=100.0),
             dict(name="节点四", status="Ready", cpu=64.0, mem=503.9, cpu_usage=80.00001, mem_usage=0.0)]
    msgs, contents = rh.run_usage_messages(cluster, nodes)
    return dict(cluster=cluster, nodes=nodes, reference_get_usage_message=msgs, reference_get_usage_content=contents)


def main():
    os.makedirs(OUT, exist_ok=True)
    files = {
        "reference_set_cluster_data.json": gen_set_cluster_data(),
        "reference_sorts.json": gen_sorts(),
        "reference_available_rate.json": gen_available_rate(),
        "reference_loki.json": gen_loki(),
        "builddefined_series.json": gen_series(),
        "handcomputed_fleet.json": gen_fleet(),
        "reference_usage_alert.json": gen_alerts(),
    }
    for name, obj in files.items():
        with open(os.path.join(OUT, name), "w") as f:
            json.dump(obj, f, separators=(",", ":"))
        print(name, os.path.getsize(os.path.join(OUT, name)), "bytes")


if __name__ == "__main__":
    main()
