"""Host-side mirror of the reference's rollup interface, with the arithmetic done by libkorollup.so.

Same names, argument meaning and result layout as the reference functions they stand in for, so the
reference can call them in place of its Python loops (patch: INTEGRATION.md):

=============================================  =====================================================
here                                            reference (core/apps/kubeops_api/…)
=============================================  =====================================================
``ClusterRollup.set_cluster_data``              ``ClusterMonitor.set_cluster_data`` cluster_monitor.py:157-192
``ClusterRollup.list_pods``                     ``ClusterMonitor.list_pods``        cluster_monitor.py:86-120
``quick_sort_pods``                             ``quick_sort_pods``                 cluster_monitor.py:450-463
``quick_sort_error_loki_container``             ``quick_sort_error_loki_container`` cluster_monitor.py:466-479
``calculate_available_rate``                    ``PrometheusClient.calculate_available_rate`` prometheus_client.py:75-85
``count_loki_errors``                           loop of ``get_msg_from_loki``       prometheus_client.py:132-148
``dashboard_merge``                             ``DashBoardView.get`` 'all' branch  api.py:491-501
``fleet_rollup``                                ``DashboardComponent.getClusterData`` ui/src/app/dashboard/dashboard.component.ts:117-153
=============================================  =====================================================

Deviations from the reference, all forced by its latent defects:
* node values are converted with ``float()`` BEFORE the zero / threshold tests (the reference raises
  ``TypeError`` on the strings Prometheus returns, cluster_monitor.py:176);
* the alert branch (cluster_monitor.py:177-179, which raises ``AttributeError``) is reported as the list of
  flagged nodes; ``alert_sink`` may turn them into the ``CLUSTER_USAGE`` message the reference meant to send
  (cluster_monitor.py:423-442);
* the quick sorts do not mutate their argument (the reference ``remove``s the pivot, cluster_monitor.py:456).
No I/O happens here: collectors, Redis and REST stay in the reference.
"""
import json

import numpy as np

_default_rollup = None


def get_rollup(device=0):
    """Process-wide lazily created context (create it AFTER the Celery fork)."""
    global _default_rollup
    if _default_rollup is None:
        from .korollup import Rollup
        _default_rollup = Rollup(device)
    return _default_rollup


# ------------------------------------------------------------------------------------------ wire format
def cluster_data_dict(cluster_id, cluster_name, token, pods, nodes, namespaces, deployments, cpu_usage, cpu_total,
                      mem_usage, mem_total, restart_pods, warn_containers, error_loki_containers, error_pods):
    """``ClusterData(...).__dict__`` — same keys in the same order as cluster_data.py:9-23."""
    return {"id": str(cluster_id), "name": cluster_name, "pods": pods, "nodes": nodes, "token": token,
            "namespaces": namespaces, "deployments": deployments, "cpu_usage": cpu_usage, "cpu_total": cpu_total,
            "mem_usage": mem_usage, "mem_total": mem_total, "restart_pods": restart_pods,
            "warn_containers": warn_containers, "error_loki_containers": error_loki_containers,
            "error_pods": error_pods}


def _order_to_list(items, order):
    return [items[int(i)] for i in order]


def quick_sort_pods(pod_list, rollup=None):
    """Stable descending sort by ``restart_count`` (cluster_monitor.py:450-463) on the GPU."""
    if len(pod_list) < 2:                                       # :451
        return pod_list
    rollup = rollup or get_rollup()
    order = rollup.stable_desc_order([int(p["restart_count"]) for p in pod_list])
    return _order_to_list(pod_list, order)


def quick_sort_error_loki_container(containers, rollup=None):
    """Stable descending sort by ``error_count`` (cluster_monitor.py:466-479) on the GPU."""
    if len(containers) < 2:                                     # :467
        return containers
    rollup = rollup or get_rollup()
    order = rollup.stable_desc_order([int(c.get("error_count")) for c in containers])
    return _order_to_list(containers, order)


def calculate_available_rate(result, rollup=None):
    """In-place like prometheus_client.py:75-85: fills ``res['rate']`` per job and ``result['rate']``."""
    rollup = rollup or get_rollup()
    data = result["data"]
    up = [sum(1 for job in res["data"] if job["value"] == "Ready") for res in data]     # :79-81
    ln = [len(res["data"]) for res in data]
    if not data:
        result["rate"] = 0                                                                # :85
        return
    job_rate, cluster_rate = rollup.available_rate(up, ln, [0, len(data)])
    for res, r, n in zip(data, job_rate, ln):
        res["rate"] = float(r) if n > 0 else 0                                            # :82
    result["rate"] = float(cluster_rate[0])


def count_loki_errors(names, lines_by_name, cluster_name, rollup=None):
    """The counting loop of get_msg_from_loki (prometheus_client.py:132-148): per container name the number
    of lines containing 'level=error'; containers with zero hits are dropped (:146).  ``lines_by_name`` maps
    a name to its list of log lines (None entries are skipped like :144)."""
    rollup = rollup or get_rollup()
    chunks, line_off, group_off = [], [0], [0]
    for name in names:
        for line in lines_by_name.get(name, []):
            if line is None:
                continue
            b = line.encode("utf-8")
            chunks.append(b)
            line_off.append(line_off[-1] + len(b))
        group_off.append(len(line_off) - 1)
    buf = np.frombuffer(b"".join(chunks), dtype=np.uint8) if chunks else np.zeros(0, np.uint8)
    counts = rollup.loki_count(buf, line_off, group_off)
    out = []
    for name, c in zip(names, counts):
        if c > 0:                                                                         # :146
            out.append({"name": name, "error_count": int(c), "cluster_name": cluster_name})   # cluster_data.py:76-81
    return out


class ClusterRollup:
    """Arithmetic of ``ClusterMonitor`` for one cluster; collectors are passed in by the caller."""

    def __init__(self, cluster_name, cluster_id="", token="", rollup=None, threshold=80, alert_sink=None,
                 item_id=None, item_name=None, status=None):
        self.cluster_name = cluster_name
        self.cluster_id = cluster_id
        self.token = token
        self.item_id, self.item_name, self.status = item_id, item_name, status   # cluster.item_id / item_name / status
        self.rollup = rollup or get_rollup()
        self.threshold = threshold
        self.alert_sink = alert_sink
        self.restart_pods = []            # cluster_monitor.py:31-33
        self.warn_containers = []
        self.error_pods = []
        self.flagged_nodes = []

    def list_pods(self, pod_items, host_name_of=None):
        """``pod_items``: the ``items`` of a V1PodList (objects with .status/.metadata as the kubernetes
        client returns them).  Builds the same ``Pod.__dict__`` list as cluster_monitor.py:90-117 and fills
        ``restart_pods`` / ``warn_containers`` / ``error_pods`` (unsorted, like the reference at that point).
        ``host_name_of(ip)`` stands in for the ``Host.objects.get`` lookup of :105-106."""
        c_restart, c_ready, offsets, phase_ok = [], [], [0], []
        for p in pod_items:
            cs = p.status.container_statuses
            if cs is not None:                                                        # :96
                for c in cs:
                    c_restart.append(int(c.restart_count))
                    c_ready.append(0 if c.ready == False else 1)                      # noqa: E712  (:101)
            offsets.append(len(c_restart))
            phase_ok.append(1 if (p.status.phase == 'Running' or p.status.phase == 'Succeeded') else 0)   # :115
        r = self.rollup.pod_rollup(c_restart, c_ready, offsets, phase_ok)
        pod_list, flat_containers = [], []
        for i, p in enumerate(pod_items):
            status = p.status
            containers = []
            if status.container_statuses is not None:
                for c in status.container_statuses:
                    containers.append({"name": c.name, "ready": c.ready, "restart_count": c.restart_count,
                                       "pod_name": p.metadata.name})                  # cluster_data.py:58-64
                flat_containers.extend(containers)
            hostname = host_ip = None
            if status.host_ip is not None:                                            # :104-107
                hostname = host_name_of(status.host_ip) if host_name_of else None
                host_ip = status.host_ip
            pod_ip = status.pod_ip if status.pod_ip is not None else None             # :108
            pod_list.append({"name": p.metadata.name, "cluster_name": self.cluster_name,
                             "restart_count": int(r["restart"][i]), "status": status.phase,
                             "namespace": p.metadata.namespace, "host_ip": host_ip, "pod_ip": pod_ip,
                             "host_name": hostname, "containers": containers})        # cluster_data.py:26-37
        self.warn_containers.extend(flat_containers[int(i)] for i in np.flatnonzero(r["warn"]))
        # members in input order (what the reference holds before :183-184) ...
        self.restart_pods.extend(pod_list[i] for i in sorted(int(x) for x in r["order_restart"]))
        self.error_pods.extend(pod_list[i] for i in sorted(int(x) for x in r["order_error"]))
        # ... and the already sorted lists, reused by set_cluster_data when nothing else was appended
        self._sorted = ([pod_list[int(i)] for i in r["order_restart"]], [pod_list[int(i)] for i in r["order_error"]],
                        len(self.restart_pods), len(self.error_pods))
        return pod_list

    def rollup_nodes(self, nodes):
        """The loop of cluster_monitor.py:163-182 for this cluster.  Returns (cpu_usage, cpu_total, mem_usage,
        mem_total) and records ``flagged_nodes`` (the nodes for which the reference enters :176)."""
        n = len(nodes)
        if n == 0:
            self.flagged_nodes = []
            return 0, 0, 0, 0                                                         # :163-166 untouched
        cols = [[float(x[k]) for x in nodes] for k in ("cpu", "mem", "cpu_usage", "mem_usage")]   # :170-173
        out = self.rollup.cluster_rollup(cols[0], cols[1], cols[2], cols[3], np.zeros(n, np.int32), 1,
                                         float(self.threshold))
        self.flagged_nodes = [nodes[int(i)] for i in np.flatnonzero(out["flags"])]
        if self.alert_sink is not None:
            for node in self.flagged_nodes:
                self.alert_sink(self.get_usage_message(node))
        return (float(out["cpu_usage"][0]), float(out["cpu_total"][0]), float(out["mem_usage"][0]),
                float(out["mem_total"][0]))

    def get_usage_content(self, node):
        """``ClusterMonitor.get_usage_content`` (cluster_monitor.py:433-442), same keys in the same order; ``node``
        is the node's dict (the reference serialises ``node.__dict__`` of cluster_data.py:47-55)."""
        return {"item_name": self.item_name, "resource": "集群", "resource_name": self.cluster_name,
                "resource_type": "CLUSTER_USAGE", "detail": json.dumps(node), "status": self.status}

    def get_usage_message(self, node):
        """``ClusterMonitor.get_usage_message`` (cluster_monitor.py:423-431) with the content of
        ``get_usage_content``: the reference passes the node to ``get_event_content`` there (:427), which labels a
        usage alert CLUSTER_EVENT; get_usage_content exists for exactly this message and is what is used here."""
        return {"item_id": self.item_id, "title": "集群资源告警", "content": self.get_usage_content(node),
                "level": "WARNING", "type": "CLUSTER"}

    def set_cluster_data(self, nodes, pods, namespaces, deployments):
        """Returns the dict the reference serialises into Redis at cluster_monitor.py:186-192.  ``pods`` is the
        result of :meth:`list_pods` (which also filled the member lists, as in the reference :159)."""
        cpu_usage, cpu_total, mem_usage, mem_total = self.rollup_nodes(nodes)
        srt = getattr(self, "_sorted", None)
        if srt is not None and srt[2] == len(self.restart_pods) and srt[3] == len(self.error_pods):
            sort_restart_pod_list, error_pods = srt[0], srt[1]
        else:
            sort_restart_pod_list = quick_sort_pods(self.restart_pods, self.rollup)   # :183
            error_pods = quick_sort_pods(self.error_pods, self.rollup)                # :184
        return cluster_data_dict(self.cluster_id, self.cluster_name, self.token, pods, nodes, namespaces,
                                 deployments, cpu_usage, cpu_total, mem_usage, mem_total, sort_restart_pod_list,
                                 self.warn_containers, [], error_pods)                # :186-191


def dashboard_merge(cluster_blobs, rollup=None):
    """The 'all' branch of DashBoardView.get (api.py:491-501) on already loaded Redis blobs: concatenate the
    four lists over clusters, then re-sort restart_pods and error_loki_containers fleet-wide."""
    restart_pods, warn_containers, error_loki_containers, error_pods, cluster_data = [], [], [], [], []
    for res in cluster_blobs:
        if len(res) != 0:                                                             # :494
            restart_pods = restart_pods + res.get('restart_pods', [])
            warn_containers = warn_containers + res.get('warn_containers', [])
            error_loki_containers = error_loki_containers + res.get('error_loki_containers', [])
            error_pods = error_pods + res.get('error_pods', [])
            cluster_data.append(json.dumps(res))                                      # :499
    restart_pods = quick_sort_pods(restart_pods, rollup)                              # :500
    error_loki_containers = quick_sort_error_loki_container(error_loki_containers, rollup)   # :501
    return {'data': cluster_data, 'warnContainers': warn_containers, 'restartPods': restart_pods,
            'errorLokiContainers': error_loki_containers, 'errorPods': error_pods}    # :513-514


def fleet_rollup(cluster_blobs, max_pods_by_name=None, nodes_by_name=None, rollup=None):
    """dashboard.component.ts:117-153 over parsed blobs.  ``max_pods_by_name`` = cluster.configs['MAX_PODS'] of
    the clusters that configure it (a missing name = undefined -> 110; 0 is a configured zero), ``nodes_by_name`` =
    len(cluster.nodes) of the cluster list the UI holds (:136-143); a cluster missing from that list adds no pod
    capacity, as in the UI loop.  Without ``nodes_by_name`` every blob's own node count is used."""
    rollup = rollup or get_rollup()
    max_pods_by_name = max_pods_by_name or {}
    counts, cols, mp = [], [[], [], [], []], []
    for d in cluster_blobs:
        ui_nodes = len(d["nodes"]) if nodes_by_name is None else nodes_by_name.get(d["name"], -1)
        counts.append([len(d["pods"]), len(d["namespaces"]), len(d["deployments"]),
                       sum(len(p["containers"]) for p in d["pods"]), len(d["nodes"]), int(ui_nodes)])
        for c, k in zip(cols, ("cpu_total", "mem_total", "cpu_usage", "mem_usage")):
            c.append(float(d[k]))
        mp.append(max_pods_by_name.get(d["name"]))
    return rollup.fleet_rollup(counts, cols[0], cols[1], cols[2], cols[3], mp)
