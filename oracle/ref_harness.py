"""TEST INFRASTRUCTURE ONLY — runs the UNMODIFIED reference rollup code under sys.modules stubs.

Used in THIS container only (it needs /root/reference, which does not exist on the GPU box) by
``oracle/gen_golden.py`` to pin ``oracle/rollup_oracle.py`` against the reference's own functions:

* ``ClusterMonitor.set_cluster_data``        core/apps/kubeops_api/cluster_monitor.py:157-192
* ``ClusterMonitor.list_pods``               core/apps/kubeops_api/cluster_monitor.py:86-120
* ``quick_sort_pods``                        core/apps/kubeops_api/cluster_monitor.py:450-463
* ``quick_sort_error_loki_container``        core/apps/kubeops_api/cluster_monitor.py:466-479
* ``PrometheusClient.calculate_available_rate`` / ``handle_targets_message``
                                             core/apps/kubeops_api/prometheus_client.py:34-85
* ``PrometheusClient.get_msg_from_loki``     core/apps/kubeops_api/prometheus_client.py:119-149
* ``ClusterMonitor.get_usage_message`` / ``get_usage_content``
                                             core/apps/kubeops_api/cluster_monitor.py:423-442

Nothing here is product code; nothing under kubeoperator_b200/ may import it.  No reference source is
copied: the modules are imported from where they lie (method: SURVEY.md §8c).
"""
import importlib
import json
import sys
import types

REF_APPS = "/root/reference/core/apps"


class _Anything(types.ModuleType):
    """Module stub: any attribute is a harmless callable class."""

    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)
        obj = type(name, (), {"__init__": lambda self, *a, **k: None,
                              "__call__": lambda self, *a, **k: None})
        setattr(self, name, obj)
        return obj


class HostTable:
    """Stand-in for the Django ``Host`` model manager (cluster_monitor.py:105, prometheus_client.py:55)."""

    def __init__(self):
        self.by_ip = {}

    def get(self, ip=None):
        return types.SimpleNamespace(name=self.by_ip.get(ip, "host-" + str(ip)), ip=ip)


HOSTS = HostTable()
SENT_MESSAGES = []


def load_reference():
    """Import kubeops_api.cluster_monitor / prometheus_client from /root/reference under stubs."""
    if "kubeops_api.cluster_monitor" in sys.modules and getattr(
            sys.modules["kubeops_api.cluster_monitor"], "_ko_harness", False):
        return sys.modules["kubeops_api.cluster_monitor"], sys.modules["kubeops_api.prometheus_client"]
    if REF_APPS not in sys.path:
        sys.path.insert(0, REF_APPS)

    class ApiException(Exception):
        def __init__(self, status=0, reason=""):
            self.status, self.reason = status, reason

    stubs = ["kubernetes", "kubernetes.client", "kubernetes.client.rest", "redis",
             "kubeoperator", "kubeoperator.settings", "log", "log.es", "django", "django.db",
             "django.db.models", "django.utils", "django.utils.timezone",
             "kubeops_api.models", "kubeops_api.models.cluster", "kubeops_api.models.host",
             "ansible_api", "ansible_api.models", "ansible_api.models.inventory",
             "common", "common.ssh", "message_center", "message_center.message_client", "requests"]
    for name in stubs:
        if name not in sys.modules or isinstance(sys.modules[name], _Anything):
            m = _Anything(name)
            m.__path__ = []
            sys.modules[name] = m
    sys.modules["kubernetes.client.rest"].ApiException = ApiException
    sys.modules["kubernetes"].client = sys.modules["kubernetes.client"]
    sys.modules["kubeops_api.models.host"].Host = types.SimpleNamespace(objects=HOSTS)

    class MessageClient:
        def insert_message(self, message):
            SENT_MESSAGES.append(message)

    sys.modules["message_center.message_client"].MessageClient = MessageClient
    cm = importlib.import_module("kubeops_api.cluster_monitor")
    pc = importlib.import_module("kubeops_api.prometheus_client")
    cm._ko_harness = True
    return cm, pc


class _FakeRedis:
    def __init__(self):
        self.store = {}

    def set(self, key, value):
        self.store[key] = value
        return True

    def get(self, key):
        v = self.store.get(key)
        return None if v is None else v.encode("utf-8")


def _pod_list_object(pods):
    """Build the V1PodList-shaped object ``list_pods`` walks (cluster_monitor.py:89-117)."""
    items = []
    for p in pods:
        cs = None
        if p.get("containers") is not None:
            cs = [types.SimpleNamespace(name=c["name"], ready=c["ready"], restart_count=c["restart_count"])
                  for c in p["containers"]]
        status = types.SimpleNamespace(container_statuses=cs, host_ip=p.get("host_ip"),
                                       pod_ip=p.get("pod_ip"), phase=p["phase"])
        meta = types.SimpleNamespace(name=p["name"], namespace=p.get("namespace", "default"))
        items.append(types.SimpleNamespace(status=status, metadata=meta))
    return types.SimpleNamespace(items=items)


def run_set_cluster_data(cluster_name, nodes, pods, namespaces=(), deployments=(), flag_mode=False):
    """Call the REAL ``set_cluster_data`` (and the REAL ``list_pods``) on synthetic collectors.

    ``nodes``: list of dicts with float cpu/mem/cpu_usage/mem_usage (strings raise).
    ``pods``: list of {name, namespace, phase, host_ip, pod_ip, containers:[{name,ready,restart_count}]}.
    ``flag_mode``: the >80 branch of the reference raises AttributeError; in flag mode
    ``get_event_message`` is stubbed so the loop completes and the flagged nodes are recorded.
    Returns (parsed Redis JSON blob, list of flagged node names).
    """
    cm, _ = load_reference()
    mon = cm.ClusterMonitor.__new__(cm.ClusterMonitor)
    mon.cluster = types.SimpleNamespace(id="00000000-0000-0000-0000-000000000001", name=cluster_name,
                                        item_id="item", item_name="item", status="RUNNING")
    mon.token = "tok"
    mon.restart_pods, mon.warn_containers, mon.error_pods = [], [], []
    mon.redis_cli = _FakeRedis()
    mon.api_instance = types.SimpleNamespace(
        list_pod_for_all_namespaces=lambda: _pod_list_object(pods))
    mon.list_nodes = lambda: [dict(n) for n in nodes]
    mon.list_namespaces = lambda: list(namespaces)
    mon.list_deployments = lambda: list(deployments)
    flagged = []
    if flag_mode:
        def _msg(node):
            flagged.append(node["name"])
            return {"flag": node["name"]}
        mon.get_event_message = _msg
    del SENT_MESSAGES[:]
    ok = mon.set_cluster_data()
    assert ok is True
    blob = json.loads(mon.redis_cli.store[cluster_name])
    return blob, flagged


def run_quick_sort_pods(pod_list):
    cm, _ = load_reference()
    old = sys.getrecursionlimit()
    sys.setrecursionlimit(max(old, 4 * len(pod_list) + 1000))
    try:
        return cm.quick_sort_pods([dict(p) for p in pod_list])
    finally:
        sys.setrecursionlimit(old)


def run_quick_sort_loki(containers):
    cm, _ = load_reference()
    old = sys.getrecursionlimit()
    sys.setrecursionlimit(max(old, 4 * len(containers) + 1000))
    try:
        return cm.quick_sort_error_loki_container([dict(c) for c in containers])
    finally:
        sys.setrecursionlimit(old)


def run_available_rate(result):
    """REAL ``calculate_available_rate`` on a prepared result dict (prometheus_client.py:75-85)."""
    _, pc = load_reference()
    cli = pc.PrometheusClient({})
    res = json.loads(json.dumps(result))
    cli.calculate_available_rate(res)
    return res


def run_handle_targets(targets_json, host_names):
    """REAL ``handle_targets_message`` (prometheus_client.py:34-73)."""
    _, pc = load_reference()
    HOSTS.by_ip = dict(host_names)
    cli = pc.PrometheusClient({})
    return cli.handle_targets_message(targets_json)


def run_loki_count(label_values, streams_by_name, cluster_name="c"):
    """REAL ``get_msg_from_loki`` against a fake AppsClient (prometheus_client.py:119-149)."""
    _, pc = load_reference()

    class Resp:
        def __init__(self, payload):
            self.ok, self._p = True, payload

        def json(self):
            return self._p

    class FakeApps:
        def __init__(self, cluster=None):
            pass

        def get(self, app, url):
            if "/label/container_name/values" in url:
                return Resp({"values": list(label_values)})
            name = url.split('container_name="', 1)[1].split('"', 1)[0]
            return Resp({"streams": streams_by_name.get(name, [])})

    old = pc.AppsClient
    pc.AppsClient = FakeApps
    try:
        cli = pc.PrometheusClient({"host": "loki", "cluster": None})
        return cli.get_msg_from_loki(cluster_name)
    finally:
        pc.AppsClient = old


def run_usage_messages(cluster, nodes):
    """REAL ``get_usage_message`` and ``get_usage_content`` (cluster_monitor.py:423-442) for every node dict.
    ``cluster``: dict with name, item_id, item_name, status.  Returns (messages, contents)."""
    cm, _ = load_reference()
    mon = cm.ClusterMonitor.__new__(cm.ClusterMonitor)
    mon.cluster = types.SimpleNamespace(**cluster)
    msgs, contents = [], []
    for n in nodes:
        node = types.SimpleNamespace(**n)            # cluster_data.py:47-55: the reference dumps node.__dict__
        msgs.append(mon.get_usage_message(node))
        contents.append(mon.get_usage_content(node))
    return msgs, contents
