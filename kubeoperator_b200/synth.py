"""Deterministic synthetic Prometheus dumps for the BASELINE.json configs, columnar:
``values`` f64[] (all series back to back), ``offsets`` i64[n_series+1], ``kind`` per series.

NumPy generation (``make_series``) is used by tests, smoke and the bounded CPU sample; bench.py builds the
full-size column on the GPU with torch from the same distributions (``make_series_torch``) because drawing
4e8 Beta variates on the host would take longer than the whole bench.
"""
import numpy as np

STEP_SECONDS = 15
KIND_GAUGE, KIND_CONST, KIND_COUNTER, KIND_DISCRETE = 0, 1, 2, 3

# name -> (nodes, gpus per node, pods, samples per series)
CONFIGS = {
    "c1": dict(nodes=4, gpus=0, pods=120, L=5760),          # 1 master + 3 workers, 24 h
    "c2": dict(nodes=100, gpus=8, pods=3000, L=40320),      # 100 nodes incl. DCGM series, 7 d
    "c3": dict(nodes=1000, gpus=0, pods=30000, L=172800),   # 1000 nodes x 30 pods, 30 d
    # 10 000-node multi-cluster fleet, 30 d: 8 node-level series per node (the 4 of the reference + 4 GPU aggregates)
    "c4": dict(nodes=10000, gpus=0, pods=0, L=172800, per_node=8, clusters=100),
}


def series_layout(cfg):
    """Kinds of the series of a config: per node cpu_usage(gauge) cpu(const) mem(const) mem_usage(gauge);
    per GPU util(discrete %) fb_used(gauge) fb_free(gauge) temp(discrete) power(gauge);
    per pod cpu(counter) mem(gauge)."""
    kinds = []
    if cfg.get("per_node", 4) == 8:     # + GPU aggregates per node: mean util (gauge), fb used (gauge), max temp (discrete), power (gauge)
        return np.array([KIND_GAUGE, KIND_CONST, KIND_CONST, KIND_GAUGE, KIND_GAUGE, KIND_GAUGE, KIND_DISCRETE,
                         KIND_GAUGE] * cfg["nodes"], np.int8)
    kinds += [KIND_GAUGE, KIND_CONST, KIND_CONST, KIND_GAUGE] * cfg["nodes"]
    kinds += [KIND_DISCRETE, KIND_GAUGE, KIND_GAUGE, KIND_DISCRETE, KIND_GAUGE] * (cfg["nodes"] * cfg["gpus"])
    kinds += [KIND_COUNTER, KIND_GAUGE] * cfg["pods"]
    return np.array(kinds, np.int8)


def mix_kinds(config, n):
    """``n`` series with the same proportions of kinds as the config (interleaved, not in blocks)."""
    full = series_layout(CONFIGS[config] if isinstance(config, str) else config)
    return np.random.default_rng(12345).choice(full, n)


def n_series(cfg):
    return cfg.get("per_node", 4) * cfg["nodes"] + 5 * cfg["nodes"] * cfg["gpus"] + 2 * cfg["pods"]


def _one_series(rng, kind, L):
    if kind == KIND_GAUGE:
        v = rng.beta(2.0, 5.0, L)
        hot = rng.random(L) < 0.01
        v = np.where(hot, 0.8 + 0.2 * v, v)              # 1 % of usage samples above the 0.8 band
    elif kind == KIND_CONST:
        v = np.full(L, float(rng.choice([4, 8, 16, 64])))
    elif kind == KIND_COUNTER:
        inc = rng.gamma(2.0, 0.05, L)
        reset = rng.random(L) < 1e-5
        v = np.empty(L)
        acc = 0.0
        start = 0
        for r in list(np.flatnonzero(reset)) + [L]:
            seg = np.cumsum(inc[start:r])
            v[start:r] = seg
            start = r
        del acc
    else:
        v = np.floor(40.0 + 40.0 * rng.beta(2.0, 2.0, L))  # integer-valued (temperature / util style)
    v = v.astype(np.float64)
    v[rng.random(L) < 1e-3] = np.nan                      # staleness markers
    return v


def make_series(config="c1", seed=1, L=None, max_series=None, all_zero_frac=0.02):
    """Columnar dump of a config (optionally truncated to ``max_series`` series / ``L`` samples)."""
    cfg = dict(CONFIGS[config]) if isinstance(config, str) else dict(config)
    if L is not None:
        cfg["L"] = L
    kinds = series_layout(cfg)
    if max_series is not None:
        kinds = kinds[:max_series]
    rng = np.random.default_rng(seed)
    n = len(kinds)
    L = cfg["L"]
    values = np.empty(n * L, np.float64)
    for i, k in enumerate(kinds):
        values[i * L:(i + 1) * L] = _one_series(rng, int(k), L)
    # 2 % of nodes all-zero (exercises cluster_monitor.py:174): zero both usage series of the node
    for node in range(min(cfg["nodes"], n // 4)):
        if rng.random() < all_zero_frac:
            for s in (4 * node, 4 * node + 3):
                values[s * L:(s + 1) * L] = 0.0
    offsets = np.arange(n + 1, dtype=np.int64) * L
    return dict(values=values, offsets=offsets, kinds=kinds, cfg=cfg)


def make_ragged(seed, lengths):
    """Ragged batch for edge-case tests: one series per entry of ``lengths``, kinds cycling."""
    rng = np.random.default_rng(seed)
    parts = [_one_series(rng, i % 4, int(L)) if L > 0 else np.zeros(0) for i, L in enumerate(lengths)]
    offsets = np.concatenate([[0], np.cumsum([len(p) for p in parts])]).astype(np.int64)
    values = np.concatenate(parts) if parts else np.zeros(0)
    return dict(values=values, offsets=offsets)


def make_stress(seed):
    """Randomised batch for parity stress: a mixture of series kinds (the four of the configs plus quantised
    values of both signs, heavy tails, random staleness rates, integer random walks, stray infinities), lengths
    from empty to 70 000 samples, optional odd alignment, and a random set of up to four quantiles."""
    rng = np.random.default_rng(1000 + seed)
    n = int(rng.integers(20, 200))
    series = []
    for _ in range(n):
        L = int(rng.choice([rng.integers(0, 40), rng.integers(200, 1100), rng.integers(1025, 9000),
                            rng.integers(9000, 70000), 40320]))
        kind = int(rng.integers(0, 8))
        if kind < 4:
            v = _one_series(rng, kind, L) if L > 0 else np.zeros(0)
        elif kind == 4:
            v = np.round(rng.normal(0, 1, L), int(rng.integers(0, 3)))
        elif kind == 5:
            v = rng.standard_cauchy(L) * 10.0 ** rng.integers(-5, 6)
        elif kind == 6:
            v = np.where(rng.random(L) < rng.random(), np.nan, rng.random(L))
        else:
            v = np.cumsum(rng.integers(-1, 2, L)).astype(np.float64)
        if L > 10 and rng.random() < 0.1:
            v[rng.integers(0, L, 3)] = [np.inf, -np.inf, np.nan]
        series.append(np.asarray(v, np.float64))
    offsets = np.concatenate([[0], np.cumsum([len(x) for x in series])]).astype(np.int64)
    values = np.concatenate(series) if series else np.zeros(0)
    if rng.random() < 0.5:
        values = np.concatenate([[0.5], values])
        offsets = np.concatenate([[0], offsets + 1])
    nq = int(rng.integers(0, 5))
    grid = [0.0, 0.01, 0.05, 0.25, 0.5, 0.75, 0.9, 0.95, 0.99, 0.999, 1.0]
    qs = sorted(set(np.round(rng.choice(grid, nq), 3).tolist()))
    rng.shuffle(qs)
    return dict(values=values, offsets=offsets, qs=qs)


def make_pods(seed, n_pods):
    """Columnar pod list: Poisson(0.3) restarts per container, 3 % not ready, 2 % pods not Running."""
    rng = np.random.default_rng(seed)
    n_cont = rng.choice([1, 1, 2, 2, 3], n_pods)
    offsets = np.concatenate([[0], np.cumsum(n_cont)]).astype(np.int64)
    c = int(offsets[-1])
    return dict(container_restart=rng.poisson(0.3, c).astype(np.int32),
                container_ready=(rng.random(c) >= 0.03).astype(np.uint8),
                pod_offsets=offsets, pod_phase_ok=(rng.random(n_pods) >= 0.02).astype(np.uint8))


def make_nodes(seed, n_nodes, n_clusters=1, zero_frac=0.02, hot_frac=0.0):
    """Instant node columns as ClusterMonitor.list_nodes delivers them (floats)."""
    rng = np.random.default_rng(seed)
    cu, mu = rng.beta(2, 5, n_nodes), rng.beta(2, 5, n_nodes)
    zero = rng.random(n_nodes) < zero_frac
    cu[zero] = 0.0
    mu[zero] = 0.0
    hot = rng.random(n_nodes) < hot_frac
    cu[hot & ~zero] = 80.0 + 20.0 * rng.random(int((hot & ~zero).sum()))
    cpu = rng.choice([4.0, 8.0, 16.0, 64.0], n_nodes)
    mem = np.floor(rng.uniform(8, 512, n_nodes) * 2 ** 30) / 1024 / 1024 / 1024
    cluster = np.sort(rng.integers(0, n_clusters, n_nodes)).astype(np.int32)
    return dict(cpu=cpu, mem=mem, cpu_usage=cu, mem_usage=mu, cluster_of_node=cluster)


def make_series_torch(config, seed, device, L=None, n_series_override=None, kinds=None, rows=256, zero_nodes=True):
    """Full-size column generated ON THE GPU (torch, float64 throughout) from the same distributions as
    make_series: Beta draws as ratios of Gamma variates, counters as cumulative Gamma(2, 0.05) increments with
    resets (p = 1e-5 per sample), staleness NaN (p = 1e-3), 2 % of the nodes all-zero.
    Returns (values tensor f64 [n*L], offsets tensor i64 [n+1], kinds ndarray).  ``kinds`` replaces the config's
    series layout (e.g. a shorter batch with the same mix of kinds; no all-zero nodes then)."""
    import torch
    cfg = dict(CONFIGS[config]) if isinstance(config, str) else dict(config)
    if L is not None:
        cfg["L"] = L
    layout = kinds is None
    kinds = series_layout(cfg) if kinds is None else np.asarray(kinds, np.int8)
    if n_series_override is not None:
        kinds = np.resize(kinds, n_series_override)
    n, L = len(kinds), cfg["L"]
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    f64 = torch.float64
    values = torch.empty(n * L, dtype=f64, device=device)
    view = values.view(n, L)
    kt = torch.from_numpy(kinds.astype(np.int64)).to(device)

    def beta(a, b, m):
        x = torch._standard_gamma(torch.full((m, L), a, dtype=f64, device=device), generator=g)
        y = torch._standard_gamma(torch.full((m, L), b, dtype=f64, device=device), generator=g)
        return x.div_(x + y)

    for r0 in range(0, n, rows):                           # row blocks bound the temporaries
        r1 = min(n, r0 + rows)
        k = kt[r0:r1]
        blk = view[r0:r1]
        for kind in (KIND_GAUGE, KIND_CONST, KIND_COUNTER, KIND_DISCRETE):
            idx = (k == kind).nonzero().flatten()
            m = int(idx.numel())
            if m == 0:
                continue
            if kind == KIND_GAUGE:
                v = beta(2.0, 5.0, m)
                hot = torch.rand((m, L), generator=g, device=device) < 0.01
                v = torch.where(hot, 0.8 + 0.2 * v, v)     # 1 % of usage samples above the 0.8 band
            elif kind == KIND_CONST:
                c = torch.tensor([4.0, 8.0, 16.0, 64.0], device=device, dtype=f64)
                v = c[torch.randint(0, 4, (m, 1), generator=g, device=device)].expand(m, L)
            elif kind == KIND_COUNTER:
                inc = torch._standard_gamma(torch.full((m, L), 2.0, dtype=f64, device=device), generator=g).mul_(0.05)
                reset = torch.rand((m, L), generator=g, device=device) < 1e-5
                c = torch.cumsum(inc, dim=1)
                base = torch.cummax(torch.where(reset, c - inc, torch.zeros_like(c)), dim=1).values
                v = c - base                               # restarts from its own increment at a reset
                del inc, c, base
            else:
                v = torch.floor(40.0 + 40.0 * beta(2.0, 2.0, m))   # integer-valued (temperature / util style)
            stale = torch.rand((m, L), generator=g, device=device) < 1e-3
            blk[idx] = torch.where(stale, torch.full_like(v, float("nan")), v)
            del v, stale
    if layout and zero_nodes:                              # exercises cluster_monitor.py:174
        nodes = min(cfg["nodes"], n // cfg.get("per_node", 4))
        per = cfg.get("per_node", 4)
        z = (torch.rand(nodes, generator=g, device=device) < 0.02).nonzero().flatten().tolist()
        for node in z:
            view[per * node] = 0.0
            view[per * node + 3] = 0.0
    offsets = torch.arange(n + 1, dtype=torch.int64, device=device) * L
    return values, offsets, kinds
