"""Fleet-wide merges across series shards (one process per GPU).  BUILD-DEFINED: the
reference has no multi-process rollup; its fleet merge is the serial concat + re-sort of DashBoardView.get
(core/apps/kubeops_api/api.py:495-501) and the browser sums of dashboard.component.ts:117-153.

Per-series statistics need no communication (series are sharded by ID).  Collectives — NCCL on GPUs, gloo
in the CPU tests — are used only for the small fleet-level exchange:
* ``fleet_quantile``: exact nearest-rank quantile of the pooled per-shard values by all-gathering 65 536-bin
  key histograms and refining 16 bits per round (4 rounds for a 64-bit key);
* ``fleet_quantile_small`` / ``fleet_quantile_device``: when every shard holds few values (<= 65 536), ONE
  all-gather of the values themselves and a selection on the host / on the GPU (no read-back per step);
* ``fleet_topk``: all-gather of every shard's top-k (key, global index) and a merge that keeps the
  reference's stable descending order (ties by global index);
* ``fleet_scalars``: all-reduce of the additive fleet scalars.

``hist_fn(shift, prefix_mask, prefix) -> int64[65536]`` supplies the local histogram: on a GPU it is
``Rollup.key_histogram`` / ``key_histogram_dev`` (the CUDA kernel); the CPU tests pass a NumPy stand-in.
"""
import math

import numpy as np

HIST_BINS = 65536


def shard_range(n_items, rank, world):
    """Contiguous balanced shard [lo, hi) of n_items for ``rank``."""
    base, rem = divmod(n_items, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def shard_by_weight(weights, world):
    """Contiguous series ranges balanced by sample count (prefix sum over series lengths)."""
    w = np.asarray(weights, np.int64)
    csum = np.concatenate([[0], np.cumsum(w)])
    total = int(csum[-1])
    cuts = [int(np.searchsorted(csum, total * r / world, side="left")) for r in range(world + 1)]
    cuts[0], cuts[-1] = 0, len(w)
    return [(cuts[r], max(cuts[r], cuts[r + 1])) for r in range(world)]


def _all_gather_sum(t, dist):
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return t
    import torch
    bufs = [torch.empty_like(t) for _ in range(dist.get_world_size())]
    dist.all_gather(bufs, t)                      # per north_star: allgather of per-shard histograms
    return torch.stack(bufs).sum(dim=0)


def key_to_double(key):
    k = np.uint64(key)
    u = (k & np.uint64(0x7FFFFFFFFFFFFFFF)) if (int(k) >> 63) else ~k
    return float(np.array([u], np.uint64).view(np.float64)[0])


def fleet_quantile(hist_fn, n_valid_local, q, dist=None, device="cpu"):
    """Exact nearest-rank quantile over all shards.  Returns NaN when the fleet has no valid value."""
    import torch
    n = torch.tensor([int(n_valid_local)], dtype=torch.int64, device=device)
    n_total = int(_all_gather_sum(n, dist).item())
    if n_total == 0:
        return float("nan")
    r = max(1, min(n_total, int(math.ceil(q * n_total))))
    prefix, mask = 0, 0
    for shift in (48, 32, 16, 0):
        h = hist_fn(shift, mask, prefix)
        if not torch.is_tensor(h):
            h = torch.from_numpy(np.ascontiguousarray(h, np.int64)).to(device)
        h = _all_gather_sum(h, dist)
        c = torch.cumsum(h, 0)
        d = int(torch.searchsorted(c, torch.tensor([r], dtype=torch.int64, device=c.device)).item())
        if d > 0:
            r -= int(c[d - 1].item())
        prefix |= d << shift
        mask |= 0xFFFF << shift
    return key_to_double(prefix) + 0.0


SMALL_FLEET_CAP = 1 << 16


def fleet_quantile_small(values, q, cap, dist=None, device="cpu", workspace=None):
    """Exact nearest-rank quantile over all shards when every shard holds at most ``cap`` values (the same
    ``cap`` on every rank, <= SMALL_FLEET_CAP): ONE all-gather of the NaN-padded values and a selection on the
    host, instead of the four histogram rounds of fleet_quantile (each a collective and a host sync).  NaN =
    no value, as everywhere in the rollup.  Same result as fleet_quantile.  ``workspace``: a dict the caller
    keeps between calls; the send / receive / pinned host buffers are created once in it."""
    import torch
    assert 0 < cap <= SMALL_FLEET_CAP
    v = torch.as_tensor(values, dtype=torch.float64, device=device).reshape(-1)
    assert v.numel() <= cap
    world = dist.get_world_size() if (dist is not None and dist.is_initialized()) else 1
    ws = workspace if workspace is not None else {}
    if ws.get("cap") != cap or ws.get("world") != world:
        ws["cap"], ws["world"] = cap, world
        ws["send"] = torch.empty(cap, dtype=torch.float64, device=device)
        ws["recv"] = torch.empty(world * cap, dtype=torch.float64, device=device) if world > 1 else ws["send"]
        pin = torch.device(device).type == "cuda"
        ws["host"] = torch.empty(world * cap, dtype=torch.float64, pin_memory=pin)
    send, recv, host = ws["send"], ws["recv"], ws["host"]
    send[:v.numel()].copy_(v)
    if v.numel() < cap:
        send[v.numel():].fill_(float("nan"))
    if world > 1:
        dist.all_gather_into_tensor(recv, send)
    host.copy_(recv)                     # one device -> host copy (synchronises)
    allv = host.numpy()
    allv = allv[allv == allv]
    if allv.size == 0:
        return float("nan")
    r = max(1, min(allv.size, int(math.ceil(q * allv.size))))
    return float(np.partition(allv, r - 1)[r - 1]) + 0.0


def fleet_quantile_device(values, q, cap, rollup, dist=None, workspace=None):
    """As fleet_quantile_small, but the selection runs on the GPU and nothing is read back: the gathered,
    NaN-padded values are handed to ``Rollup.series_rollup_dev`` as ONE series (NaN = no value there too), so
    the nearest-rank quantile is the library's own exact order statistic.  Returns a 1-element CUDA tensor
    that is valid once the current stream has run; a caller in a loop reads it when it needs it instead of
    synchronising every step.  ``values``: CUDA float64 tensor, at most ``cap`` entries, same ``cap`` on
    every rank."""
    import torch
    assert 0 < cap <= SMALL_FLEET_CAP and values.is_cuda and values.dtype == torch.float64
    v = values.reshape(-1)
    assert v.numel() <= cap
    world = dist.get_world_size() if (dist is not None and dist.is_initialized()) else 1
    ws = workspace if workspace is not None else {}
    if ws.get("dcap") != cap or ws.get("dworld") != world:
        ws["dcap"], ws["dworld"] = cap, world
        ws["dsend"] = torch.empty(cap, dtype=torch.float64, device=v.device)
        ws["drecv"] = torch.empty(world * cap, dtype=torch.float64, device=v.device) if world > 1 else ws["dsend"]
        ws["doffs"] = torch.tensor([0, world * cap], dtype=torch.int64, device=v.device)
        ws["dq"] = torch.empty(1, dtype=torch.float64, device=v.device)
    send, recv = ws["dsend"], ws["drecv"]
    send[:v.numel()].copy_(v)
    if v.numel() < cap:
        send[v.numel():].fill_(float("nan"))
    if world > 1:
        dist.all_gather_into_tensor(recv, send)
    rollup.series_rollup_dev(recv.data_ptr(), ws["doffs"].data_ptr(), 1, [q], d_quant=ws["dq"].data_ptr(),
                             stream=torch.cuda.current_stream().cuda_stream)
    return ws["dq"]


def fleet_topk(local_keys, local_global_idx, k, dist=None, device="cpu"):
    """Global indices of the fleet's top-k by (key descending, global index ascending) — the order
    quick_sort_pods gives on the concatenated list (api.py:495-500), truncated to k.  ``local_keys`` /
    ``local_global_idx`` are this shard's top-k already in that order."""
    import torch
    keys = torch.full((k,), torch.iinfo(torch.int64).min, dtype=torch.int64, device=device)
    gidx = torch.full((k,), torch.iinfo(torch.int64).max, dtype=torch.int64, device=device)
    m = min(k, len(local_keys))
    if m:
        keys[:m] = torch.as_tensor(np.asarray(local_keys[:m], np.int64), device=device)
        gidx[:m] = torch.as_tensor(np.asarray(local_global_idx[:m], np.int64), device=device)
    pair = torch.stack([keys, gidx])
    if dist is not None and dist.is_initialized() and dist.get_world_size() > 1:
        bufs = [torch.empty_like(pair) for _ in range(dist.get_world_size())]
        dist.all_gather(bufs, pair)
        pair = torch.cat(bufs, dim=1)
    ks, gs = pair[0].cpu().numpy(), pair[1].cpu().numpy()
    keep = gs != np.iinfo(np.int64).max
    ks, gs = ks[keep], gs[keep]
    order = np.lexsort((gs, -ks))
    return gs[order][:k].tolist()


def fleet_scalars(values, dist=None, device="cpu"):
    """All-reduce(sum) of a small vector of additive fleet scalars (f64)."""
    import torch
    t = torch.as_tensor(np.asarray(values, np.float64), device=device)
    if dist is not None and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t)
    return t.cpu().numpy()


# ---- device-side merges: nothing is read back between the steps of a merge (the bench loop stays asynchronous) ----
def fleet_quantiles_device(values, qs, rollup, dist=None, workspace=None):
    """Exact nearest-rank quantiles ``qs`` of the pooled per-shard ``values`` (CUDA float64 tensor, NaN = no
    value), all ranks get the same result: a CUDA float64 tensor of len(qs), valid once the current stream has
    run.  Four rounds of: local 65 536-bin key histogram under the current prefix (``ko_key_histogram_state_dev``),
    ONE all-reduce of the bins of all quantiles over NCCL, digit selection on the GPU (``ko_key_select_dev``).
    The refinement state lives in device memory, so no round waits for the host."""
    import torch
    dev = values.device
    v = values.reshape(-1).contiguous()
    nq = len(qs)
    world = dist.get_world_size() if (dist is not None and dist.is_initialized()) else 1
    ws = workspace if workspace is not None else {}
    if ws.get("fq_qs") != tuple(qs):
        ws["fq_qs"] = tuple(qs)
        ws["fq_state"] = torch.zeros((nq, 4), dtype=torch.int64, device=dev)
        ws["fq_hist"] = torch.zeros((nq, HIST_BINS), dtype=torch.int64, device=dev)
        ws["fq_q"] = torch.tensor(list(qs), dtype=torch.float64, device=dev)
    state, hist, qt = ws["fq_state"], ws["fq_hist"], ws["fq_q"]
    n = (v == v).sum().reshape(1)
    if world > 1:
        dist.all_reduce(n)
    nf = n.double()
    rank = torch.clamp(torch.ceil(qt * nf), min=1.0).minimum(nf).long()      # 0 when the fleet has no valid value
    state.zero_()
    state[:, 2] = rank
    stream = torch.cuda.current_stream().cuda_stream
    for shift in (48, 32, 16, 0):
        for j in range(nq):
            rollup.key_histogram_state_dev(v.data_ptr(), v.numel(), shift, state[j].data_ptr(), hist[j].data_ptr(), stream)
        if world > 1:
            dist.all_reduce(hist)
        for j in range(nq):
            rollup.key_select_dev(hist[j].data_ptr(), shift, state[j].data_ptr(), stream)
    return state[:, 3].view(torch.float64)


def fleet_topk_device(values, global_offset, k, dist=None):
    """Fleet top-k of per-series ``values`` (CUDA float64, NaN = no value) by (value descending, global index
    ascending): every shard's own top-k goes through ONE all-gather and is merged on the GPU.  Returns
    (values[k], global_index[k]) CUDA tensors, padded with -inf / -1 when the fleet holds fewer than k values.
    Shards hold ascending, disjoint ranges of global indices (rank order), so two stable descending sorts - one per
    shard, one over the gathered lists in rank order - give the index-ascending order among equal values."""
    import torch
    dev = values.device
    v = values.reshape(-1)
    v = torch.where(v != v, torch.full_like(v, float("-inf")), v)            # NaN only: +inf stays +inf
    kk = min(k, v.numel())
    if kk:
        sv, si = torch.sort(v, descending=True, stable=True)                  # ties: lower index first
    if kk == k:
        top_v, top_i = sv[:k].contiguous(), si[:k] + int(global_offset)
    else:
        top_v = torch.full((k,), float("-inf"), dtype=torch.float64, device=dev)
        top_i = torch.full((k,), -1, dtype=torch.int64, device=dev)
        if kk:
            top_v[:kk] = sv[:kk]
            top_i[:kk] = si[:kk] + int(global_offset)
    world = dist.get_world_size() if (dist is not None and dist.is_initialized()) else 1
    if world > 1:
        gv = torch.empty(world * k, dtype=torch.float64, device=dev)
        gi = torch.empty(world * k, dtype=torch.int64, device=dev)
        dist.all_gather_into_tensor(gv, top_v)
        dist.all_gather_into_tensor(gi, top_i)
        # (padding entries are -inf / -1 and sort behind every value; a real -inf keeps its index)
        out_v, order = torch.sort(gv, descending=True, stable=True)
        out_v, out_i = out_v[:k], gi[order[:k]]
    else:
        out_v, out_i = top_v, top_i
    # "no value" (NaN, mapped to -inf above) and padding carry index -1
    out_i = torch.where(out_v == float("-inf"), torch.full_like(out_i, -1), out_i)
    return out_v, out_i


def fleet_scalars_device(t, dist=None):
    """All-reduce(sum) of a small CUDA vector of additive fleet scalars, in place; returns it."""
    if dist is not None and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t)
    return t
