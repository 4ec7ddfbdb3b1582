"""Fleet-wide merges on the device (kubeoperator_b200/fleet.py): exact quantiles by four rounds of
65 536-bin key histograms with the refinement state in device memory (ko_key_histogram_state_dev +
ko_key_select_dev), top-k and additive scalars - against the NumPy oracle, on one GPU and, when the box has two,
over NCCL with one process per GPU."""
import os
import socket

import numpy as np
import pytest

from oracle import rollup_oracle as ro

pytestmark = pytest.mark.gpu


def _bits(a):
    return np.ascontiguousarray(a, np.float64).view(np.uint64)


def _fleet_values(seed, n):
    rng = np.random.default_rng(seed)
    v = np.concatenate([rng.beta(2, 5, n) * 100.0, -rng.gamma(2.0, 3.0, n // 7), np.repeat(42.5, n // 5),
                        [0.0, -0.0, np.inf, -np.inf, 5e-324, -5e-324, 1.7976931348623157e308]])
    v[rng.random(len(v)) < 0.03] = np.nan
    rng.shuffle(v)
    return v


QS = [0.0, 0.01, 0.5, 0.95, 0.99, 1.0]


def test_fleet_quantiles_device_one_gpu_equals_the_oracle():
    import torch
    from kubeoperator_b200 import Rollup, fleet
    with Rollup(0) as r:
        ws = {}
        for seed, n in ((1, 100_000), (2, 1), (3, 37)):
            v = _fleet_values(seed, n)
            for lo in range(0, len(QS), 3):               # up to KO_MAX_QUANTILES per call; the workspace is reused
                qs = QS[lo:lo + 3]
                got = fleet.fleet_quantiles_device(torch.from_numpy(v).cuda(), qs, r, None, ws).cpu().numpy()
                exp = np.array([ro.fleet_quantile([v], q) for q in qs])
                assert np.array_equal(_bits(got), _bits(exp)), (seed, qs, got, exp)
        # no valid value in the whole fleet: NaN, as the oracle says
        v = np.full(1000, np.nan)
        got = fleet.fleet_quantiles_device(torch.from_numpy(v).cuda(), [0.5, 0.99], r, None, ws).cpu().numpy()
        assert np.isnan(got).all() and np.isnan(ro.fleet_quantile([v], 0.5))
        # an empty shard
        got = fleet.fleet_quantiles_device(torch.zeros(0, dtype=torch.float64, device="cuda"), [0.5], r, None, ws)
        assert np.isnan(got.cpu().numpy()).all()


def test_fleet_topk_and_scalars_device_one_gpu():
    import torch
    from kubeoperator_b200 import fleet
    rng = np.random.default_rng(5)
    v = rng.integers(0, 9, 400).astype(np.float64)
    v[rng.random(400) < 0.1] = np.nan
    v[17], v[301] = np.inf, -np.inf                       # infinities are values, only NaN means "no value"
    tv, ti = fleet.fleet_topk_device(torch.from_numpy(v).cuda(), 1000, 10)
    keys = np.where(np.isnan(v), -np.inf, v)
    order = np.argsort(-keys, kind="stable")[:10]         # value descending, index ascending among ties
    assert np.array_equal(ti.cpu().numpy(), order + 1000)
    assert np.array_equal(tv.cpu().numpy(), keys[order])
    # fewer values than k: padded with -inf / -1
    tv, ti = fleet.fleet_topk_device(torch.tensor([np.nan, 3.0, np.nan], dtype=torch.float64, device="cuda"), 0, 4)
    assert ti.cpu().tolist() == [1, -1, -1, -1] and tv.cpu().tolist()[0] == 3.0
    t = fleet.fleet_scalars_device(torch.tensor([1.5, 2.0], dtype=torch.float64, device="cuda"))
    assert t.cpu().tolist() == [1.5, 2.0]


# ---- two processes, two GPUs, NCCL ------------------------------------------------------------------------------
def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _nccl_worker(rank, world, port, q_out):
    import torch
    import torch.distributed as dist
    from kubeoperator_b200 import Rollup, fleet
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    allv = _fleet_values(11, 50_000)
    lo, hi = fleet.shard_range(len(allv), rank, world)
    mine = torch.from_numpy(allv[lo:hi]).cuda()
    res = {}
    with Rollup(rank) as r:
        ws = {}
        res["q"] = fleet.fleet_quantiles_device(mine, [0.5, 0.95, 0.99], r, dist, ws).cpu().numpy()
        res["q2"] = fleet.fleet_quantiles_device(mine, [0.0, 1.0], r, dist, ws).cpu().numpy()
        tv, ti = fleet.fleet_topk_device(mine, lo, 10, dist)
        res["topk_v"], res["topk_i"] = tv.cpu().numpy(), ti.cpu().numpy()
        res["sum"] = fleet.fleet_scalars_device(torch.tensor([float(rank + 1), 2.0], dtype=torch.float64,
                                                             device="cuda"), dist).cpu().tolist()
    if rank == 0:
        q_out.put(res)
    dist.barrier()
    dist.destroy_process_group()


def test_fleet_merges_over_nccl_two_gpus():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs (run with gpurun --gpus 2)")
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_nccl_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = q.get(timeout=300)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    allv = _fleet_values(11, 50_000)
    exp = np.array([ro.fleet_quantile([allv], qv) for qv in (0.5, 0.95, 0.99)])
    assert np.array_equal(_bits(res["q"]), _bits(exp))
    exp2 = np.array([ro.fleet_quantile([allv], qv) for qv in (0.0, 1.0)])
    assert np.array_equal(_bits(res["q2"]), _bits(exp2))
    keys = np.where(np.isnan(allv), -np.inf, allv)
    order = np.argsort(-keys, kind="stable")[:10]         # value descending, index ascending among ties
    assert np.array_equal(res["topk_i"], order)
    assert np.array_equal(res["topk_v"], keys[order])
    assert res["sum"] == [3.0, 4.0]
