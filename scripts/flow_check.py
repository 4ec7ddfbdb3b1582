"""Development check of series_flow on a GPU box: parity on small batches of every kind against the C oracle,
how many series needed the second attempt / the slow path, and device-resident speed at 7-day and 30-day length.
Usage: python scripts/flow_check.py [parity] [speed]"""
import sys
import time

import numpy as np

sys.path.insert(0, ".")
import torch
from kubeoperator_b200 import Rollup, synth
from oracle import c_oracle

QS = [0.5, 0.95, 0.99]
what = set(sys.argv[1:]) or {"parity", "speed"}
r = Rollup(0)


def bits(a):
    return np.ascontiguousarray(a, np.float64).view(np.uint64)


def check(name, values, offsets, qs):
    n = len(offsets) - 1
    dv, do = torch.from_numpy(values).cuda(), torch.from_numpy(offsets).cuda()
    outs = {k: torch.zeros(n, dtype=torch.float64, device="cuda") for k in ("sum", "mean", "min", "max", "last")}
    cnt = torch.zeros(n, dtype=torch.int64, device="cuda")
    qt = torch.full((n, max(1, len(qs))), -7.0, dtype=torch.float64, device="cuda")
    t0 = time.time()
    r.series_rollup_dev(dv.data_ptr(), do.data_ptr(), n, qs, outs["sum"].data_ptr(), outs["mean"].data_ptr(),
                        outs["min"].data_ptr(), outs["max"].data_ptr(), outs["last"].data_ptr(), cnt.data_ptr(),
                        qt.data_ptr(), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    st = r.series_status()
    exp = c_oracle.series_rollup(values, offsets, qs, nthreads=8)
    bad = []
    if not np.array_equal(cnt.cpu().numpy(), exp["count"]):
        bad.append("count")
    for k in ("min", "max", "last"):
        g, e = outs[k].cpu().numpy(), exp[k]
        both = np.isnan(g) & np.isnan(e)
        if not np.array_equal(bits(g)[~both], bits(e)[~both]) or not np.array_equal(np.isnan(g), np.isnan(e)):
            bad.append(k)
    if len(qs):
        g, e = qt.cpu().numpy()[:, :len(qs)], exp["quantiles"]
        both = np.isnan(g) & np.isnan(e)
        wrong = (bits(g) != bits(e)) & ~both
        if wrong.any():
            bad.append("quantiles@%s" % np.argwhere(wrong)[:6].tolist())
    g, e = outs["sum"].cpu().numpy(), exp["sum"]
    absum = np.array([np.nansum(np.abs(values[offsets[i]:offsets[i + 1]])) for i in range(n)])
    fin = np.isfinite(e)
    if not np.all(np.abs(g - e)[fin] <= 1e-12 * absum[fin]):
        bad.append("sum")
    print("%-28s n=%5d %s  status=%s  (%.1fs)" % (name, n, "OK" if not bad else "MISMATCH " + str(bad), st,
                                                   time.time() - t0), flush=True)
    return not bad


if "parity" in what:
    ok = True
    d = synth.make_series("c2", seed=3, max_series=48)
    ok &= check("c2 first 48 (gauge/const)", d["values"], d["offsets"], QS)
    d = synth.make_series(dict(nodes=2, gpus=2, pods=4, L=40320), seed=4)
    ok &= check("7-day all kinds", d["values"], d["offsets"], QS)
    d = synth.make_series(dict(nodes=2, gpus=1, pods=3, L=172800), seed=5)
    ok &= check("30-day all kinds", d["values"], d["offsets"], QS)
    d = synth.make_series("c1", seed=1)
    ok &= check("c1 (24 h)", d["values"], d["offsets"], QS)
    ok &= check("c1 stats only", d["values"], d["offsets"], [])
    rg = synth.make_ragged(2, [0, 1, 20, 240, 1024, 1025, 1026, 1535, 1537, 2048, 5000, 40320, 40321, 70001])
    ok &= check("ragged", rg["values"], rg["offsets"], QS)
    rg = synth.make_ragged(3, [1025, 1027, 3001, 9999])
    v = np.concatenate([[0.25], rg["values"]])
    ok &= check("ragged, odd alignment", v, rg["offsets"] + 1, [0.0, 0.5, 1.0])
    for seed in range(8):
        s = synth.make_stress(seed)
        ok &= check("stress %d q=%s" % (seed, s["qs"]), s["values"], s["offsets"], s["qs"])
    print("PARITY", "OK" if ok else "FAILED", flush=True)

if "speed" in what:
    for name, cfg, n in (("7-day c2", "c2", None), ("30-day c3 slice", "c3", 5920)):
        vals, offs, kinds = synth.make_series_torch(cfg, 1, "cuda", kinds=None if n is None else synth.mix_kinds(cfg, n))
        n = len(kinds)
        q = torch.zeros((n, 3), dtype=torch.float64, device="cuda")
        s = torch.zeros(n, dtype=torch.float64, device="cuda")
        for it in range(4):
            r.series_rollup_dev(vals.data_ptr(), offs.data_ptr(), n, QS, s.data_ptr(), 0, 0, 0, 0, 0, q.data_ptr(),
                                torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        ms, pms = r.last_series_kernel_ms(), r.last_series_pipe_ms()
        gb = vals.numel() * 8 / 1e9
        print("%-16s n=%d %.2f GB: all long kernels %.3f ms = %.0f GB/s; series_flow %.3f ms = %.0f GB/s; status %s"
              % (name, n, gb, ms, gb / ms * 1e3, pms, gb / pms * 1e3, r.series_status()), flush=True)
        del vals, offs, q, s
        torch.cuda.empty_cache()
