"""Randomised parity of the no-allocation entry point (ko_series_rollup_static: small spill space, so more series take
the second attempt and the exact multi-pass kernel) against the C oracle: python scripts/stress_static.py first n"""
import sys
sys.path.insert(0, ".")
sys.path.insert(0, "tests")
import numpy as np
import torch
from kubeoperator_b200 import korollup, synth
from oracle import c_oracle
from test_gpu_series import assert_series_parity

first = int(sys.argv[1]) if len(sys.argv) > 1 else 0
count = int(sys.argv[2]) if len(sys.argv) > 2 else 10
for seed in range(first, first + count):
    d = synth.make_stress(seed)
    if seed % 5 == 0:      # now and then a 30-day series of every kind in the batch
        rng = np.random.default_rng(seed)
        extra = [synth._one_series(rng, k, 172800) for k in range(4)]
        d["values"] = np.concatenate([d["values"]] + extra)
        d["offsets"] = np.concatenate([d["offsets"], d["offsets"][-1] + np.cumsum([len(e) for e in extra])]).astype(np.int64)
    n = len(d["offsets"]) - 1
    qs = d["qs"]
    dv, do = torch.from_numpy(d["values"]).cuda(), torch.from_numpy(d["offsets"]).cuda()
    o = {k: torch.full((n,), -7.0, dtype=torch.float64, device="cuda") for k in ("sum", "mean", "min", "max", "last")}
    cnt = torch.full((n,), -7, dtype=torch.int64, device="cuda")
    qt = torch.full((n, max(1, len(qs))), -7.0, dtype=torch.float64, device="cuda")
    korollup.series_rollup_static(0, dv.data_ptr(), do.data_ptr(), n, qs, o["sum"].data_ptr(), o["mean"].data_ptr(),
                                  o["min"].data_ptr(), o["max"].data_ptr(), o["last"].data_ptr(), cnt.data_ptr(),
                                  qt.data_ptr(), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    got = {k: v.cpu().numpy() for k, v in o.items()}
    got["count"] = cnt.cpu().numpy()
    got["quantiles"] = qt.cpu().numpy()[:, :len(qs)]
    exp = c_oracle.series_rollup(d["values"], d["offsets"], qs, nthreads=8)
    assert_series_parity(got, exp, d["values"], d["offsets"], qs)
    print("seed", seed, "ok: n=%d samples=%d qs=%s" % (n, len(d["values"]), qs), flush=True)
print("all ok")
