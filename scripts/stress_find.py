"""Find and describe the first failing seed of the randomised stress: python scripts/stress_find.py first count"""
import sys
sys.path.insert(0, ".")
sys.path.insert(0, "tests")
import numpy as np
from kubeoperator_b200 import Rollup, synth
from oracle import c_oracle

first = int(sys.argv[1]); count = int(sys.argv[2])
r = Rollup(0)
for seed in range(first, first + count):
    d = synth.make_stress(seed)
    got = r.series_rollup(d["values"], d["offsets"], d["qs"])
    exp = c_oracle.series_rollup(d["values"], d["offsets"], d["qs"], nthreads=8)
    if not len(d["qs"]):
        continue
    g, e = got["quantiles"], exp["quantiles"]
    both = np.isnan(g) & np.isnan(e)
    bad = (np.ascontiguousarray(g).view(np.uint64) != np.ascontiguousarray(e).view(np.uint64)) & ~both
    if bad.any():
        print("seed", seed, "qs", d["qs"], "status", r.series_status() if hasattr(r, "series_status") else None)
        for i, j in np.argwhere(bad)[:5]:
            v = d["values"][d["offsets"][i]:d["offsets"][i + 1]]
            ok = v[~np.isnan(v)]
            print(" series", i, "L", len(v), "valid", len(ok), "col", j, "q", d["qs"][j], "got", g[i, j], "exp", e[i, j],
                  "all got", g[i], "all exp", e[i])
            print("  distinct", len(np.unique(ok)), "min", ok.min() if len(ok) else None, "max", ok.max() if len(ok) else None,
                  "head", v[:8], "n_inf", int(np.isinf(ok).sum()), "n_zero", int((ok == 0).sum()), "neg zero", int(np.signbit(ok[ok == 0]).sum()))
            np.save("gpurun_out/stress_fail_series.npy", v)
        break
else:
    print("none failed")
