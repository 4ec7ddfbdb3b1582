"""Randomised parity of 30-day-class series (100 000 .. 200 000 samples, every kind, random stale fractions, odd
alignments, random quantile sets) against the C oracle: python scripts/stress_long.py first_seed n_seeds"""
import sys
sys.path.insert(0, ".")
sys.path.insert(0, "tests")
import numpy as np
from kubeoperator_b200 import Rollup, synth
from oracle import c_oracle
import test_gpu_series
from test_gpu_series import assert_series_parity

first = int(sys.argv[1]) if len(sys.argv) > 1 else 0
count = int(sys.argv[2]) if len(sys.argv) > 2 else 5
QSETS = ([0.5, 0.95, 0.99], [0.01, 0.9, 0.05], [0.0, 1.0, 0.25, 0.75], [0.999, 0.001], [0.5], [])
r = Rollup(0)
for seed in range(first, first + count):
    rng = np.random.default_rng(10_000 + seed)
    parts = [rng.random(int(rng.integers(0, 4)))]                    # odd alignment of what follows
    for _ in range(int(rng.integers(24, 48))):
        L = int(rng.integers(100_000, 200_001))
        v = synth._one_series(rng, int(rng.integers(0, 4)), L)
        mode = rng.integers(0, 6)
        if mode == 0:
            v[rng.random(L) < rng.choice([0.3, 0.9, 0.99])] = np.nan    # mostly stale
        elif mode == 1:
            v = np.round(v, 1)                                            # quantised
        elif mode == 2:
            v[: L // 2] = v[0]                                            # long constant run
        parts.append(v)
        if rng.random() < 0.3:
            parts.append(rng.random(int(rng.integers(0, 2000))))         # a short series in between
    values = np.concatenate(parts)
    offsets = np.concatenate([[0], np.cumsum([len(p) for p in parts])]).astype(np.int64)
    qs = QSETS[seed % len(QSETS)]
    # Sum tolerance: the oracle adds left to right like the reference's loops, and over a run of 10^5 equal samples
    # that sum itself drifts from the exact one by up to n * 2^-53 relative (seed 19: oracle 1.3e-12 off, GPU sum
    # equal to math.fsum to the last place).  So: the a-priori bound of the oracle's own error, not less than 1e-12.
    test_gpu_series.SUM_RTOL = max(1e-12, 200_000 * 2.0 ** -53)
    got = r.series_rollup(values, offsets, qs)
    exp = c_oracle.series_rollup(values, offsets, qs, nthreads=16)
    try:
        assert_series_parity(got, exp, values, offsets, qs)
    except AssertionError:
        absum = np.array([np.nansum(np.abs(values[offsets[i]:offsets[i + 1]])) for i in range(len(parts))])
        rel = np.abs(got["sum"] - exp["sum"]) / np.maximum(absum, 1e-300)
        i = int(np.nanargmax(rel))
        v = values[offsets[i]:offsets[i + 1]]
        import math
        print("seed", seed, "worst sum: series", i, "L", len(v), "rel", rel[i], "got", got["sum"][i], "exp", exp["sum"][i],
              "fsum", math.fsum(v[~np.isnan(v)]), "count", got["count"][i], exp["count"][i], "head", v[:4])
        raise
    print("seed", seed, "ok: n=%d samples=%d qs=%s status=%s" % (len(parts), len(values), qs, r.counters()["series_fallback"]), flush=True)
print("all ok")
