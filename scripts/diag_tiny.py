"""Small device-resident rollup for sanitizer runs: a few long series of each kind."""
import sys
import numpy as np
sys.path.insert(0, ".")
from kubeoperator_b200 import Rollup, synth
from oracle import c_oracle
r = Rollup(0)
rng = np.random.default_rng(3)
L = 40320
vals = np.concatenate([synth._one_series(rng, i % 4, L) for i in range(16)])
offs = np.arange(17, dtype=np.int64) * L
got = r.series_rollup(vals, offs, [0.5, 0.95, 0.99])
exp = c_oracle.series_rollup(vals, offs, [0.5, 0.95, 0.99], nthreads=4)
assert np.array_equal(got["quantiles"].view(np.uint64), exp["quantiles"].view(np.uint64))
print("tiny ok")
