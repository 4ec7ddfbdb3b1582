"""Randomised parity stress (GPU): synth.make_stress batches against the C oracle.
Usage: python scripts/stress_parity.py [first_seed] [n_seeds]   (640 seeds passed on the round-1 build)"""
import sys
sys.path.insert(0, ".")
sys.path.insert(0, "tests")
from kubeoperator_b200 import Rollup, synth
from oracle import c_oracle
from test_gpu_series import assert_series_parity

first = int(sys.argv[1]) if len(sys.argv) > 1 else 0
count = int(sys.argv[2]) if len(sys.argv) > 2 else 10
r = Rollup(0)
for seed in range(first, first + count):
    d = synth.make_stress(seed)
    got = r.series_rollup(d["values"], d["offsets"], d["qs"])
    exp = c_oracle.series_rollup(d["values"], d["offsets"], d["qs"], nthreads=8)
    assert_series_parity(got, exp, d["values"], d["offsets"], d["qs"])
    print("seed", seed, "ok: n=%d samples=%d qs=%s fallback=%d" % (len(d["offsets"]) - 1, len(d["values"]), d["qs"],
                                                                   r.counters()["series_fallback"]), flush=True)
print("all ok")
