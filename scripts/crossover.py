"""Config 5: where does the GPU path overtake the CPU for SHORT windows?

For 1k .. 1M series of 20 and 240 samples (5 min / 1 h at 15 s) this times
  * the C oracle with OpenMP on all host cores (the CPU path),
  * ko_series_rollup with host buffers (H2D + kernels + D2H: what a caller with host data pays),
  * ko_series_rollup_dev with the samples already in HBM,
and writes profiles/<tag>_crossover.json.  Run on the GPU box:  python scripts/crossover.py r01
"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from kubeoperator_b200 import Rollup  # noqa: E402
from oracle import c_oracle  # noqa: E402

QS = [0.5, 0.95, 0.99]
TAG = sys.argv[1] if len(sys.argv) > 1 else "r01"


def best_of(fn, reps):
    best = 1e30
    for _ in range(reps):
        t0 = time.perf_counter()
        fn()
        best = min(best, time.perf_counter() - t0)
    return best


def main():
    c_oracle.build()
    cores = os.cpu_count() or 1
    r = Rollup(0)
    rng = np.random.default_rng(5)
    rows = []
    for L in (20, 240):
        for n in (1000, 10000, 100000, 1000000):
            values = rng.beta(2.0, 5.0, n * L)
            values[rng.random(n * L) < 1e-3] = np.nan
            offsets = np.arange(n + 1, dtype=np.int64) * L
            r.pin(values)
            cpu = best_of(lambda: c_oracle.series_rollup(values, offsets, QS, nthreads=cores), 3)
            got = r.series_rollup(values, offsets, QS)      # warm (buffers, first-touch)
            exp = c_oracle.series_rollup(values, offsets, QS, nthreads=cores)
            assert np.array_equal(got["count"], exp["count"])
            assert np.array_equal(got["quantiles"].view(np.uint64), exp["quantiles"].view(np.uint64))
            host = best_of(lambda: r.series_rollup(values, offsets, QS), 5)
            dv, do = torch.from_numpy(values).cuda(), torch.from_numpy(offsets).cuda()
            outs = [torch.zeros(n, dtype=torch.float64, device="cuda") for _ in range(5)]
            cnt = torch.zeros(n, dtype=torch.int64, device="cuda")
            qt = torch.zeros((n, 3), dtype=torch.float64, device="cuda")
            stream = torch.cuda.current_stream().cuda_stream

            def dev():
                r.series_rollup_dev(dv.data_ptr(), do.data_ptr(), n, QS, *[o.data_ptr() for o in outs],
                                    cnt.data_ptr(), qt.data_ptr(), stream)
                torch.cuda.synchronize()
            dev()
            devt = best_of(dev, 5)
            r.unpin(values)
            rows.append({"samples_per_series": L, "n_series": n, "cpu_ms": cpu * 1e3, "cpu_threads": cores,
                         "gpu_host_buffers_ms": host * 1e3, "gpu_device_resident_ms": devt * 1e3,
                         "speedup_host_buffers": cpu / host, "speedup_device_resident": cpu / devt})
            print(rows[-1], flush=True)
            del dv, do, outs, cnt, qt
    out = {"what": "config 5 crossover: window statistics + p50/p95/p99 of short windows; CPU = C oracle, OpenMP",
           "rows": rows}
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(out, open(os.path.join(ROOT, "gpurun_out", "%s_crossover.json" % TAG), "w"), indent=1)


if __name__ == "__main__":
    main()
