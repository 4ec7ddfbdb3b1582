"""Step time of the device-resident rollup on config 2 for several data seeds (one GPU)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from kubeoperator_b200 import Rollup, synth  # noqa: E402

QS = [0.5, 0.95, 0.99]
dev = torch.device("cuda:0")
r = Rollup(0)
for seed in (1, 2, 3, 4, 1):
    d_values, d_offsets, kinds = synth.make_series_torch("c2", seed, dev)
    n = len(kinds)
    outs = [torch.zeros(n, dtype=torch.float64, device=dev) for _ in range(5)]
    cnt = torch.zeros(n, dtype=torch.int64, device=dev)
    qt = torch.zeros((n, 3), dtype=torch.float64, device=dev)
    stream = torch.cuda.current_stream().cuda_stream

    def local():
        r.series_rollup_dev(d_values.data_ptr(), d_offsets.data_ptr(), n, QS, *[o.data_ptr() for o in outs],
                            cnt.data_ptr(), qt.data_ptr(), stream)
    for _ in range(5):
        local()
    torch.cuda.synchronize()
    c0 = r.counters()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        local()
    e1.record()
    torch.cuda.synchronize()
    c1 = r.counters()
    diff = {k: (c1[k] - c0[k]) / 20 for k in c1 if isinstance(c1[k], (int, float)) and c1[k] != c0[k]}
    print("seed %d  %.3f ms  pipe %.3f ms  all %.3f ms  %s" % (seed, e0.elapsed_time(e1) / 20, r.last_series_pipe_ms(),
                                                    r.last_series_kernel_ms(), diff), flush=True)
    del d_values, d_offsets, outs, cnt, qt
