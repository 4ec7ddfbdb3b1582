"""What the fleet step costs on one GPU without the collective: the local rollup alone against the local rollup
followed by index_select + fleet_quantile_device (dist=None).  Run on the GPU box: python scripts/diag_fleet.py"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from kubeoperator_b200 import Rollup, fleet, synth  # noqa: E402

QS = [0.5, 0.95, 0.99]


def main():
    dev = torch.device("cuda:0")
    r, fr = Rollup(0), Rollup(0)
    d_values, d_offsets, kinds = synth.make_series_torch("c2", 1, dev)
    n = len(kinds)
    outs = [torch.zeros(n, dtype=torch.float64, device=dev) for _ in range(5)]
    cnt = torch.zeros(n, dtype=torch.int64, device=dev)
    qt = torch.zeros((n, 3), dtype=torch.float64, device=dev)
    node_idx = torch.arange(0, 400, 4, device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    ws = {}

    def local():
        r.series_rollup_dev(d_values.data_ptr(), d_offsets.data_ptr(), n, QS, *[o.data_ptr() for o in outs],
                            cnt.data_ptr(), qt.data_ptr(), stream)

    def with_fleet(mult):
        local()
        nv = outs[1].index_select(0, node_idx)
        if mult > 1:
            nv = nv.repeat(mult)
        fleet.fleet_quantile_device(nv, 0.95, nv.numel(), fr, None, ws)

    def timed(fn, reps=30):
        for _ in range(5):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps

    print("local rollup only        %.3f ms" % timed(local))
    for mult in (1, 2, 8):
        print("+ fleet step, %4d values %.3f ms" % (100 * mult, timed(lambda: with_fleet(mult))))

    def only_fleet():
        nv = outs[1].index_select(0, node_idx)
        fleet.fleet_quantile_device(nv, 0.95, nv.numel(), fr, None, ws)
    print("fleet step alone (launch-bound) %.3f ms" % timed(only_fleet))


if __name__ == "__main__":
    main()
