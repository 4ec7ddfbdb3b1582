"""ko_series_rollup_static through the C ABI against ko_series_rollup_dev.

The static entry point exists for callers that cannot allocate GPU memory when they need a rollup (their
memory is owned by another component's pool, or they must avoid the device-wide synchronisation of
cudaMalloc/cudaFree).  It runs the same kernels with the work list and counters in a fixed array inside the
library, so its outputs must equal the regular device-pointer entry point's: bit for bit for count, min, max,
last and quantiles, and within the 1e-12 sum tolerance of the other series tests (in fact equal, the
reduction order does not depend on where the scratch lives).  Two calls in a row on DIFFERENT batches check
that nothing left in the shared array by the first call (work list, cursor, flags) leaks into the second.
This file sorts last on purpose: a run that stops at the first failure has then run everything else."""
import numpy as np
import pytest

from kubeoperator_b200 import synth

pytestmark = pytest.mark.gpu
QS = [0.5, 0.95, 0.99]


def _run_both(rollup, values, offsets, qs):
    import torch
    from kubeoperator_b200 import korollup
    n = len(offsets) - 1
    dv = torch.from_numpy(values).cuda()
    do = torch.from_numpy(offsets).cuda()
    stream = torch.cuda.current_stream().cuda_stream
    res = []
    for which in ("dev", "static"):
        o = {k: torch.full((n,), -7.0, dtype=torch.float64, device="cuda") for k in ("sum", "mean", "min", "max", "last")}
        cnt = torch.full((n,), -7, dtype=torch.int64, device="cuda")
        qt = torch.full((n, len(qs)), -7.0, dtype=torch.float64, device="cuda")
        args = (dv.data_ptr(), do.data_ptr(), n, qs, o["sum"].data_ptr(), o["mean"].data_ptr(), o["min"].data_ptr(),
                o["max"].data_ptr(), o["last"].data_ptr(), cnt.data_ptr(), qt.data_ptr(), stream)
        if which == "dev":
            rollup.series_rollup_dev(*args)
        else:
            korollup.series_rollup_static(0, *args)
        torch.cuda.synchronize()
        out = {k: v.cpu().numpy() for k, v in o.items()}
        out["count"], out["quantiles"] = cnt.cpu().numpy(), qt.cpu().numpy()
        res.append(out)
    return res


def _same(a, b):
    return np.array_equal(np.ascontiguousarray(a).view(np.uint64), np.ascontiguousarray(b).view(np.uint64))


def test_static_entry_point_matches_dev_entry_point_twice():
    from kubeoperator_b200 import KoRollupError, Rollup, korollup
    rollup = Rollup(0)
    try:
        # call 1: many long series + short ones (fills the shared work list); call 2: a different, smaller batch
        a = synth.make_ragged(31, [5760] * 40 + [20, 240, 0, 1024, 1025, 40320, 7, 172800])
        b = synth.make_ragged(32, [33, 2049, 12345, 0, 1, 8192])
        for d in (a, b):
            dev, sta = _run_both(rollup, d["values"], d["offsets"], QS)
            assert np.array_equal(dev["count"], sta["count"])
            for k in ("min", "max", "last", "quantiles"):
                assert _same(dev[k], sta[k]), k
            absum = np.array([np.nansum(np.abs(d["values"][d["offsets"][i]:d["offsets"][i + 1]]))
                              for i in range(len(d["offsets"]) - 1)])
            assert np.all(np.abs(dev["sum"] - sta["sum"]) <= 1e-12 * absum)
            assert np.all(np.abs(dev["mean"] - sta["mean"]) <= 1e-12 * absum)
        # a batch beyond the fixed scratch is refused with a clear message, not truncated
        with pytest.raises(KoRollupError) as e:
            korollup.series_rollup_static(0, 8, 8, korollup.STATIC_MAX_SERIES + 1, QS)
        assert "KO_STATIC_MAX_SERIES" in str(e.value)
    finally:
        rollup.close()
