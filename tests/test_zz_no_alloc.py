"""ko_series_rollup_static through the C ABI: against the oracle, against ko_series_rollup_dev, and against
cudaMemGetInfo.

The static entry point exists for callers that cannot allocate GPU memory when they need a rollup (their
memory is owned by another component's pool, or they must avoid the device-wide synchronisation of
cudaMalloc/cudaFree).  It runs the same kernels with the work lists, counters and spill space in a fixed array
inside the library, so
  * all seven outputs (sum, mean, min, max, last, count, quantiles) must equal the oracle's on a ragged batch -
    bit for bit for count, min, max, last and quantiles, within 1e-12 * sum|x| for sum and mean;
  * they must equal the regular device-pointer entry point's;
  * the free device memory cudaMemGetInfo reports must be the same before and after the calls;
  * two calls in a row on DIFFERENT batches must not see each other's leftovers (work lists, cursors, flags);
  * a batch beyond KO_STATIC_MAX_SERIES must be refused with a message that says so.
This file sorts last on purpose: a run that stops at the first failure has then run everything else, and the
test must pass both alone and after every other test has used the GPU."""
import numpy as np
import pytest

from kubeoperator_b200 import synth

pytestmark = pytest.mark.gpu
QS = [0.5, 0.95, 0.99]
KEYS = ("sum", "mean", "min", "max", "last")


def _bits(a):
    return np.ascontiguousarray(a, np.float64).view(np.uint64)


def _same(a, b):
    return np.array_equal(_bits(a), _bits(b))


class _Batch:
    """Inputs and poisoned outputs of one call, all on the device BEFORE any free-memory reading is taken."""

    def __init__(self, d, qs):
        import torch
        self.values, self.offsets, self.qs = d["values"], d["offsets"], qs
        self.n = len(self.offsets) - 1
        self.dv = torch.from_numpy(self.values).cuda()
        self.do = torch.from_numpy(self.offsets).cuda()
        self.out = {}
        for which in ("dev", "static"):
            o = {k: torch.full((self.n,), -7.0, dtype=torch.float64, device="cuda") for k in KEYS}
            o["count"] = torch.full((self.n,), -7, dtype=torch.int64, device="cuda")
            o["quantiles"] = torch.full((self.n, len(qs)), -7.0, dtype=torch.float64, device="cuda")
            self.out[which] = o

    def args(self, which, stream):
        o = self.out[which]
        return (self.dv.data_ptr(), self.do.data_ptr(), self.n, self.qs, o["sum"].data_ptr(), o["mean"].data_ptr(),
                o["min"].data_ptr(), o["max"].data_ptr(), o["last"].data_ptr(), o["count"].data_ptr(),
                o["quantiles"].data_ptr(), stream)

    def host(self, which):
        return {k: v.cpu().numpy() for k, v in self.out[which].items()}

    def check(self):
        from oracle import c_oracle
        exp = c_oracle.series_rollup(self.values, self.offsets, self.qs)
        dev, sta = self.host("dev"), self.host("static")
        absum = np.array([np.nansum(np.abs(self.values[self.offsets[i]:self.offsets[i + 1]])) for i in range(self.n)])
        # the static entry point against the oracle: all seven outputs
        assert np.array_equal(sta["count"], exp["count"])
        for k in ("min", "max", "last", "quantiles"):
            assert _same(sta[k], exp[k]), k
        for k in ("sum", "mean"):
            nan = np.isnan(exp[k])
            assert np.array_equal(np.isnan(sta[k]), nan), k
            tol = 1e-12 * (absum if k == "sum" else absum / np.maximum(exp["count"], 1))
            assert np.all(np.abs(sta[k] - exp[k])[~nan] <= tol[~nan]), k
        # ... and against the regular device-pointer entry point: the same kernels, so the same bits
        assert np.array_equal(sta["count"], dev["count"])
        for k in KEYS + ("quantiles",):
            assert _same(sta[k], dev[k]), k


def test_static_entry_point_oracle_parity_no_allocation_two_calls():
    import torch
    from kubeoperator_b200 import KoRollupError, Rollup, korollup
    rollup = Rollup(0)
    try:
        stream = torch.cuda.current_stream().cuda_stream
        # batch 1: many long series + short ones + empty ones (fills the shared work lists); batch 2: different, smaller
        a = _Batch(synth.make_ragged(31, [5760] * 40 + [20, 240, 0, 1024, 1025, 40320, 7, 172800, 0, 3001]), QS)
        b = _Batch(synth.make_ragged(32, [33, 2049, 12345, 0, 1, 8192]), QS)
        # warm-up on a third batch: the first launch of a kernel loads its code (CUDA loads modules lazily), and the
        # first static call on a device looks up the array's address; neither belongs to the steady state measured below
        w = _Batch(synth.make_ragged(33, [100, 5000, 41000]), QS)
        korollup.series_rollup_static(0, *w.args("static", stream))
        rollup.series_rollup_dev(*w.args("dev", stream))
        torch.cuda.synchronize()
        w.check()

        free0, total0 = torch.cuda.mem_get_info(0)          # cudaMemGetInfo
        korollup.series_rollup_static(0, *a.args("static", stream))
        korollup.series_rollup_static(0, *b.args("static", stream))
        torch.cuda.synchronize()
        free1, total1 = torch.cuda.mem_get_info(0)
        assert (free1, total1) == (free0, total0), "ko_series_rollup_static changed the free device memory"

        rollup.series_rollup_dev(*a.args("dev", stream))
        torch.cuda.synchronize()
        rollup.series_rollup_dev(*b.args("dev", stream))
        torch.cuda.synchronize()
        a.check()
        b.check()

        # a batch beyond the fixed scratch is refused with a clear message, not truncated
        with pytest.raises(KoRollupError) as e:
            korollup.series_rollup_static(0, 8, 8, korollup.STATIC_MAX_SERIES + 1, QS)
        assert "KO_STATIC_MAX_SERIES" in str(e.value)
        free2, _ = torch.cuda.mem_get_info(0)
        assert free2 == free1
    finally:
        rollup.close()
