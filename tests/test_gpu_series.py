"""GPU parity of the window-statistics path (ko_series_rollup through the C ABI) against the oracle.

Bar (north_star): counts, min/max, last, order statistics bit-exact; sums and means within 1e-12 relative
(relative to sum|x|, the condition the left-to-right oracle sum itself satisfies)."""
import json
import os

import numpy as np
import pytest

from oracle import c_oracle, rollup_oracle as ro
from kubeoperator_b200 import synth

pytestmark = pytest.mark.gpu
QS = [0.5, 0.95, 0.99]
SUM_RTOL = 1e-12


@pytest.fixture(scope="module")
def rollup():
    from kubeoperator_b200 import Rollup
    r = Rollup(0)
    yield r
    r.close()


def _bits(a):
    return np.ascontiguousarray(a, np.float64).view(np.uint64)


def assert_series_parity(got, exp, values, offsets, qs):
    n = len(offsets) - 1
    assert np.array_equal(got["count"], exp["count"])
    for k in ("min", "max", "last"):
        both_nan = np.isnan(got[k]) & np.isnan(exp[k])
        assert np.array_equal(_bits(got[k])[~both_nan], _bits(exp[k])[~both_nan]), k
        assert np.array_equal(np.isnan(got[k]), np.isnan(exp[k])), k
    if len(qs):
        g, e = got["quantiles"], exp["quantiles"]
        assert g.shape == e.shape
        both_nan = np.isnan(g) & np.isnan(e)
        bad = (_bits(g) != _bits(e)) & ~both_nan
        assert not bad.any(), "quantile mismatch at %s: got %s expected %s" % (
            np.argwhere(bad)[:5].tolist(), g[bad][:5], e[bad][:5])
    absum = np.array([np.nansum(np.abs(values[offsets[i]:offsets[i + 1]])) for i in range(n)])
    with np.errstate(invalid="ignore"):
        finite = np.isfinite(exp["sum"])
        assert np.all(np.abs(got["sum"] - exp["sum"])[finite] <= SUM_RTOL * absum[finite] + 0.0)
        assert np.array_equal(np.isnan(got["sum"][~finite]), np.isnan(exp["sum"][~finite]))
        assert np.array_equal(got["sum"][~finite & ~np.isnan(exp["sum"])], exp["sum"][~finite & ~np.isnan(exp["sum"])])
        cnt = np.maximum(exp["count"], 1)
        assert np.all(np.abs(got["mean"] - exp["mean"])[finite] <= SUM_RTOL * absum[finite] / cnt[finite])


def test_golden_vectors(rollup, golden_dir):
    g = json.load(open(os.path.join(golden_dir, "builddefined_series.json")))
    dec = lambda v: float("nan") if v is None else float(v)
    vals, offs = [], [0]
    for s in g["series"]:
        vals += [dec(x) for x in s["values"]]
        offs.append(len(vals))
    values, offsets = np.array(vals), np.array(offs, np.int64)
    got = rollup.series_rollup(values, offsets, g["qs"])
    exp = ro.series_rollup_np(values, offsets, g["qs"])
    assert_series_parity(got, exp, values, offsets, g["qs"])
    for i, s in enumerate(g["series"]):
        for j in range(len(g["qs"])):
            a, b = got["quantiles"][i, j], dec(s["quantiles"][j])
            assert (a != a and b != b) or _bits([a])[0] == _bits([b])[0]
        assert int(got["count"][i]) == s["expect"]["count"]


def test_config1_full(rollup):
    d = synth.make_series("c1", seed=1)
    got = rollup.series_rollup(d["values"], d["offsets"], QS)
    exp = c_oracle.series_rollup(d["values"], d["offsets"], QS, nthreads=8)
    assert_series_parity(got, exp, d["values"], d["offsets"], QS)


def test_config2_slice_long_series(rollup):
    # 48 series of the 100-node config at its full length (40 320 samples): every kind incl. DCGM-like
    d = synth.make_series("c2", seed=2, max_series=16)
    cfg = dict(nodes=0, gpus=0, pods=0, L=40320)
    cfg.update(nodes=2, gpus=2, pods=6)
    e = synth.make_series(cfg, seed=3)
    values = np.concatenate([d["values"], e["values"]])
    offsets = np.concatenate([d["offsets"], d["offsets"][-1] + e["offsets"][1:]])
    got = rollup.series_rollup(values, offsets, QS)
    exp = c_oracle.series_rollup(values, offsets, QS, nthreads=8)
    assert_series_parity(got, exp, values, offsets, QS)


def test_ragged_and_edge_lengths(rollup):
    lengths = [0, 1, 2, 3, 31, 32, 33, 20, 240, 1023, 1024, 1025, 1026, 2047, 2048, 2049, 4097, 8191, 8192, 8193,
               0, 5, 12345, 40321, 7]
    d = synth.make_ragged(4, lengths)
    got = rollup.series_rollup(d["values"], d["offsets"], QS)
    exp = ro.series_rollup_np(d["values"], d["offsets"], QS)
    assert_series_parity(got, exp, d["values"], d["offsets"], QS)
    # odd start offsets: the same batch behind one leading sample shifts every 16-byte alignment
    v2 = np.concatenate([[1.5], d["values"]])
    o2 = np.concatenate([[0], d["offsets"] + 1])
    got = rollup.series_rollup(v2, o2, QS)
    exp = ro.series_rollup_np(v2, o2, QS)
    assert_series_parity(got, exp, v2, o2, QS)


def test_ties_constants_nan_inf(rollup):
    rng = np.random.default_rng(5)
    L = 20000
    series = [
        np.full(L, 16.0),                                          # constant
        np.where(rng.random(L) < 0.5, 0.0, rng.random(L)),         # half the mass on one value
        np.where(rng.random(L) < 0.6, 100.0, 100 * rng.random(L)),  # saturated upper tie group
        np.floor(rng.normal(60, 3, L)),                            # few distinct integers
        np.full(L, np.nan),                                        # all stale
        np.where(rng.random(L) < 0.9, np.nan, rng.random(L)),      # mostly stale
        np.concatenate([np.full(L // 2, -0.0), np.full(L - L // 2, 0.0)]),
        np.where(rng.random(L) < 0.02, np.inf, rng.random(L)),     # +inf above the p99 rank
        np.where(rng.random(L) < 0.6, -np.inf, rng.random(L)),
        np.sort(rng.random(L)),                                    # monotone (counter-like)
        np.sort(rng.random(L))[::-1].copy(),
        rng.normal(0, 1, L) * 10.0 ** rng.integers(-300, 300, L),  # huge dynamic range, mixed sign
        np.concatenate([np.zeros(L - 3), [1.0, 2.0, 3.0]]),        # the sample never sees the tail
    ]
    offsets = np.arange(len(series) + 1, dtype=np.int64) * L
    values = np.concatenate(series)
    for qs in (QS, [0.0, 1.0, 0.001, 0.999]):
        got = rollup.series_rollup(values, offsets, qs)
        exp = ro.series_rollup_np(values, offsets, qs)
        assert_series_parity(got, exp, values, offsets, qs)


def test_few_distinct_values(rollup):
    """Integer-valued and quantised gauges (DCGM temperature / utilisation style): the ranks sit inside heavy tie
    groups, which the kernel answers by counting per listed level instead of capturing candidates."""
    rng = np.random.default_rng(11)
    series = []
    for L in (5000, 40320, 40321, 172801):
        series += [
            np.floor(40.0 + 40.0 * rng.beta(2.0, 2.0, L)),                 # 40 integer levels
            np.floor(100.0 * rng.beta(2.0, 5.0, L)) / 10.0,                # steps of 0.1 (not exact in binary)
            np.floor(rng.normal(0.0, 2.0, L)),                             # negative and positive levels, +-0
            np.where(rng.random(L) < 0.5, -0.0, np.floor(rng.normal(0.0, 1.5, L))),
            np.where(rng.random(L) < 0.97, np.floor(rng.normal(60, 2, L)), rng.random(L) * 100),   # levels + stray values
            np.where(rng.random(L) < 0.03, np.inf, np.floor(6 * rng.random(L))),   # +inf is one of the levels
            np.where(rng.random(L) < 0.3, np.nan, np.floor(4 * rng.random(L))),    # many stale samples
            np.floor(rng.random(L) * 12) * 1e-310,                         # denormal levels
            np.concatenate([np.full(L - 40, 7.0), np.arange(40.0)]),       # levels the pre-sample never sees
        ]
    offsets = np.concatenate([[0], np.cumsum([len(x) for x in series])]).astype(np.int64)
    values = np.concatenate(series)
    for qs in (QS, [0.5], [0.0, 1.0, 0.25, 0.75], [0.9, 0.1]):
        got = rollup.series_rollup(values, offsets, qs)
        exp = c_oracle.series_rollup(values, offsets, qs, nthreads=8)
        assert_series_parity(got, exp, values, offsets, qs)


def test_more_long_series_than_one_plan_window(rollup):
    """20 000 long series (many per CTA of the persistent one-pass kernel, short ones: three stages each), and the
    static entry point over 5 000 of them."""
    import torch
    from kubeoperator_b200 import korollup
    rng = np.random.default_rng(12)
    n, L = 20000, 1100
    values = rng.beta(2.0, 5.0, n * L)
    values[rng.random(n * L) < 1e-3] = np.nan
    values[::7] = np.floor(values[::7] * 50)          # some repeated values
    offsets = np.arange(n + 1, dtype=np.int64) * L
    got = rollup.series_rollup(values, offsets, QS)
    exp = c_oracle.series_rollup(values, offsets, QS, nthreads=8)
    assert_series_parity(got, exp, values, offsets, QS)
    # the no-allocation entry point over 5 000 of them
    m = 5000
    dv, do = torch.from_numpy(values[:m * L]).cuda(), torch.from_numpy(offsets[:m + 1]).cuda()
    qt = torch.zeros((m, 3), dtype=torch.float64, device="cuda")
    cnt = torch.zeros(m, dtype=torch.int64, device="cuda")
    korollup.series_rollup_static(0, dv.data_ptr(), do.data_ptr(), m, QS, 0, 0, 0, 0, 0, cnt.data_ptr(),
                                  qt.data_ptr(), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    assert np.array_equal(cnt.cpu().numpy(), exp["count"][:m])
    assert np.array_equal(_bits(qt.cpu().numpy()), _bits(exp["quantiles"][:m]))


def test_brackets_that_miss_overflow_or_are_skewed(rollup):
    """Series built to defeat the pre-sample (1024 aligned runs of four samples, one every 40 samples here): the
    wanted ranks lie where the sample points say nothing (second attempt with wide intervals, then the slow path),
    candidates pile up in a sliver of the value range, and three separate intervals at once."""
    rng = np.random.default_rng(13)
    L = 40960                                           # sample runs start at 20, 60, 100, ... (rounded to 4)
    idx = np.arange(L)
    on_sample = ((idx // 4) % 10) == 5                  # the run of four that holds position 40 i + 20
    series = [
        np.where(on_sample, rng.random(L), 5.0 + rng.random(L)),          # the sample sees only the low 2.5 %
        np.where(on_sample, 5.0 + rng.random(L), rng.random(L)),          # ... only the high 2.5 %
        np.where(on_sample, 1.0, rng.random(L) * 1e-3),                   # sample is constant, data is not
        np.concatenate([rng.random(L // 2), 1.0 + 1e-9 * rng.random(L - L // 2)]),   # half the mass in a sliver
        np.where(rng.random(L) < 0.5, 0.25 + 1e-12 * rng.integers(0, 300, L), rng.random(L)),   # dense cluster at p50
        np.exp(rng.normal(0, 4, L)),                                      # heavy tail: tail sub-brackets are skewed
        -np.exp(rng.normal(0, 4, L)),
        rng.standard_cauchy(L),
        np.where(rng.random(L) < 0.4, np.floor(rng.random(L) * 3), rng.random(L) * 3),   # levels mixed with noise
    ]
    offsets = np.arange(len(series) + 1, dtype=np.int64) * L
    values = np.concatenate(series)
    for qs in (QS, [0.1, 0.5, 0.99], [0.02, 0.5, 0.98, 0.999], [0.5], [0.999], [0.3, 0.35]):
        got = rollup.series_rollup(values, offsets, qs)
        exp = c_oracle.series_rollup(values, offsets, qs, nthreads=8)
        assert_series_parity(got, exp, values, offsets, qs)


def test_thirty_day_series_every_kind_one_pass(rollup):
    """30 days at 15 s (config 3 / 4 length): every series kind, parity with the oracle, and the one-pass kernel
    answers them itself -- none of them may need the exact multi-pass slow path."""
    d = synth.make_series(dict(nodes=2, gpus=2, pods=3, L=172800), seed=6)
    assert set(d["kinds"].tolist()) == {synth.KIND_GAUGE, synth.KIND_CONST, synth.KIND_COUNTER, synth.KIND_DISCRETE}
    for qs in (QS, [0.5], [0.01, 0.25, 0.75, 0.999]):
        got = rollup.series_rollup(d["values"], d["offsets"], qs)
        exp = c_oracle.series_rollup(d["values"], d["offsets"], qs, nthreads=8)
        assert_series_parity(got, exp, d["values"], d["offsets"], qs)
    import torch
    n = len(d["offsets"]) - 1
    dv, do = torch.from_numpy(d["values"]).cuda(), torch.from_numpy(d["offsets"]).cuda()
    qt = torch.zeros((n, 3), dtype=torch.float64, device="cuda")
    rollup.series_rollup_dev(dv.data_ptr(), do.data_ptr(), n, QS, 0, 0, 0, 0, 0, 0, qt.data_ptr(),
                             torch.cuda.current_stream().cuda_stream)
    st = rollup.series_status()
    assert st["long_series"] == n and st["slow_path"] == 0 and st["error_flag"] == 0
    assert st["second_attempt"] <= 2          # a 3-sigma interval misses its rank now and then


def test_series_the_one_pass_kernel_must_defer(rollup):
    """Series the pre-sample says nothing about (almost only stale samples; a spread below the resolution of the
    32-bit high key) go to the exact slow path: same results, and the device says so."""
    import torch
    rng = np.random.default_rng(21)
    L = 30000
    a = rng.random(L)
    a[rng.random(L) < 0.995] = np.nan                     # ~150 valid samples: the pre-sample sees a handful
    b = 8589934592.0 + rng.integers(0, 1000, L).astype(np.float64)    # byte counts: equal high words, different values
    c = rng.beta(2, 5, L)
    values = np.concatenate([a, b, c])
    offsets = np.array([0, L, 2 * L, 3 * L], np.int64)
    got = rollup.series_rollup(values, offsets, QS)
    exp = c_oracle.series_rollup(values, offsets, QS, nthreads=4)
    assert_series_parity(got, exp, values, offsets, QS)
    dv, do = torch.from_numpy(values).cuda(), torch.from_numpy(offsets).cuda()
    qt = torch.zeros((3, 3), dtype=torch.float64, device="cuda")
    rollup.series_rollup_dev(dv.data_ptr(), do.data_ptr(), 3, QS, 0, 0, 0, 0, 0, 0, qt.data_ptr(),
                             torch.cuda.current_stream().cuda_stream)
    st = rollup.series_status()
    assert st["slow_path"] == 2 and st["error_flag"] == 0
    assert np.array_equal(_bits(qt.cpu().numpy()), _bits(exp["quantiles"]))


def test_device_error_flag_is_reported(rollup):
    """ko_series_rollup_dev cannot look at device offsets on the host: a decreasing pair is skipped on the GPU and
    ko_series_status reports it."""
    import torch
    values = torch.rand(5000, dtype=torch.float64, device="cuda")
    offs = torch.tensor([0, 3000, 2000, 5000], dtype=torch.int64, device="cuda")
    cnt = torch.full((3,), -1, dtype=torch.int64, device="cuda")
    rollup.series_rollup_dev(values.data_ptr(), offs.data_ptr(), 3, [], 0, 0, 0, 0, 0, cnt.data_ptr(), 0,
                             torch.cuda.current_stream().cuda_stream)
    st = rollup.series_status()
    assert st["error_flag"] != 0
    assert cnt.cpu().tolist() == [3000, -1, 3000]          # the bad series is left untouched


def test_short_windows_config5(rollup):
    rng = np.random.default_rng(7)
    for L in (20, 240):
        n = 20000
        values = rng.beta(2, 5, n * L)
        values[rng.random(n * L) < 1e-3] = np.nan
        offsets = np.arange(n + 1, dtype=np.int64) * L
        got = rollup.series_rollup(values, offsets, QS)
        exp = c_oracle.series_rollup(values, offsets, QS, nthreads=8)
        assert_series_parity(got, exp, values, offsets, QS)
        got = rollup.series_rollup(values, offsets, ())        # statistics only
        assert_series_parity(got, exp, values, offsets, ())


def test_short_windows_every_length_ties_and_mixed_quads(rollup):
    """series_small: four series of <= 32 samples side by side in one warp (eight lanes each), one warp per series
    with the histogram select for 33..256 samples (register sort when a histogram bin holds more than 32 keys),
    series_medium above.  Every length 0..300, tie-heavy / constant / NaN-heavy / infinite values, and batches whose
    groups of four consecutive series mix short, medium and long ones."""
    rng = np.random.default_rng(21)
    qsets = ([0.5, 0.95, 0.99], [0.0, 1.0, 0.25, 0.75], [0.5])

    def draw(L, mode):
        if mode == 0:
            v = rng.beta(2, 5, L) * 100.0
        elif mode == 1:
            v = rng.integers(0, 4, L).astype(np.float64)            # few distinct values: bins far above 32 keys
        elif mode == 2:
            v = np.full(L, 7.25)
        elif mode == 3:
            v = rng.normal(0.0, 1e-300, L)                           # tiny magnitudes, both signs, +-0 among them
            v[rng.random(L) < 0.2] = 0.0
            v[rng.random(L) < 0.1] = -0.0
        elif mode == 4:
            v = rng.normal(0.0, 1.0, L)
            v[rng.random(L) < 0.05] = np.inf
            v[rng.random(L) < 0.05] = -np.inf
        else:
            v = np.concatenate([np.full(L - L // 3, 50.0), 50.0 + rng.random(L // 3) * 1e-9])   # one crowded bin
            rng.shuffle(v)
        v[rng.random(L) < (0.6 if mode == 0 and L % 7 == 0 else 0.02)] = np.nan
        return v

    # 1. every length, every mode, uniform batches (all four series of a warp alike)
    for mode in range(6):
        lens = np.repeat(np.arange(0, 301), 2)
        vals = [draw(int(L), mode) for L in lens]
        values = np.concatenate(vals) if len(vals) else np.zeros(0)
        offsets = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
        qs = qsets[mode % 3]
        got = rollup.series_rollup(values, offsets, qs)
        exp = c_oracle.series_rollup(values, offsets, qs, nthreads=8)
        assert_series_parity(got, exp, values, offsets, qs)
    # 2. lengths shuffled: a warp's four series mix <= 32, <= 256, <= 1024 and long ones; all-NaN and empty series
    lens = np.concatenate([rng.integers(0, 33, 400), rng.integers(33, 257, 200), rng.integers(257, 1025, 60),
                           rng.integers(1025, 3000, 12), [0, 0, 0, 1, 1, 32, 32, 33, 256, 257, 1024, 1025]])
    rng.shuffle(lens)
    vals = [draw(int(L), int(rng.integers(0, 6))) for L in lens]
    for i in range(0, len(vals), 37):
        vals[i][:] = np.nan
    values = np.concatenate(vals)
    offsets = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    got = rollup.series_rollup(values, offsets, QS)
    exp = c_oracle.series_rollup(values, offsets, QS, nthreads=8)
    assert_series_parity(got, exp, values, offsets, QS)
    # 3. a batch of 20-sample windows whose count is not a multiple of four (last warp partly empty)
    n, L = 1003, 20
    values = rng.beta(2, 5, n * L)
    offsets = np.arange(n + 1, dtype=np.int64) * L
    got = rollup.series_rollup(values, offsets, QS)
    exp = c_oracle.series_rollup(values, offsets, QS, nthreads=8)
    assert_series_parity(got, exp, values, offsets, QS)


def test_quantile_order_and_subsets(rollup):
    d = synth.make_series("c1", seed=8, max_series=12)
    for qs in ([0.99, 0.5], [0.5], [0.25, 0.25, 0.75, 0.1]):
        got = rollup.series_rollup(d["values"], d["offsets"], qs)
        exp = ro.series_rollup_np(d["values"], d["offsets"], qs)
        assert_series_parity(got, exp, d["values"], d["offsets"], qs)
    got = rollup.series_rollup(d["values"], d["offsets"], QS, want=("count",))
    assert set(got) == {"count", "quantiles"}


def test_chunked_host_pipeline_matches_device_path(rollup):
    """> 256 MB of samples exercises the double-buffered H2D pipeline; properties at full length:
    count/min/max/quantile invariants instead of a CPU oracle over the whole batch."""
    import torch
    n, L = 1100, 40320                                   # 355 MB
    vals, offs, _ = synth.make_series_torch("c2", 9, "cuda", n_series_override=n)
    host = vals.cpu().numpy()
    offsets = offs.cpu().numpy()
    got = rollup.series_rollup(host, offsets, QS)
    # device-pointer entry point on the same data
    outs = {k: torch.zeros(n, dtype=torch.float64, device="cuda") for k in ("sum", "mean", "min", "max", "last")}
    cnt = torch.zeros(n, dtype=torch.int64, device="cuda")
    qt = torch.zeros((n, 3), dtype=torch.float64, device="cuda")
    rollup.series_rollup_dev(vals.data_ptr(), offs.data_ptr(), n, QS, outs["sum"].data_ptr(), outs["mean"].data_ptr(),
                             outs["min"].data_ptr(), outs["max"].data_ptr(), outs["last"].data_ptr(), cnt.data_ptr(),
                             qt.data_ptr(), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    for k in outs:
        assert np.array_equal(_bits(got[k]), _bits(outs[k].cpu().numpy())), k
    assert np.array_equal(got["count"], cnt.cpu().numpy())
    assert np.array_equal(_bits(got["quantiles"]), _bits(qt.cpu().numpy()))
    # oracle on a sample of the series, properties on all
    pick = np.arange(0, n, 37)
    sub_off = np.concatenate([[0], np.cumsum(np.full(len(pick), L))]).astype(np.int64)
    sub_val = np.concatenate([host[offsets[i]:offsets[i + 1]] for i in pick])
    exp = c_oracle.series_rollup(sub_val, sub_off, QS, nthreads=8)
    sub = {k: (v[pick] if k != "quantiles" else v[pick]) for k, v in got.items()}
    assert_series_parity(sub, exp, sub_val, sub_off, QS)
    q = got["quantiles"]
    assert np.all(got["min"] <= q[:, 0]) and np.all(q[:, 0] <= q[:, 1]) and np.all(q[:, 1] <= q[:, 2])
    assert np.all(q[:, 2] <= got["max"])


def test_counters_add_up_over_staged_chunks(rollup):
    """A host-pointer call larger than one 256 MB staging buffer runs chunk by chunk on the same scratch:
    series_fallback (and the device error flag) must cover every chunk, not only the last one."""
    n, L = 820, 40320                                    # 264 MB of ordinary series ...
    vals, offs, _ = synth.make_series_torch("c2", 19, "cuda", n_series_override=n)
    body = vals.cpu().numpy()
    rng = np.random.default_rng(19)
    long30 = rng.beta(2.0, 5.0, 172800)
    long30[rng.random(172800) < 0.997] = np.nan          # ... behind one 30-day series the fast path must defer
    tail = rng.beta(2.0, 5.0, 12 * L)                    # and enough after them to open a second chunk
    values = np.concatenate([long30, body, tail])
    offsets = np.concatenate([[0], 172800 + offs.cpu().numpy(),
                              172800 + n * L + np.arange(1, 13, dtype=np.int64) * L]).astype(np.int64)
    assert values.nbytes > 256 * 1024 * 1024
    got = rollup.series_rollup(values, offsets, QS)
    assert rollup.counters()["series_fallback"] >= 1     # the 30-day series sits in the FIRST chunk
    pick = np.array([0, 1, n // 2, n, n + 12])
    sub_off = np.concatenate([[0], np.cumsum(offsets[pick + 1] - offsets[pick])]).astype(np.int64)
    sub_val = np.concatenate([values[offsets[i]:offsets[i + 1]] for i in pick])
    exp = c_oracle.series_rollup(sub_val, sub_off, QS, nthreads=8)
    sub = {k: v[pick] for k, v in got.items()}
    assert_series_parity(sub, exp, sub_val, sub_off, QS)


@pytest.mark.parametrize("seed", list(range(24)) + [422, 1344])
def test_randomised_mixtures(rollup, seed):
    """synth.make_stress: random kinds, lengths, alignment and quantile sets (scripts/stress_parity.py runs
    the same generator over hundreds of seeds; 4 000 seeds pass on the round-2 build).  Seed 422 is the regression
    case of round 2: a mostly stale series whose p5 lies above its interval while that interval is open below, so
    that the first selection round's bins also covered the candidates of the p90 interval; seed 1344: an
    integer-valued series where a tie level of ANOTHER interval happened to hold exactly as many samples as the bin."""
    d = synth.make_stress(seed)
    got = rollup.series_rollup(d["values"], d["offsets"], d["qs"])
    exp = c_oracle.series_rollup(d["values"], d["offsets"], d["qs"], nthreads=8)
    assert_series_parity(got, exp, d["values"], d["offsets"], d["qs"])


def test_errors_are_loud(rollup):
    from kubeoperator_b200 import KoRollupError
    with pytest.raises(KoRollupError):
        rollup.series_rollup(np.zeros(4), np.array([0, 3, 2, 4]), QS)          # decreasing offsets
    with pytest.raises(KoRollupError):
        rollup.series_rollup(np.zeros(4), np.array([0, 4]), [0.1, 0.2, 0.3, 0.4, 0.5])   # too many quantiles
    with pytest.raises(KoRollupError):
        rollup.series_rollup(np.zeros(4), np.array([0, 4]), [1.5])
    out = rollup.series_rollup(np.zeros(0), np.array([0]), QS)                 # zero series is fine
    assert len(out["count"]) == 0
