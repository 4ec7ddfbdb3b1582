"""Prometheus response ingest (ko_prom_open / ko_prom_fill, host C++) against the oracle's json + float().
No GPU is involved: these run in the CPU suite."""
import json

import numpy as np
import pytest

from kubeoperator_b200 import KoRollupError, prom_ingest
from oracle import rollup_oracle as ro


def _fmt(v):
    if v != v:
        return "NaN"
    if v == float("inf"):
        return "+Inf"
    if v == float("-inf"):
        return "-Inf"
    return repr(float(v))


def _matrix(rng, n_series, lengths, compact=True):
    result = []
    for i in range(n_series):
        L = int(lengths[i % len(lengths)])
        vals = rng.normal(0, 1, L) * 10.0 ** rng.integers(-8, 9, L)
        vals[rng.random(L) < 0.05] = np.nan
        if L > 3:
            vals[1], vals[2] = np.inf, -np.inf
        pairs = [[1435781430.781 + 15 * k, _fmt(v)] for k, v in enumerate(vals)]
        result.append({"metric": {"__name__": "node_cpu", "instance": "10.0.0.%d:9100" % i, "q": 'a"b\\c]}'},
                       "values": pairs})
    doc = {"status": "success", "data": {"resultType": "matrix", "result": result}}
    return json.dumps(doc, separators=(",", ":") if compact else None, indent=None if compact else 2).encode()


def _check(body, threads):
    got = prom_ingest.parse_response(body, want_timestamps=True, n_threads=threads)
    values, offsets, stamps, metrics = ro.prom_columnar(body)
    assert got["offsets"].tolist() == offsets
    exp = np.array(values, np.float64)
    assert np.array_equal(np.isnan(got["values"]), np.isnan(exp))
    ok = ~np.isnan(exp)
    assert np.array_equal(got["values"][ok].view(np.uint64), exp[ok].view(np.uint64))      # bit-exact
    assert np.array_equal(got["timestamps"].view(np.uint64), np.array(stamps, np.float64).view(np.uint64))
    assert got["metrics"] == metrics


@pytest.mark.parametrize("threads", [1, 5])
def test_matrix_matches_json_and_float(threads):
    rng = np.random.default_rng(21)
    _check(_matrix(rng, 37, [0, 1, 2, 240, 7, 1025]), threads)
    _check(_matrix(rng, 3, [20], compact=False), threads)            # pretty-printed: whitespace everywhere


def test_number_forms_round_like_python_float():
    strs = ["0", "-0", "1", "0.1", "1e+06", "1E-7", "123456789012345678901234567890", "4.9e-324", "2.2250738585072014e-308",
            "1.7976931348623157e+308", "1e400", "-1e400", "1e-400", "-1e-400", "0.30000000000000004", "9007199254740993",
            "NaN", "+Inf", "-Inf", "Inf", "5e-324", "2.4703282292062327e-324", "2.4703282292062328e-324"]
    doc = {"status": "success", "data": {"resultType": "matrix",
                                         "result": [{"metric": {}, "values": [[i, s] for i, s in enumerate(strs)]}]}}
    body = json.dumps(doc).encode()
    got = prom_ingest.parse_response(body)["values"]
    exp = np.array([float(s) for s in strs])
    assert np.array_equal(np.isnan(got), np.isnan(exp))
    ok = ~np.isnan(exp)
    assert np.array_equal(got[ok].view(np.uint64), exp[ok].view(np.uint64))


def test_instant_vector_as_the_reference_reads_it():
    # prometheus_client.py:94-95: result[0]['value'][1] -> here: one sample per series
    body = (b'{"status":"success","data":{"resultType":"vector","result":['
            b'{"metric":{"kubernetes_io_hostname":"node1"},"value":[1435781451.781,"0.4375"]},'
            b'{"metric":{"kubernetes_io_hostname":"node2"},"value":[1435781451.781,"NaN"]}]}}')
    got = prom_ingest.parse_response(body)
    assert got["offsets"].tolist() == [0, 1, 2]
    assert got["values"][0] == float(json.loads(body)["data"]["result"][0]["value"][1])
    assert np.isnan(got["values"][1])
    assert got["metrics"][1] == {"kubernetes_io_hostname": "node2"}
    _check(body, 2)


def test_empty_result_and_caller_buffer():
    body = b'{"status":"success","data":{"resultType":"matrix","result":[]}}'
    got = prom_ingest.parse_response(body)
    assert got["offsets"].tolist() == [0] and got["values"].size == 0
    rng = np.random.default_rng(22)
    body = _matrix(rng, 4, [50])
    buf = np.full(1000, -1.0)
    got = prom_ingest.parse_response(body, values_out=buf)
    assert got["values"].base is buf or got["values"].ctypes.data == buf.ctypes.data
    assert np.all(buf[200:] == -1.0)


def test_errors_are_loud():
    with pytest.raises(KoRollupError) as e:        # the reference's guard: status != success -> nothing is read
        prom_ingest.parse_response(b'{"status":"error","errorType":"bad_data","error":"parse error at char 5"}')
    assert "parse error at char 5" in str(e.value)
    for bad in (b'', b'[1,2]', b'{"status":"success","data":{"resultType":"matrix","result":[{"values":[[1,"x"]]}]}}',
                b'{"status":"success","data":{"resultType":"matrix","result":[{"values":[[1,"2"]',
                b'{"status":"success","data":{"resultType":"scalar","result":[1,"2"]}}'):
        with pytest.raises(KoRollupError):
            prom_ingest.parse_response(bad)
