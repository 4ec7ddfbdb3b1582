"""CPU: libkorollup.so loads and exports every symbol include/korollup.h declares (no compute calls),
and the product path fails loudly instead of falling back to the CPU."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "korollup.h")


def _declared():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(ko_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_exported_and_bound():
    from kubeoperator_b200 import build, korollup
    build.build()
    names = _declared()
    assert len(names) >= 17
    lib = ctypes.CDLL(build.LIB)
    for n in names:
        assert hasattr(lib, n), "libkorollup.so does not export %s" % n
    assert sorted(korollup.ABI.keys()) == names, "ctypes ABI table and header disagree"
    bound = korollup.load_library()
    assert bound.ko_version() == 100


def test_every_entry_point_cites_the_reference():
    src = open(HEADER).read()
    for anchor in ("cluster_monitor.py:163-182", "cluster_monitor.py:93-117", "cluster_monitor.py:450-463",
                   "prometheus_client.py:75-85", "dashboard.component.ts:117-153", "prometheus_client.py:140-146",
                   "api.py:500-501"):
        assert anchor in src


def test_no_cpu_fallback_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from kubeoperator_b200 import KoRollupError, Rollup
    with pytest.raises(KoRollupError) as e:
        Rollup(0)
    assert "no CPU fallback" in str(e.value) or "CUDA" in str(e.value)


def test_missing_library_is_loud(monkeypatch, tmp_path):
    from kubeoperator_b200 import korollup
    monkeypatch.setattr(korollup, "_lib", None)
    monkeypatch.setenv(korollup.LIB_ENV, str(tmp_path / "nope.so"))
    with pytest.raises(korollup.KoRollupError):
        korollup.load_library()
    monkeypatch.setattr(korollup, "_lib", None)


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "kubeoperator_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", text, flags=re.M), f
                assert "libkooracle" not in text and "c_oracle" not in text, f
