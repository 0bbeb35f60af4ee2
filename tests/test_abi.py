"""CPU-side checks of the C-ABI boundary: the library builds for sm_100a, loads, and exports every
symbol include/b200km.h declares.  No compute call is made here (there is no GPU in this container)."""
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    from spark_rapids_ml_nai_b200 import _native, build

    build.build()
    return _native.load()


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "b200km.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(b200km_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_are_exported_and_bound(lib):
    from spark_rapids_ml_nai_b200 import _native

    declared = _declared_symbols()
    assert len(declared) >= 12
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in include/b200km.h but not exported"
    assert sorted(_native.exported_symbols()) == declared


def test_version_and_error_channel(lib):
    from spark_rapids_ml_nai_b200 import _native

    assert "sm_100a" in _native.version()
    assert _native.query_workspace(1000, 20, 8) > 0
    with pytest.raises(RuntimeError, match="bad shape"):
        _native.query_workspace(-1, 20, 8)


def test_no_cpu_fallback_when_library_missing(monkeypatch):
    from spark_rapids_ml_nai_b200 import _native

    monkeypatch.setattr(_native, "_lib", None)
    monkeypatch.setattr(_native, "LIB_PATH", "/nonexistent/libb200km.so")
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        _native.load()


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "spark_rapids_ml_nai_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in src, f"{f} mentions the oracle; the product path must not use it"
