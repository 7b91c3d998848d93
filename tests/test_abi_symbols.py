"""CPU checks of the drop-in boundary: the library loads without a GPU and exports every
symbol include/pilott_b200.h declares (no compute calls here)."""
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "pilott_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(pilott_[a-z0-9_]+)\s*\(", src)))


@pytest.fixture(scope="module")
def lib():
    from pilottai_b200 import build, _lib
    build.build()
    return _lib.load()


def test_every_declared_symbol_is_exported_and_bound(lib):
    from pilottai_b200 import _lib
    declared = _declared()
    assert len(declared) >= 18
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in the header but not exported"
    assert sorted(_lib.SIGNATURES) == declared, "ctypes table and header disagree"


def test_version_and_error_string(lib):
    assert lib.pilott_version() == 100
    assert lib.pilott_kp() == 32
    assert isinstance(lib.pilott_last_error(), bytes)


def test_null_handle_is_rejected_without_a_gpu(lib):
    from pilottai_b200 import _lib
    assert lib.pilott_mem_clear(None, None) == _lib.EINVAL
    assert b"NULL" in lib.pilott_last_error()
    with pytest.raises(_lib.PilottError):
        _lib.check(lib.pilott_mem_set_option(None, 0, 0))


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    from pilottai_b200 import _lib
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", str(tmp_path / "nope.so"))
    with pytest.raises(ImportError, match="no CPU fallback"):
        _lib.load()
