"""CPU-side checks of the C-ABI boundary: the library builds/loads, exports every symbol that
include/tsm_b200.h declares, and the ctypes table covers the same set (no compute calls)."""
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "tsm_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(tsm_[a-z0-9_]+)\s*\(", src)))


@pytest.fixture(scope="module")
def libpath():
    from tianshou_marl_b200 import _lib

    if not os.path.exists(_lib.LIB_PATH):
        _lib.build()
    return _lib.LIB_PATH


def test_header_symbols_exported(libpath):
    out = subprocess.run(["nm", "-D", "--defined-only", libpath], capture_output=True, text=True, check=True).stdout
    exported = set(re.findall(r"\bT (tsm_[a-z0-9_]+)", out))
    missing = [s for s in _declared() if s not in exported]
    assert not missing, f"declared in tsm_b200.h but not exported: {missing}"


def test_ctypes_table_matches_header(libpath):
    from tianshou_marl_b200 import _lib

    assert sorted(_lib.exported_symbols()) == _declared()
    l = _lib.lib()  # binds every symbol; AttributeError if one is absent
    assert l.tsm_abi_version() == 1


def test_sass_is_sm100(libpath):
    out = subprocess.run(["cuobjdump", "-lelf", libpath], capture_output=True, text=True).stdout
    assert "sm_100a" in out, out


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    from tianshou_marl_b200 import _lib

    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", str(tmp_path / "nope.so"))
    with pytest.raises(_lib.TsmError):
        _lib.lib()
