"""CPU tests of the C-ABI library: it loads without a GPU and exports every symbol include/b200vit.h declares."""
import ctypes
import re
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent


@pytest.fixture(scope="module")
def libpath():
    import __graft_entry__ as ge
    ge.build()
    from b200vit import _C
    p = _C.library_path()
    assert p.exists()
    return p


def declared_symbols():
    text = (ROOT / "include" / "b200vit.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(b200vit_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_are_exported(libpath):
    lib = ctypes.CDLL(str(libpath))
    syms = declared_symbols()
    assert len(syms) >= 20
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in include/b200vit.h but not exported"


def test_python_binding_covers_header(libpath):
    from b200vit import _C
    assert sorted(_C.exported_symbols()) == declared_symbols()
    assert _C.lib().b200vit_abi_version() == _C.ABI_VERSION


def test_argument_errors_are_reported_without_a_gpu(libpath):
    from b200vit import _C
    lib = _C.lib()
    rc = lib.b200vit_gemm(9, 0, 1, 1, 1, None, 8, None, 8, None, 8, None, 0, None, None, 0, 0.0, 0, 0, 0, 1, None)
    assert rc < 0
    assert b"bad op" in lib.b200vit_last_error()
    rc = lib.b200vit_layernorm_fwd(None, 4, None, None, None, None, None, None, 1, 4, 1e-5, None)
    assert rc < 0 and b"null" in lib.b200vit_last_error()
