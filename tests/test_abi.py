"""The C-ABI library loads on a CPU-only box and exports every symbol the header declares."""

import ctypes
import re
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent


def declared_symbols():
    text = (ROOT / "include" / "nerftex_b200.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(ntx_[a-z0-9_]+)\s*\(", text)))


def test_library_builds_loads_and_exports_header_symbols():
    import __graft_entry__
    __graft_entry__.build()
    from nerf_tex_b200 import _lib
    lib = ctypes.CDLL(str(_lib.LIB_PATH))
    names = declared_symbols()
    assert "ntx_mlp_forward" in names and "ntx_composite_forward" in names
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, f"header declares symbols the library does not export: {missing}"
    unbound = [n for n in names if n not in _lib.SIGNATURES]
    assert not unbound, f"ctypes table misses: {unbound}"
    lib.ntx_version.restype = ctypes.c_int
    assert lib.ntx_version() == 100
