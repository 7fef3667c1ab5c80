"""CPU-only: the C-ABI library loads and exports every symbol include/opb200.h declares."""
import os
import re

from openpose_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    names = set()
    inc = os.path.join(ROOT, "include")
    for f in os.listdir(inc):
        if f.endswith(".h"):
            text = open(os.path.join(inc, f)).read()
            text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
            names.update(re.findall(r"\b(opb200_[a-z0-9_]+)\s*\(", text))
    return names


def test_library_exports_every_declared_symbol():
    import ctypes
    assert os.path.exists(_lib.LIB_PATH), "build first: python -c 'import __graft_entry__ as g; g.build()'"
    handle = ctypes.CDLL(_lib.LIB_PATH)
    declared = _declared_symbols()
    assert len(declared) >= 10
    missing = [n for n in sorted(declared) if not hasattr(handle, n)]
    assert not missing, missing


def test_python_binding_covers_header():
    import openpose_b200.net  # noqa: F401  (registers its symbols)
    import openpose_b200.filestream  # noqa: F401
    import openpose_b200.engine  # noqa: F401
    declared = _declared_symbols()
    assert declared <= set(_lib.SYMBOLS), sorted(declared - set(_lib.SYMBOLS))
    L = _lib.lib()
    assert L.opb200_version() == 1
    assert isinstance(L.opb200_last_error(), bytes)
