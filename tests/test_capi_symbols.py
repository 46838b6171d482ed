"""The C-ABI library loads (no device needed) and exports every symbol include/hyteg_b200.h declares."""
import ctypes
import os
import re

from hyteg_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "hyteg_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    names = set(re.findall(r"\b(hb_[a-z0-9_]+)\s*\(", src))
    names.discard("hb_allreduce_fn")
    return sorted(names)


def test_header_and_binding_agree():
    assert set(declared_symbols()) == set(capi.EXPORTED_SYMBOLS)


def test_library_exports_all_symbols():
    import __graft_entry__
    __graft_entry__.build()
    dll = ctypes.CDLL(capi.DEFAULT_LIB)
    for name in declared_symbols():
        assert hasattr(dll, name), name
    dll.hb_version.restype = ctypes.c_char_p
    assert b"sm_100a" in dll.hb_version()


def test_product_has_no_cpu_fallback():
    """the package never references the oracle or the emulation build"""
    pkg = os.path.join(ROOT, "hyteg_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cpp", ".h", ".hpp", ".inc")):
                text = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in text and "from oracle" not in text, f
                assert "libhyteg_b200_emul" not in text, f
