"""The C-ABI library must load on a machine without a GPU and export every function include/*.h declares."""
import ctypes
import glob
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_functions():
    names = []
    for h in sorted(glob.glob(os.path.join(ROOT, "include", "*.h"))):
        text = open(h).read()
        text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
        names += re.findall(r"\b(cyl_[a-z0-9_]+)\s*\(", text)
        names += re.findall(r"\b(cylon_[a-z0-9_]+)\s*\(", text)
    return sorted(set(names))


def test_library_exports_every_declared_symbol():
    path = os.path.join(ROOT, "cylon_b200", "lib", "libcylon_b200.so")
    assert os.path.exists(path), "build the library first: python -c 'import __graft_entry__ as g; g.build()'"
    lib = ctypes.CDLL(path)
    names = declared_functions()
    assert len(names) >= 20
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, "declared in include/*.h but not exported: %s" % missing


def test_version_and_error_string_without_gpu():
    lib = ctypes.CDLL(os.path.join(ROOT, "cylon_b200", "lib", "libcylon_b200.so"))
    assert lib.cyl_version() >= 100
    lib.cyl_last_error.restype = ctypes.c_char_p
    assert isinstance(lib.cyl_last_error(), bytes)
