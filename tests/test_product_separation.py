"""The product (cylon_b200/, include/) must not use the checker (oracle/) in any form, and must fail loudly -- not fall
back to a CPU path -- when the CUDA library is missing."""
import os
import re
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _product_sources():
    for base in ("cylon_b200",):
        for dirpath, _dirs, files in os.walk(os.path.join(ROOT, base)):
            if "__pycache__" in dirpath:
                continue
            for f in files:
                if f.endswith((".py", ".cu", ".cuh", ".cpp", ".hpp", ".h")) or f == "Makefile":
                    yield os.path.join(dirpath, f)


def test_no_product_source_mentions_the_oracle():
    pat = re.compile(r"\boracle\b|pyoracle|pyref|liboracle|libcylon_ref|libref_", re.IGNORECASE)
    hits = []
    for path in _product_sources():
        with open(path, errors="replace") as fh:
            for ln, line in enumerate(fh, 1):
                if pat.search(line):
                    hits.append("%s:%d: %s" % (os.path.relpath(path, ROOT), ln, line.strip()))
    assert not hits, "\n".join(hits)


def test_shared_library_does_not_link_the_checkers():
    so = os.path.join(ROOT, "cylon_b200", "lib", "libcylon_b200.so")
    if not os.path.exists(so):
        pytest.skip("library not built")
    needed = subprocess.run(["readelf", "-d", so], capture_output=True, text=True).stdout
    libs = re.findall(r"\(NEEDED\)\s+Shared library: \[(.*?)\]", needed)
    assert libs, "readelf printed no NEEDED entries"
    assert not [x for x in libs if "oracle" in x or "cylon_ref" in x or "arrow" in x], libs


def test_missing_library_is_a_hard_error():
    """Importing the package with the .so path pointing nowhere must raise at first use, not compute on the CPU."""
    code = ("import os, sys; sys.path.insert(0, %r)\n"
            "from cylon_b200 import capi\n"
            "capi.LIB_PATH = os.path.join(%r, 'definitely_missing.so')\n"
            "capi._lib = None\n"
            "try:\n"
            "    capi.lib()\n"
            "except Exception as e:\n"
            "    print('RAISED', type(e).__name__)\n"
            "else:\n"
            "    print('NO ERROR')\n") % (ROOT, ROOT)
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=120).stdout
    assert "RAISED" in out, out
