"""C++ user code over Arrow host tables (tests/cpp/arrow_bridge.cpp): include/cylon/arrow_bridge.hpp (Table::FromArrowTable /
ToArrowTable / get_table, table.hpp:76,96,146 of the reference) and the std::vector forms of MapToHashPartitions / Split /
PartitionByHashing (partition/partition.hpp:32-90). The CPU test builds and links it against pyarrow's Arrow C++; the GPU
test runs it on cuda:0."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO_DIR = os.path.join(ROOT, "cylon_b200", "lib")


def _build(tmp_path):
    try:
        import pyarrow as pa
    except ImportError:
        pytest.skip("needs pyarrow's Arrow C++ headers")
    if shutil.which("g++") is None or not os.path.exists(os.path.join(SO_DIR, "libcylon_b200.so")):
        pytest.skip("needs g++ and the built library")
    libdir = pa.get_library_dirs()[0]
    libs = [f for f in os.listdir(libdir) if f.startswith("libarrow.so")]
    if not libs:
        pytest.skip("pyarrow ships no libarrow.so here")
    exe = str(tmp_path / "arrow_bridge")
    cmd = ["g++", "-std=c++20", "-O1", "-I" + os.path.join(ROOT, "include"), "-I" + pa.get_include(), os.path.join(ROOT, "tests", "cpp", "arrow_bridge.cpp"),
           "-o", exe, "-L" + SO_DIR, "-lcylon_b200", "-Wl,-rpath," + SO_DIR, os.path.join(libdir, sorted(libs)[-1]), "-Wl,-rpath," + libdir]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-4000:]
    return exe


def test_arrow_bridge_builds_and_links(tmp_path):
    exe = _build(tmp_path)
    assert subprocess.run([exe]).returncode == 0  # loads both libraries, returns before any GPU call


@pytest.mark.gpu
def test_arrow_bridge_runs_on_gpu(tmp_path):
    exe = _build(tmp_path)
    r = subprocess.run([exe, "run"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "ARROW_BRIDGE_OK" in r.stdout
