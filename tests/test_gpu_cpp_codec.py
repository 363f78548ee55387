"""GPU test of the C++ host mirror of the reference's codec seam (include/tfs_codec.hpp): a small C++ program
drives GpuCodec::encoder() / decoder() through the C ABI and checks Appendix-B vectors, the oracle, round trips,
batching and error mapping.  The program is built by __graft_entry__.build()."""
import os
import subprocess

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "tests", "cpp", "test_codec")


def test_cpp_codec_program():
    assert os.path.exists(BIN), "tests/cpp/test_codec is missing: run __graft_entry__.build()"
    r = subprocess.run([BIN], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "cpp codec ok" in r.stdout, r.stdout + r.stderr
