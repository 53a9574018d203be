"""The C++ side of the boundary: shim/ChCollisionSystemB200 (compiled against the reference's real ChCollisionData.h)
drives a bed through the C ABI and must leave in the reference's data contract exactly what the reference's own collision
code computes (shim/test_shim.cpp; built by __graft_entry__.build() where /root/reference is available)."""
import os
import subprocess

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "shim", "_build", "test_shim")


def test_cpp_shim_fills_reference_containers_bit_exact():
    if not os.path.exists(BIN):
        pytest.skip("shim/_build/test_shim not built (needs the reference headers at build time)")
    r = subprocess.run([BIN], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and r.stdout.strip().splitlines()[-1].startswith("PASS"), r.stdout + r.stderr
