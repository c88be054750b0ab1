"""Runs the reference-style C++ caller built over include/dietgpu_b200.hpp (tests/cpp/adaptor_roundtrip.cpp,
the batch round trips of the reference's ANSTest.cu / FloatTest.cu through namespace dietgpu)."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "tests", "cpp", "adaptor_roundtrip.bin")


def test_adaptor_binary_is_built():
    assert os.path.exists(BIN), "run __graft_entry__.build() (make -C dietgpu_fork_b200/csrc)"


@pytest.mark.gpu
def test_adaptor_round_trips():
    r = subprocess.run([BIN], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "adaptor round trips OK" in r.stdout
