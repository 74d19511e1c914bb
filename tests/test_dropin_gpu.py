"""The C++ drop-in on the GPU: tests/cpp/dropin_test.cpp is the reference's own test code
(test_multiexp.cpp:112-256,344-390) compiled against the UNMODIFIED libff headers plus
include/libff_b200.hpp, so libff::multi_exp / batch_exp / batch_to_special run through the b200msm
C ABI.  The binary is built where /root/reference exists (`make -C oracle dropin`) and travels to
the GPU box inside oracle/_ref/."""
import os
import subprocess

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "oracle", "_ref", "dropin_test")


def test_cpp_dropin_runs_reference_tests_on_gpu():
    if not os.path.exists(EXE):
        pytest.skip("oracle/_ref/dropin_test not built (needs /root/reference: make -C oracle dropin)")
    p = subprocess.run([EXE], capture_output=True, text=True, timeout=600)
    print(p.stdout[-2000:], p.stderr[-2000:])
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-2000:]
    assert " 0 failures" in p.stdout
