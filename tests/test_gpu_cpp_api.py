"""Runs the C++ gate test (tests/cpp/test_gate_gpu.cc): the reference's test_gate_gpu.cc
user code recompiled against include/cufhe_b200.hpp (per-gate Ctxt/Stream API + the
batched calls), for lvl1 and lvl0 parameters, all 13 gates + NMux."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.gpu


def test_cpp_gate_api():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    exe = os.path.join(ROOT, "tests", "cpp", "test_gate_gpu")
    if not os.path.exists(exe):
        import __graft_entry__ as g
        g.build_cpp_test()
    r = subprocess.run([exe], capture_output=True, text=True, timeout=900)
    print(r.stdout[-3000:], r.stderr[-2000:])
    assert r.returncode == 0 and "ALL PASS" in r.stdout
