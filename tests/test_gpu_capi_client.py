"""Runs the C++ client that was compiled against the reference's own headers (tests/capi/mpm88_client.cpp: the flow of
c_api/tests/mpm88_test.cpp through ti::Runtime / ti::AotModule / ti::NdArray / ti::Kernel) on the GPU.  The binary is
built where /root/reference exists (__graft_entry__.build / tests/test_cabi_reference.py) and travels with the repo."""
import os
import re
import subprocess

import pytest

from _b2 import have_gpu

pytestmark = pytest.mark.gpu
BIN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "capi", "bin", "mpm88_client")


def test_reference_cpp_client_runs():
    if not have_gpu():
        pytest.skip("no CUDA device")
    if not os.path.exists(BIN):
        pytest.skip("tests/capi/bin/mpm88_client not built (needs the reference headers at build time)")
    r = subprocess.run([BIN], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    m = re.search(r"finite=1 mean_y0=([0-9.]+) mean_y1=([0-9.]+) capabilities=0 tcm_module=0 tcm_error=-1 launches=(\d+)", r.stdout)
    assert m, r.stdout
    assert float(m.group(2)) < float(m.group(1))          # 50 substeps of gravity: the cloud has fallen
    assert int(m.group(3)) >= 200
