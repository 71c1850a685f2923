"""The C ABI driven from plain C (tests/c/abi_smoke.c): decode -> assign -> loss fwd+bwd -> NMS ->
count read, checked inside the program against scalar restatements of the reference formulas."""
import os
import subprocess

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_abi_smoke_c_program(dev):
    exe = os.path.join(ROOT, "tests", "c", "abi_smoke")
    assert os.path.exists(exe), "build it: make -C yolo_dl_b200/csrc abi_smoke"
    r = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    print(r.stdout, r.stderr)
    assert r.returncode == 0, (r.returncode, r.stdout, r.stderr)
    assert "abi_smoke OK" in r.stdout
