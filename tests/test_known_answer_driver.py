"""tools/test_model: the stand-alone C++ known-answer driver over the C ABI (reference: src/ani_csrc/test_model.cpp:64-174,
run by tests/test_all.sh:7-12 for {cpu,cuda} x {double,single})."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def binary():
    from lammps_ani_b200 import build
    build.build()
    return build.build_test_model()


def test_driver_builds_and_refuses_cpu(binary, ani1xnr_path):
    r = subprocess.run([binary, ani1xnr_path, "cpu", "single"], capture_output=True, text=True, timeout=120)
    assert r.returncode != 0 and "without a CPU path" in r.stderr       # no CPU fallback behind the reference's `cpu` token
    r = subprocess.run([binary, ani1xnr_path, "tpu", "single"], capture_output=True, text=True, timeout=120)
    assert r.returncode == 255 and "2nd argument must be <cpu/cuda>" in r.stderr


@pytest.mark.gpu
@pytest.mark.parametrize("precision", ["double", "single"])
def test_driver_known_answer_ani1xnr(binary, ani1xnr_path, precision):
    """ANI-1xnr weights (in tree): expectations = the fp64 oracle on the reference's inputs; thresholds 1e-8 / 3e-4 as
    test_model.cpp:164.  With an ANI-2x model the same binary checks the reference's own numbers (test_model.cpp:121-137)."""
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    r = subprocess.run([binary, ani1xnr_path, "cuda", precision], capture_output=True, text=True, timeout=300)
    print(r.stdout, r.stderr)
    assert r.returncode == 0 and r.stdout.strip().endswith("OK")
