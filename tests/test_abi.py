"""CPU-side checks of the boundary: the C-ABI library loads, exports every symbol include/ani_b200.h declares,
parses parameter files, and refuses to run without a CUDA device (no CPU fallback)."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    from lammps_ani_b200 import build, runner
    build.build()
    return runner.load_library()


def test_header_symbols_are_exported(lib):
    hdr = open(os.path.join(ROOT, "include", "ani_b200.h")).read()
    declared = set(re.findall(r"^(?:int|void|const char\*)\s+(ani_\w+)\s*\(", hdr, flags=re.M))
    from lammps_ani_b200 import runner
    assert declared == set(runner.EXPORTED_SYMBOLS)
    for name in declared:
        assert hasattr(lib, name), name
    assert lib.ani_abi_version() == 1


def test_model_load_and_info(lib, ani1xnr_path):
    m = C.c_void_p()
    assert lib.ani_model_load(ani1xnr_path.encode(), C.byref(m)) == 0
    ns, al, nm = C.c_int(), C.c_int(), C.c_int()
    rcr, rca = C.c_double(), C.c_double()
    assert lib.ani_model_info(m, C.byref(ns), C.byref(al), C.byref(nm), C.byref(rcr), C.byref(rca)) == 0
    assert (ns.value, al.value, nm.value, rcr.value, rca.value) == (4, 448, 8, 5.2, 3.5)
    lib.ani_model_free(m)


def test_bad_model_files_fail_loudly(lib, tmp_path):
    m = C.c_void_p()
    assert lib.ani_model_load(str(tmp_path / "missing.aniw").encode(), C.byref(m)) != 0
    assert b"cannot open" in lib.ani_last_error()
    bad = tmp_path / "bad.aniw"
    bad.write_bytes(b"not a model" * 20)
    assert lib.ani_model_load(str(bad).encode(), C.byref(m)) != 0
    pt = tmp_path / "ani2x.pt"
    pt.write_bytes(b"PK\x03\x04")
    assert lib.ani_model_load(str(pt).encode(), C.byref(m)) != 0
    assert b"lammps_ani_b200.extract" in lib.ani_last_error()   # tells the user how to extract once


def test_no_cpu_fallback(lib, ani1xnr_path):
    import torch
    m = C.c_void_p()
    assert lib.ani_model_load(ani1xnr_path.encode(), C.byref(m)) == 0
    ctx = C.c_void_p()
    # `device cpu` (local_rank -1 in the reference) is rejected outright
    assert lib.ani_ctx_create(m, -1, -1, 1, 1, 1, C.byref(ctx)) != 0
    assert b"CPU" in lib.ani_last_error() or b"cpu" in lib.ani_last_error()
    if not torch.cuda.is_available():
        assert lib.ani_ctx_create(m, 0, -1, 1, 1, 1, C.byref(ctx)) != 0
        assert b"no CUDA device" in lib.ani_last_error()
        from lammps_ani_b200.runner import ANI
        with pytest.raises(RuntimeError):
            ANI(ani1xnr_path, 0)
    lib.ani_model_free(m)


def test_kokkos_branch_of_the_pair_style_type_checks():
    """`ani/kk` on Kokkos views (-DLMP_KOKKOS, reference src/pair_ani_kokkos.cpp:67-235) against the in-tree LAMMPS KOKKOS headers
    and Kokkos 3.7.01 (Serial backend), g++ -fsyntax-only.  Needs /root/reference (skipped on the GPU box)."""
    import os
    import pytest
    from lammps_ani_b200 import build
    if not os.path.exists(os.path.join(os.path.dirname(build.LAMMPS_SRC), "lib", "kokkos", "core", "src", "Kokkos_Core.hpp")):
        pytest.skip("LAMMPS / Kokkos sources not present")
    build.check_kokkos_branch()
