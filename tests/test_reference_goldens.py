"""Auto-pinning against the reference's OWN absolute known-answer vectors.

Every absolute number the reference tree holds for this path is an ANI-2x (8 members) result on the 30-atom water case:
  * src/ani_csrc/test_model.cpp:121-137  -- energy -763.9931790697472 Ha and 90 force components (half list, non-periodic),
    thresholds 1e-8 (fp64) / 3e-4 (fp32)  (test_model.cpp:164);
  * the 15 YAML files under tests/lammps-unittest/ -- init_vdwl, init_forces, init_stress per pair_style line, each with its
    own `epsilon` (relative error per component, as LAMMPS's EXPECT_FP_LE_WITH_EPS computes it).
They are carried in tests/golden/water30_reference.json (tools/make_golden.py).  The ANI-2x parameters themselves are NOT in
/root/reference (torchani downloads them at export time) and there is no network, so these tests SKIP with that reason until
a model file appears; the moment one does -- `ani2x.aniw`, an exported `ani2x.pt`, or torchani's `ani-2x_8x.info` NeuroChem
directory, in any of the places listed in `_CANDIDATES` or named by $ANI2X_MODEL -- the oracle (CPU, `-m "not gpu"`) and the
CUDA path (`-m gpu`) are both checked against every vector, and "parity unpinned"  can be struck."""
import glob
import json
import os

import numpy as np
import pytest

from lammps_ani_b200 import domain

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HARTREE2KCALMOL = 627.5094738898777

_CANDIDATES = [
    "ani2x.aniw", "lammps_ani_b200/data/ani2x.aniw", "ani2x.pt", "models/ani2x.pt", "lammps_ani_b200/data/ani2x.pt",
    "baseline/_ref/**/ani2x.pt", "baseline/_ref/**/ani-2x_8x.info", "oracle/_ref/**/ani2x.pt",
]


def find_ani2x_source():
    env = os.environ.get("ANI2X_MODEL")
    if env and os.path.exists(env):
        return env
    for pat in _CANDIDATES:
        hits = sorted(glob.glob(os.path.join(ROOT, pat), recursive=True))
        if hits:
            return hits[0]
    try:                                   # a torchani installation carries the NeuroChem resources once downloaded
        import importlib.util
        spec = importlib.util.find_spec("torchani")
        if spec and spec.submodule_search_locations:
            for base in spec.submodule_search_locations:
                hits = sorted(glob.glob(os.path.join(base, "resources", "**", "ani-2x_8x.info"), recursive=True))
                if hits:
                    return hits[0]
    except Exception:
        pass
    return None


@pytest.fixture(scope="module")
def ani2x_path(tmp_path_factory):
    src = find_ani2x_source()
    if src is None:
        pytest.skip("no ANI-2x parameters available offline (looked for $ANI2X_MODEL, " + ", ".join(_CANDIDATES) +
                    ", torchani resources): the reference's absolute goldens cannot be evaluated -- parity stays unpinned")
    if src.endswith(".aniw"):
        return src
    from lammps_ani_b200 import extract
    out = str(tmp_path_factory.mktemp("ani2x") / "ani2x.aniw")
    extract.extract(src, out)
    return out


@pytest.fixture(scope="module")
def gold():
    with open(os.path.join(ROOT, "tests", "golden", "water30_reference.json")) as f:
        return json.load(f)


def _rel_err(val, ref):
    """LAMMPS unittest EXPECT_FP_LE_WITH_EPS: |a-b| / min(|a|,|b|)  (absolute when that is zero)."""
    val, ref = np.asarray(val, dtype=float), np.asarray(ref, dtype=float)
    diff = np.abs(val - ref)
    div = np.minimum(np.abs(val), np.abs(ref))
    return np.where(div == 0.0, diff, diff / np.where(div == 0.0, 1.0, div))


def _style(case):
    """pair_style ani 5.1 <model> <device> <num_models> <aev> <nbr> <prec>  ->  (use_cuaev, use_fullnbr, use_single)"""
    tok = case["pair_style"].split()
    return tok[5] == "cuaev", tok[6] == "full", tok[7] == "single"


def _system(gold, periodic):
    x = np.array(gold["coords"]).reshape(-1, 3)
    sp = np.array(gold["species"], dtype=np.int32)
    if periodic:
        sub = domain.decompose(x + 4.0, sp, np.array([8.0, 8.0, 8.0]))[0]
    else:
        sub = domain.decompose(x + 4.0, sp, np.array([8.0, 8.0, 8.0]), periodic=False)[0]
    return sub


def _lists(sub, full):
    if full:
        return domain.full_neighbor_list(sub.x, sub.nlocal)
    return (domain.half_neighbor_list(sub.x, sub.nlocal),)


def _check_case(case, energy, forces_folded, virial):
    eps = case["epsilon"]
    assert _rel_err(energy, case["init_vdwl"]) <= eps, (energy, case["init_vdwl"])
    assert _rel_err(forces_folded, np.array(case["init_forces"])).max() <= eps
    if case["init_stress"] is not None and virial is not None:
        W = virial
        got = np.array([W[0, 0], W[1, 1], W[2, 2], W[0, 1], W[0, 2], W[1, 2]])
        assert _rel_err(got, np.array(case["init_stress"])).max() <= eps


# ------------------------------------------------------------------------------------------------ CPU: the oracle
def test_oracle_reproduces_test_model_cpp(ani2x_path, gold):
    """test_model.cpp:64-174 -- hard-coded half list, use_cuaev = false (no radial re-filter, note N3), fp64: 1e-8."""
    from oracle.oracle import OracleANI
    x = np.array(gold["coords"]).reshape(-1, 3)
    sp = np.array(gold["species"], dtype=np.int32)
    idx = np.array(gold["atom_index12"], dtype=np.int64).reshape(2, -1)
    r = OracleANI(ani2x_path).compute_half(sp, x, idx, 30, radial_filter=False)
    tm = gold["test_model"]
    assert abs(r["energy"] / HARTREE2KCALMOL - tm["expected_energy_hartree"]) <= tm["threshold_fp64"]
    assert np.abs(r["forces"].reshape(-1) - np.array(tm["expected_force_kcalmol_A"])).max() <= tm["threshold_fp64"]


def test_oracle_reproduces_lammps_unittest_yaml(ani2x_path, gold):
    from oracle.oracle import OracleANI
    orc = OracleANI(ani2x_path)
    for case in gold["yaml"]:
        cuaev, full, single = _style(case)
        sub = _system(gold, case["periodic"])
        nbr = _lists(sub, full)
        kw = dict(single_coords=single, radial_filter=(cuaev and full))
        r = orc.compute_full(sub.species, sub.x, *nbr, sub.nlocal, **kw) if full else orc.compute_half(sub.species, sub.x, nbr[0], sub.nlocal, **kw)
        # the fp32 goldens were produced by fp32 arithmetic: the fp64 oracle must sit inside their epsilon too
        _check_case(case, r["energy"], domain.fold_ghost_forces(sub, r["forces"]), r["virial"])


# ------------------------------------------------------------------------------------------------ GPU: the CUDA path
@pytest.mark.gpu
@pytest.mark.parametrize("single", [False, True])
def test_cuda_reproduces_test_model_cpp(ani2x_path, gold, single):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from lammps_ani_b200 import runner
    x = np.array(gold["coords"]).reshape(-1, 3)
    sp = np.array(gold["species"], dtype=np.int32)
    idx = np.array(gold["atom_index12"], dtype=np.int64).reshape(2, -1)
    ani = runner.ANI(ani2x_path, 0, -1, False, False, single)            # test_model.cpp:83: use_cuaev=false, use_fullnbr=false
    e, f, _, _ = ani.compute(sp, x, idx, nlocal=30, ago=0)
    tm = gold["test_model"]
    thr = tm["threshold_fp32"] if single else tm["threshold_fp64"]
    assert abs(e / HARTREE2KCALMOL - tm["expected_energy_hartree"]) <= thr
    assert np.abs(f.reshape(-1) - np.array(tm["expected_force_kcalmol_A"])).max() <= thr


@pytest.mark.gpu
def test_cuda_reproduces_lammps_unittest_yaml(ani2x_path, gold):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from lammps_ani_b200 import runner
    for case in gold["yaml"]:
        cuaev, full, single = _style(case)
        sub = _system(gold, case["periodic"])
        nbr = _lists(sub, full)
        ani = runner.ANI(ani2x_path, 0, -1, cuaev, full, single)
        if full:
            il, nn, jl = nbr
            e, f, _, vir = ani.compute(sub.species, sub.x, il, jl, nn, nlocal=sub.nlocal, ago=0, vflag=True)
        else:
            e, f, _, vir = ani.compute(sub.species, sub.x, nbr[0], nlocal=sub.nlocal, ago=0, vflag=True)
        _check_case(case, e, domain.fold_ghost_forces(sub, f), vir if not cuaev else None)   # cuaev computes no virial (lammps_ani.py:208-211)
