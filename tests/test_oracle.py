"""CPU tests of the oracle itself (no GPU): two independent restatements agree, analytic forces and virial
match finite differences, neighbour-list formats agree, committed goldens reproduce."""
import json
import os

import numpy as np
import pytest

from lammps_ani_b200 import aniw, domain
from oracle.oracle import OracleANI, HARTREE2KCALMOL
from oracle import pyaev_torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_list_is_the_7p1_half_list(water30):
    """test_model.cpp's hard-coded atom_index12 is the non-periodic half list at cutoff+skin=7.1 A."""
    mine = domain.half_neighbor_list(water30["x"], 30, 7.1)
    ref = water30["atom_index12"]
    assert mine.shape == ref.shape
    assert np.array_equal(mine, ref)


@pytest.mark.parametrize("model", ["ani1xnr", "ani2x_synth"])
@pytest.mark.parametrize("radial_filter", [True, False])
def test_c_oracle_matches_torch_pyaev(water30, ani1xnr_path, ani2x_synth_path, model, radial_filter):
    path = ani1xnr_path if model == "ani1xnr" else ani2x_synth_path
    params = aniw.load(path)
    orc = OracleANI(path)
    a = orc.compute_half(water30["species"], water30["x"], water30["atom_index12"], 30,
                         radial_filter=radial_filter, want_aev=True)
    b = pyaev_torch.pyaev_forward(params, water30["species"], water30["x"], water30["atom_index12"], 30,
                                  radial_filter=radial_filter, want_aev=True)
    assert np.allclose(a["aev"], b["aev"], rtol=1e-12, atol=1e-14)
    assert abs(a["energy"] - b["energy"]) < 1e-9 * abs(b["energy"])
    assert np.abs(a["forces"] - b["forces"]).max() < 1e-9
    assert np.abs(a["eatom"] - b["eatom"]).max() < 1e-8
    assert np.abs(a["virial"] - b["virial"]).max() < 1e-8


def test_periodic_with_ghosts_matches_torch_pyaev(water30, ani1xnr_path):
    """8 A box with a 7.1 A halo: every atom appears many times as a ghost."""
    params = aniw.load(ani1xnr_path)
    sub = domain.decompose(water30["x"] + 4.0, water30["species"], np.array([8.0, 8.0, 8.0]))[0]
    assert sub.nlocal == 30 and sub.nghost > 300
    il, nn, jl = domain.full_neighbor_list(sub.x, sub.nlocal)
    half = domain.half_neighbor_list(sub.x, sub.nlocal)
    orc = OracleANI(ani1xnr_path)
    a = orc.compute_full(sub.species, sub.x, il, nn, jl, sub.nlocal)
    h = orc.compute_half(sub.species, sub.x, half, sub.nlocal)
    b = pyaev_torch.pyaev_forward(params, sub.species, sub.x, half, sub.nlocal)
    for other in (h, b):
        assert abs(a["energy"] - other["energy"]) < 1e-9 * abs(a["energy"])
        assert np.abs(a["forces"] - other["forces"]).max() < 1e-9
        assert np.abs(a["virial"] - other["virial"]).max() < 1e-8
    # translational invariance: folded forces sum to zero
    f = domain.fold_ghost_forces(sub, a["forces"])
    assert np.abs(f.sum(0)).max() < 1e-8


def test_forces_and_virial_by_finite_differences(water30, ani2x_synth_path):
    orc = OracleANI(ani2x_synth_path)
    x = water30["x"]; sp = water30["species"]
    il, nn, jl = domain.full_neighbor_list(x, 30)
    base = orc.compute_full(sp, x, il, nn, jl, 30)
    h = 1e-5
    rng = np.random.default_rng(0)
    for (i, c) in [(0, 0), (7, 1), (13, 2), (29, 0)]:
        xp = x.copy(); xp[i, c] += h
        xm = x.copy(); xm[i, c] -= h
        ep = orc.compute_full(sp, xp, il, nn, jl, 30)["energy"]
        em = orc.compute_full(sp, xm, il, nn, jl, 30)["energy"]
        fd = -(ep - em) / (2 * h)
        assert abs(fd - base["forces"][i, c]) < 1e-5 * max(1.0, abs(fd))
    # virial: E(x (I + eps)) derivative, non-periodic cluster => V = -dE/d eps (symmetrised)
    for (a, b) in [(0, 0), (1, 1), (0, 1), (1, 2)]:
        eps = np.zeros((3, 3)); eps[a, b] = h
        ep = orc.compute_full(sp, x @ (np.eye(3) + eps).T, il, nn, jl, 30)["energy"]
        em = orc.compute_full(sp, x @ (np.eye(3) - eps).T, il, nn, jl, 30)["energy"]
        eps2 = np.zeros((3, 3)); eps2[b, a] = h
        ep2 = orc.compute_full(sp, x @ (np.eye(3) + eps2).T, il, nn, jl, 30)["energy"]
        em2 = orc.compute_full(sp, x @ (np.eye(3) - eps2).T, il, nn, jl, 30)["energy"]
        fd = -0.5 * ((ep - em) + (ep2 - em2)) / (2 * h)
        assert abs(fd - base["virial"][a, b]) < 2e-5 * max(1.0, abs(fd))


def test_num_models_selection_and_atomic_sum(water30, ani1xnr_path):
    x, sp = water30["x"], water30["species"]
    il, nn, jl = domain.full_neighbor_list(x, 30)
    e_all = OracleANI(ani1xnr_path).compute_full(sp, x, il, nn, jl, 30)
    e_4 = OracleANI(ani1xnr_path, use_num_models=4).compute_full(sp, x, il, nn, jl, 30)
    assert abs(e_all["eatom"].sum() - e_all["energy"]) < 1e-8
    assert abs(e_all["energy"] - e_4["energy"]) > 1e-6      # different ensembles
    params = aniw.load(ani1xnr_path)
    b = pyaev_torch.pyaev_forward(params, sp, x, domain.half_neighbor_list(x, 30), 30, num_models=4)
    assert abs(b["energy"] - e_4["energy"]) < 1e-9 * abs(b["energy"])


def test_committed_ani1xnr_goldens_reproduce(water30, ani1xnr_path):
    with open(os.path.join(ROOT, "tests", "golden", "water30_ani1xnr.json")) as f:
        g = json.load(f)
    orc = OracleANI(ani1xnr_path)
    c0 = g["cases"][0]
    r = orc.compute_half(water30["species"], water30["x"], water30["atom_index12"], 30)
    assert abs(r["energy"] - c0["energy"]) < 1e-9 * abs(c0["energy"])
    assert np.abs(r["forces"] - np.array(c0["forces"])).max() < 1e-9
    c1 = g["cases"][1]
    sub = domain.decompose(water30["x"] + 4.0, water30["species"], np.array([8.0, 8.0, 8.0]))[0]
    il, nn, jl = domain.full_neighbor_list(sub.x, sub.nlocal)
    assert sub.ntotal == c1["ntotal"] and jl.size == c1["npairs"]
    r = orc.compute_full(sub.species, sub.x, il, nn, jl, sub.nlocal)
    assert abs(r["energy"] - c1["energy"]) < 1e-9 * abs(c1["energy"])
    assert np.abs(domain.fold_ghost_forces(sub, r["forces"]) - np.array(c1["forces_folded"])).max() < 1e-9
