"""RepulsionXTB pair term (models/lammps_ani.py:186-192,301-330).  PARITY UNPINNED against the reference:
the torchani source and a model file carrying `rep_calc` are not available offline, so the oracle restates the published GFN2-xTB
functional form and is checked here by finite differences; the CUDA kernel is then checked against the oracle."""
import numpy as np
import pytest

from lammps_ani_b200 import aniw, domain, extract
from oracle.oracle import OracleANI

EV2KCAL = 23.060548


@pytest.fixture(scope="module")
def models(tmp_path_factory, ani1xnr_path):
    d = tmp_path_factory.mktemp("rep")
    base = aniw.load(ani1xnr_path)
    rep = extract.with_repulsion(base)
    p0, p1 = str(d / "base.aniw"), str(d / "rep.aniw")
    aniw.save(base, p0)
    aniw.save(rep, p1)
    return p0, p1


def _system(scale=0.9):
    x, sp, box = domain.make_water_box((3, 3, 3), seed=5)
    x, box = x * scale, np.asarray(box) * scale                    # compressed: short contacts make the repulsion matter
    sub = domain.decompose(x, sp, box)[0]
    return sub, box


def test_oracle_repulsion_energy_and_finite_differences(models):
    p0, p1 = models
    sub, box = _system()
    il, nn, jl = domain.full_neighbor_list(sub.x, sub.nlocal)
    o0, o1 = OracleANI(p0), OracleANI(p1)
    r0 = o0.compute_full(sub.species, sub.x, il, nn, jl, sub.nlocal)
    r1 = o1.compute_full(sub.species, sub.x, il, nn, jl, sub.nlocal)
    erep = r1["energy"] - r0["energy"]
    assert erep > 1.0                                              # kcal/mol: a repulsive, clearly non-zero term
    assert np.allclose(r1["eatom"], r0["eatom"])                   # added to the total only (lammps_ani.py:186-192)
    # direct evaluation of the pair sum over the periodic images
    tabs = aniw.load(p1).repulsion
    A2B = 1.8897261258369282
    e = 0.0
    for i in range(sub.nlocal):
        d = sub.x - sub.x[i]
        r = np.sqrt((d * d).sum(1))
        for j in np.nonzero((r > 0) & (r < tabs["cutoff"]))[0]:
            a, b = sub.species[i], sub.species[j]
            rb = r[j] * A2B
            fc = np.exp(1.0 - 1.0 / (1.0 - (r[j] / tabs["cutoff"]) ** 2))
            e += 0.5 * tabs["y_ab"][a, b] / rb * np.exp(-tabs["sqrt_alpha_ab"][a, b] * rb ** tabs["k_rep_ab"][a, b]) * fc
    assert abs(e * 627.5094738898777 - erep) <= 1e-9 * abs(erep)
    # forces of the folded periodic system by central differences of the total energy
    f = domain.fold_ghost_forces(sub, r1["forces"])
    h = 1e-4
    rng = np.random.default_rng(0)
    for i in rng.choice(sub.nlocal, size=4, replace=False):
        for c in range(3):
            ep = []
            for sgn in (+1, -1):
                xl = sub.x[:sub.nlocal].copy()
                xl[i, c] += sgn * h
                s2 = domain.decompose(xl, sub.species[:sub.nlocal], box)[0]
                l2 = domain.full_neighbor_list(s2.x, s2.nlocal)
                ep.append(o1.compute_full(s2.species, s2.x, *l2, s2.nlocal)["energy"])
            fd = -(ep[0] - ep[1]) / (2 * h)
            assert abs(fd - f[i, c]) <= 2e-5 * max(1.0, abs(f[i, c])), (i, c, fd, f[i, c])


@pytest.mark.gpu
@pytest.mark.parametrize("single,full", [(True, True), (False, True), (True, False), (False, False)])
def test_cuda_repulsion_matches_oracle(models, single, full):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from lammps_ani_b200 import runner
    _, p1 = models
    sub, box = _system()
    orc = OracleANI(p1)
    ani = runner.ANI(p1, 0, -1, True, full, single)
    if full:
        il, nn, jl = domain.full_neighbor_list(sub.x, sub.nlocal)
        ref = orc.compute_full(sub.species, sub.x, il, nn, jl, sub.nlocal, single_coords=single)
        e, f, eatom, vir = ani.compute(sub.species, sub.x, il, jl, nn, nlocal=sub.nlocal, ago=0, atomic=True, vflag=True)
    else:
        idx = domain.half_neighbor_list(sub.x, sub.nlocal)
        ref = orc.compute_half(sub.species, sub.x, idx, sub.nlocal, single_coords=single)
        e, f, eatom, vir = ani.compute(sub.species, sub.x, idx, nlocal=sub.nlocal, ago=0, atomic=True, vflag=True)
    if single:
        assert abs(e - ref["energy"]) <= 1e-6 * abs(ref["energy"])
        assert np.abs(f - ref["forces"]).max() <= 1e-4 * EV2KCAL
        assert np.abs(vir - ref["virial"]).max() <= 1e-4 * max(1.0, np.abs(ref["virial"]).max())
    else:
        assert abs(e - ref["energy"]) <= 1e-10 * abs(ref["energy"])
        assert np.abs(f - ref["forces"]).max() <= 1e-10 * max(1.0, np.abs(ref["forces"]).max())
        assert np.abs(vir - ref["virial"]).max() <= 1e-10 * max(1.0, np.abs(ref["virial"]).max())
    assert np.abs(eatom - ref["eatom"]).max() <= (1e-5 if single else 1e-10) * np.abs(ref["eatom"]).max()
