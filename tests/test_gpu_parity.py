"""GPU parity tests (B200): the CUDA path through the C ABI against the CPU oracle on identical inputs.

Tolerances are the north star's (BASELINE.json): fp32 total energy 1e-6 relative, per-atom forces 1e-4 eV/A
max-abs (= 2.30605e-3 kcal/mol/A); `double` mode 1e-10 relative."""
import json
import os

import numpy as np
import pytest

from lammps_ani_b200 import domain
from oracle.oracle import OracleANI

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EV2KCAL = 23.060548
F_TOL_FP32 = 1e-4 * EV2KCAL     # kcal/mol/A
E_RTOL_FP32 = 1e-6
RTOL_FP64 = 1e-10


@pytest.fixture(scope="module")
def ANI():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from lammps_ani_b200 import build, runner
    build.build()
    return runner.ANI


def periodic_system(x, sp, box):
    sub = domain.decompose(x, sp, np.asarray(box, dtype=float))[0]
    il, nn, jl = domain.full_neighbor_list(sub.x, sub.nlocal)
    return sub, il, nn, jl


def check(gpu, ref, single, nlocal, f_scale=1.0):
    e, f, eatom, virial = gpu
    if single:
        assert abs(e - ref["energy"]) <= E_RTOL_FP32 * abs(ref["energy"])
        assert np.abs(f - ref["forces"]).max() <= F_TOL_FP32 * f_scale
        if eatom is not None:
            assert np.abs(eatom - ref["eatom"]).max() <= 1e-5 * np.abs(ref["eatom"]).max()
        if virial is not None:
            assert np.abs(virial - ref["virial"]).max() <= 1e-4 * max(1.0, np.abs(ref["virial"]).max())   # tcgen05 accumulates with truncation
    else:
        assert abs(e - ref["energy"]) <= RTOL_FP64 * abs(ref["energy"])
        assert np.abs(f - ref["forces"]).max() <= RTOL_FP64 * max(1.0, np.abs(ref["forces"]).max())
        if eatom is not None:
            assert np.abs(eatom - ref["eatom"]).max() <= RTOL_FP64 * np.abs(ref["eatom"]).max()
        if virial is not None:
            assert np.abs(virial - ref["virial"]).max() <= RTOL_FP64 * max(1.0, np.abs(ref["virial"]).max())


@pytest.mark.parametrize("single", [False, True])
def test_water30_reference_half_list(ANI, water30, ani1xnr_path, single):
    """The reference's own known-answer input (test_model.cpp:84-120): non-periodic, half list, all members."""
    orc = OracleANI(ani1xnr_path)
    ref = orc.compute_half(water30["species"], water30["x"], water30["atom_index12"], 30, single_coords=single)
    ani = ANI(ani1xnr_path, 0, -1, use_cuaev=True, use_fullnbr=False, use_single=single)
    out = ani.compute(water30["species"], water30["x"], water30["atom_index12"], nlocal=30, ago=0, atomic=True, vflag=True)
    check(out, ref, single, 30)
    with open(os.path.join(ROOT, "tests", "golden", "water30_ani1xnr.json")) as f:
        g = json.load(f)["cases"][0]
    tol = E_RTOL_FP32 if single else RTOL_FP64
    assert abs(out[0] - g["energy"]) <= tol * abs(g["energy"])
    assert np.abs(out[1] - np.array(g["forces"])).max() <= (F_TOL_FP32 if single else 1e-8)


def active_columns(species, S, nR, ang_sub):
    """Columns of the full AEV that belong to species (pairs) present in the system: the library drops the others
    (structurally zero AEV entries), so its dE/dAEV is defined on these columns only."""
    act = sorted(set(int(s) for s in species))
    triu = {}
    p = 0
    for a in range(S):
        for b in range(a, S):
            triu[(a, b)] = p
            p += 1
    cols = [a * nR + k for a in act for k in range(nR)]
    for i, a in enumerate(act):
        for b in act[i:]:
            cols += [S * nR + triu[(a, b)] * ang_sub + q for q in range(ang_sub)]
    return np.array(sorted(cols))


@pytest.mark.parametrize("dense", [False, True])
@pytest.mark.parametrize("single", [False, True])
@pytest.mark.parametrize("model", ["ani1xnr", "ani2x_synth"])
def test_water30_periodic_full_list_with_ghosts(ANI, water30, ani1xnr_path, ani2x_synth_path, single, model, dense, monkeypatch):
    path = ani1xnr_path if model == "ani1xnr" else ani2x_synth_path
    monkeypatch.setenv("ANI_B200_DENSE_AEV", "1" if dense else "0")
    sub, il, nn, jl = periodic_system(water30["x"] + 4.0, water30["species"], [8.0, 8.0, 8.0])
    ref = OracleANI(path).compute_full(sub.species, sub.x, il, nn, jl, sub.nlocal, single_coords=single, want_aev=True,
                                       want_daev=True)
    ani = ANI(path, 0, -1, True, True, single)
    out = ani.compute(sub.species, sub.x, il, jl, nn, nlocal=sub.nlocal, ago=0, atomic=True, vflag=True)
    aev, daev = ani.debug_aev()
    if not dense:   # dE/dAEV exists only on the columns of the species present (H, O)
        S, nR = (4, 32) if model == "ani1xnr" else (7, 16)
        cols = active_columns(sub.species, S, nR, 32)
        assert len(cols) == 2 * nR + 3 * 32
        off = np.setdiff1d(np.arange(aev.shape[1]), cols)
        assert not aev[:, off].any() and not ref["aev"][:, off].any() and not daev[:, off].any()
        ref["daev"][:, off] = 0.0
    if single:
        assert np.abs(aev - ref["aev"]).max() <= 2e-5 * max(1.0, np.abs(ref["aev"]).max())
        assert np.abs(daev - ref["daev"]).max() <= 2e-5 * max(1.0, np.abs(ref["daev"]).max())
    else:
        assert np.abs(aev - ref["aev"]).max() <= 1e-12 * max(1.0, np.abs(ref["aev"]).max())
        assert np.abs(daev - ref["daev"]).max() <= 1e-11 * max(1.0, np.abs(ref["daev"]).max())
    check(out, ref, single, sub.nlocal)
    # ago != 0: cached neighbour data, slightly moved coordinates (ani.cpp:213-229 semantics)
    x2 = sub.x + 1e-3 * np.sin(np.arange(sub.x.size).reshape(sub.x.shape))
    # ghosts must move with their owners
    x2[sub.nlocal:] = sub.x[sub.nlocal:] + (x2[sub.ghost_owner_index] - sub.x[sub.ghost_owner_index])
    ref2 = OracleANI(path).compute_full(sub.species, x2, il, nn, jl, sub.nlocal, single_coords=single)
    out2 = ani.compute(None, x2, nlocal=sub.nlocal, ago=1, atomic=True, vflag=True)
    check(out2, ref2, single, sub.nlocal)


@pytest.mark.parametrize("single", [False, True])
def test_water_box_2k_atoms(ANI, ani2x_synth_path, single):
    """ANI-2x geometry, all 8 members, 2 160-atom periodic water box (8x9x10 molecules)."""
    x, sp, box = domain.make_water_box((8, 9, 10), seed=7)
    sub, il, nn, jl = periodic_system(x, sp, box)
    ref = OracleANI(ani2x_synth_path).compute_full(sub.species, sub.x, il, nn, jl, sub.nlocal, single_coords=single)
    ani = ANI(ani2x_synth_path, 0, -1, True, True, single)
    out = ani.compute(sub.species, sub.x, il, jl, nn, nlocal=sub.nlocal, ago=0, atomic=True, vflag=True)
    check(out, ref, single, sub.nlocal)
    # folded forces sum to zero (translation invariance) -- a size-independent property
    f = domain.fold_ghost_forces(sub, out[1])
    assert np.abs(f.sum(0)).max() < (2e-2 if single else 1e-7)


def test_half_equals_full_and_num_models(ANI, ani1xnr_path):
    x, sp, box = domain.make_water_box((4, 4, 4), seed=3)
    sub, il, nn, jl = periodic_system(x, sp, box)
    half = domain.half_neighbor_list(sub.x, sub.nlocal)
    full = ANI(ani1xnr_path, 0, 4, True, True, False).compute(sub.species, sub.x, il, jl, nn, nlocal=sub.nlocal, vflag=True)
    hal = ANI(ani1xnr_path, 0, 4, True, False, False).compute(sub.species, sub.x, half, nlocal=sub.nlocal, vflag=True)
    assert abs(full[0] - hal[0]) < 1e-10 * abs(full[0])
    assert np.abs(full[1] - hal[1]).max() < 1e-8
    ref = OracleANI(ani1xnr_path, use_num_models=4).compute_full(sub.species, sub.x, il, nn, jl, sub.nlocal)
    check(full, ref, False, sub.nlocal)
    allm = ANI(ani1xnr_path, 0, -1, True, True, False).compute(sub.species, sub.x, il, jl, nn, nlocal=sub.nlocal)
    assert abs(allm[0] - full[0]) > 1e-6


def test_pyaev_mode_has_no_radial_refilter(ANI, ani1xnr_path, water30):
    """use_cuaev=False reproduces the behaviour the reference's pyaev goldens encode."""
    x, sp = water30["x"], water30["species"]
    ref = OracleANI(ani1xnr_path).compute_half(sp, x, water30["atom_index12"], 30, radial_filter=False)
    out = ANI(ani1xnr_path, 0, -1, False, False, False).compute(sp, x, water30["atom_index12"], nlocal=30, vflag=True)
    check(out, ref, False, 30)


def test_mixed_species_buckets(ANI, ani2x_synth_path):
    """All 7 ANI-2x species present, some buckets tiny or empty-then-filled (alanine-dipeptide-like mix)."""
    rng = np.random.default_rng(5)
    x, sp, box = domain.make_water_box((5, 5, 5), seed=11)
    sp = sp.copy()
    pick = rng.choice(len(sp), size=40, replace=False)
    sp[pick] = rng.integers(0, 7, size=40)
    sub, il, nn, jl = periodic_system(x, sp, box)
    for single in (False, True):
        ref = OracleANI(ani2x_synth_path).compute_full(sub.species, sub.x, il, nn, jl, sub.nlocal, single_coords=single)
        out = ANI(ani2x_synth_path, 0, -1, True, True, single).compute(sub.species, sub.x, il, jl, nn, nlocal=sub.nlocal,
                                                                      atomic=True, vflag=True)
        check(out, ref, single, sub.nlocal)


@pytest.mark.parametrize("single", [False, True])
def test_active_species_subset_and_growth(ANI, ani2x_synth_path, single):
    """Alanine-dipeptide-like mix (H, C, N, O of ANI-2x's 7 species: 384 of 1008 AEV columns active), then the same context
    gets a rebuild in which a new species (S) appears: the column layout must grow and stay exact."""
    rng = np.random.default_rng(17)
    x, sp, box = domain.make_water_box((5, 5, 4), seed=3)
    sp = sp.copy()
    pick = rng.choice(len(sp), size=30, replace=False)
    sp[pick] = rng.choice([1, 2], size=30)            # C, N among the H (0) / O (3) of the water
    ani = ANI(ani2x_synth_path, 0, -1, True, True, single)
    orc = OracleANI(ani2x_synth_path)
    sub, il, nn, jl = periodic_system(x, sp, box)
    ref = orc.compute_full(sub.species, sub.x, il, nn, jl, sub.nlocal, single_coords=single)
    out = ani.compute(sub.species, sub.x, il, jl, nn, nlocal=sub.nlocal, ago=0, atomic=True, vflag=True)
    check(out, ref, single, sub.nlocal)
    sp2 = sp.copy()
    sp2[rng.choice(len(sp), size=5, replace=False)] = 4   # S
    sub, il, nn, jl = periodic_system(x, sp2, box)
    ref = orc.compute_full(sub.species, sub.x, il, nn, jl, sub.nlocal, single_coords=single)
    out = ani.compute(sub.species, sub.x, il, jl, nn, nlocal=sub.nlocal, ago=0, atomic=True, vflag=True)
    check(out, ref, single, sub.nlocal)
    # and back to the smaller set: the (now larger) layout stays valid
    sub, il, nn, jl = periodic_system(x, sp, box)
    ref = orc.compute_full(sub.species, sub.x, il, nn, jl, sub.nlocal, single_coords=single)
    out = ani.compute(sub.species, sub.x, il, jl, nn, nlocal=sub.nlocal, ago=0, atomic=True, vflag=True)
    check(out, ref, single, sub.nlocal)


def test_stable_caller_buffers_are_pinned_in_place(ANI, ani2x_synth_path):
    """x and out_force at a stable address (LAMMPS's atom->x and the pair style's out_force vector) are page-locked in place
    from the third call on and moved by DMA: results must be bit-identical to the staged path."""
    x, sp, box = domain.make_water_box((6, 6, 5), seed=9)
    sub, il, nn, jl = periodic_system(x, sp, box)
    ani = ANI(ani2x_synth_path, 0, -1, True, True, True)
    ani.set_option("pin_user", 1)      # opt-in: the caller promises release_host_buffers() before the arrays go away
    e0, f0, _, _ = ani.compute(sub.species, sub.x.copy(), il, jl, nn, nlocal=sub.nlocal, ago=0)
    xs = np.ascontiguousarray(sub.x, dtype=np.float64)
    fs = np.full((sub.x.shape[0], 3), np.nan)
    for k in range(5):
        e, f, _, _ = ani.compute(None, xs, nlocal=sub.nlocal, ago=1, out_force=fs)
        assert f is fs
        # forces leave the scatter kernel through fp32 atomics: equal up to summation order
        assert abs(e - e0) <= 1e-9 * abs(e0) and np.abs(fs - f0).max() <= 1e-4, f"call {k}"
        fs[:] = np.nan
    xs[:sub.nlocal] += 1e-3                       # in-place update of the caller's array must be picked up
    xs[sub.nlocal:] += 1e-3
    e1, _, _, _ = ani.compute(None, xs, nlocal=sub.nlocal, ago=1, out_force=fs)
    e2, f2, _, _ = ani.compute(None, xs.copy(), nlocal=sub.nlocal, ago=1)
    assert abs(e1 - e2) <= 1e-9 * abs(e2) and np.abs(fs - f2).max() <= 1e-4 and abs(e1 - e0) > 1e-6


@pytest.mark.parametrize("single", [False, True])
def test_two_pass_aev_defers_rows_that_outgrow_their_buffers(ANI, ani2x_synth_path, single):
    """The AEV kernels size their first-pass shared-memory buffers from the survivor counts of the previous steps and hand a row
    that does not fit to a second pass with list-sized buffers.  Here the SAME list is first used on an expanded geometry (few
    neighbours inside Rcr / Rca -> small buffers), then on the original one (every atom outgrows them): energies, forces and
    virial must equal those of a fresh context, whose first call is single-pass."""
    x, sp, box = domain.make_water_box((6, 6, 5), seed=21)
    sub, il, nn, jl = periodic_system(x, sp, box)
    ani = ANI(ani2x_synth_path, 0, -1, True, True, single)
    far = sub.x * 1.35                                     # every pair inside 5.1 A here is inside 3.8 A there: still listed
    ani.compute(sub.species, far, il, jl, nn, nlocal=sub.nlocal, ago=0)
    for _ in range(2):
        ani.compute(None, far, nlocal=sub.nlocal, ago=1)
    e, f, _, v = ani.compute(None, sub.x, nlocal=sub.nlocal, ago=1, vflag=True)
    launches_two_pass = ani.last_timing()[1]
    fresh = ANI(ani2x_synth_path, 0, -1, True, True, single)
    e0, f0, _, v0 = fresh.compute(sub.species, sub.x, il, jl, nn, nlocal=sub.nlocal, ago=0, vflag=True)
    # one extra AEV launch per direction (+1 when the force array happened to sit at a stable address and is widened on the device)
    assert launches_two_pass - fresh.last_timing()[1] in (2, 3)
    tol = 1e-4 if single else 1e-9
    assert abs(e - e0) <= 1e-9 * abs(e0) and np.abs(f - f0).max() <= tol and np.abs(v - v0).max() <= 1e-3 * max(1.0, np.abs(v0).max())
    e1, f1, _, _ = ani.compute(None, sub.x, nlocal=sub.nlocal, ago=1)          # buffers have caught up: no deferred rows now
    assert abs(e1 - e0) <= 1e-9 * abs(e0) and np.abs(f1 - f0).max() <= tol


def test_error_paths(ANI, ani1xnr_path, water30):
    ani = ANI(ani1xnr_path, 0)
    sp = water30["species"].copy()
    sp[3] = 6     # species outside the 4-species model
    il, nn, jl = domain.full_neighbor_list(water30["x"], 30)
    with pytest.raises(RuntimeError):
        ani.compute(sp, water30["x"], il, jl, nn, nlocal=30)
    with pytest.raises(RuntimeError):
        ANI(ani1xnr_path, 0, 9)        # more members than the ensemble has
    with pytest.raises(RuntimeError):
        ANI(ani1xnr_path, -1)          # "cpu"
    # empty system
    e, f, _, _ = ani.compute(np.zeros(0, dtype=np.int32), np.zeros((0, 3)), np.zeros(0, np.int32), np.zeros(0, np.int32),
                             np.zeros(0, np.int32), nlocal=0)
    assert e == 0.0 and f.shape == (0, 3)


@pytest.mark.parametrize("layout_left", [False, True])
def test_padded_neighbor_table_is_bit_identical_to_flat_list(ANI, ani1xnr_path, layout_left):
    """ani/kk hand-off (pair_ani_kokkos.cpp:143-169): the padded 2-D table, de-padded on the device, must give
    exactly the flat list's results (integer work: bit-exact), with special bits set in the table entries."""
    x, sp, box = domain.make_water_box((4, 4, 5), seed=9)
    sub, il, nn, jl = periodic_system(x, sp, box)
    maxn = int(nn.max()) + 3
    nrows = sub.nlocal + 5                                   # kokkos_nlocal > nlocal on multi-domain runs
    table = np.full((nrows, maxn), 0x7FFFFFFF, dtype=np.int32)
    off = np.concatenate([[0], np.cumsum(nn)])
    for i in range(sub.nlocal):
        table[i, :nn[i]] = jl[off[i]:off[i + 1]]
    table[: sub.nlocal: 7, 0] |= (1 << 30)                   # special-bond bits above NEIGHMASK
    a = ANI(ani1xnr_path, 0, -1, True, True, False)
    ref = a.compute(sub.species, sub.x, il, jl, nn, nlocal=sub.nlocal, ago=0, atomic=True, vflag=True)
    b = ANI(ani1xnr_path, 0, -1, True, True, False)
    b.set_neighbors_table(sub.species + 1, il, nn, table, sub.nlocal, layout_left=layout_left)
    out = b.compute(None, sub.x, nlocal=sub.nlocal, ago=1, atomic=True, vflag=True)
    assert out[0] == ref[0]
    assert np.array_equal(out[2], ref[2])
    assert np.abs(out[1] - ref[1]).max() < 1e-9             # force atomics are order-dependent in the last bits


@pytest.mark.parametrize("x_is_float,f_is_float", [(0, 0), (1, 0), (0, 1), (1, 1)])
def test_compute_device_matches_oracle(ANI, ani2x_synth_path, x_is_float, f_is_float):
    """The device-tensor entry (`ANI::compute` tensor overload, ani.cpp:268-316 -> ani_compute_device): x, f and eatom are
    device arrays, f and eatom are ACCUMULATED into (`f +=`, pair_ani_kokkos.cpp:190-191), in both Kokkos precisions."""
    import torch
    x, sp, box = domain.make_water_box((6, 6, 6), seed=13)
    sub, il, nn, jl = periodic_system(x, sp, box)
    ref = OracleANI(ani2x_synth_path).compute_full(sub.species, sub.x, il, nn, jl, sub.nlocal, single_coords=True)
    ani = ANI(ani2x_synth_path, 0, -1, True, True, True)
    ani.set_neighbors_full(sub.species, il, nn, jl, sub.nlocal)
    dev = torch.device("cuda", 0)
    xd = torch.tensor(sub.x, dtype=torch.float32 if x_is_float else torch.float64, device=dev)
    f0 = torch.full((sub.x.shape[0], 3), 0.25, dtype=torch.float32 if f_is_float else torch.float64, device=dev)
    e0 = torch.full((sub.nlocal,), 1.5, dtype=torch.float64, device=dev)
    f, ea = f0.clone(), e0.clone()
    torch.cuda.synchronize()
    e, vir = ani.compute_device(xd.data_ptr(), bool(x_is_float), f.data_ptr(), bool(f_is_float), ea.data_ptr(), want_energy=True,
                                want_virial=True)
    ani.synchronize()
    df = (f - f0).double().cpu().numpy()
    dea = (ea - e0).cpu().numpy()
    assert abs(e - ref["energy"]) <= E_RTOL_FP32 * abs(ref["energy"])
    assert np.abs(df - ref["forces"]).max() <= F_TOL_FP32 + (2e-5 if f_is_float else 0.0)   # + fp32 rounding of 0.25 + f
    assert np.abs(dea - ref["eatom"]).max() <= 1e-5 * np.abs(ref["eatom"]).max()
    assert np.abs(vir - ref["virial"]).max() <= 1e-4 * max(1.0, np.abs(ref["virial"]).max())
    # a second call without energy / virial (the off-thermo step of the device-resident styles): forces accumulate again
    ani.compute_device(xd.data_ptr(), bool(x_is_float), f.data_ptr(), bool(f_is_float), 0, want_energy=False)
    ani.synchronize()
    df2 = (f - f0).double().cpu().numpy()
    assert np.abs(df2 - 2.0 * ref["forces"]).max() <= 2 * F_TOL_FP32 + (4e-5 if f_is_float else 0.0)


@pytest.mark.parametrize("repulsion", [False, True])
def test_combustion_config5_half_list(ANI, ani1xnr_path, tmp_path, repulsion):
    """BASELINE.json config 5 at parity size: CH4 + 2 O2 at 0.25 g/cc (19 998 atoms of the reference's 300 006-atom recipe,
    examples/combustion/prepare_system/system-300k.inp), real ANI-1xnr weights, HALF neighbour list (`ani_neighbor half`,
    13 stored pairs per atom), fp32, with and without the RepulsionXTB term of the `*_repulsion` models; H, C, O present:
    288 active AEV columns, a 2 222-atom C bucket."""
    from lammps_ani_b200 import aniw, extract
    path = ani1xnr_path
    if repulsion:
        path = str(tmp_path / "ani1x_nr_repulsion.aniw")
        aniw.save(extract.with_repulsion(aniw.load(ani1xnr_path)), path)
    x, sp, box = domain.make_combustion_box(2222, seed=3)
    sub = domain.decompose(x, sp, box)[0]
    idx = domain.half_neighbor_list(sub.x, sub.nlocal)
    assert 12.0 < idx.shape[1] / sub.nlocal < 14.5
    ref = OracleANI(path).compute_half(sub.species, sub.x, idx, sub.nlocal, single_coords=True)
    out = ANI(path, 0, -1, True, False, True).compute(sub.species, sub.x, idx, nlocal=sub.nlocal, ago=0, atomic=True, vflag=True)
    check(out, ref, True, sub.nlocal)


def test_library_halo_plan_matches_torch_reference(ANI, ani2x_synth_path):
    """ani_halo_forward / ani_halo_reverse (csrc/halo.cuh: halo_pack_kernel, the self-image device copy, halo_unpack_add_kernel)
    on one rank against the torch index-op reference of ddstep.DomainStep: ghost positions after moving the local atoms, and
    the fold of ghost forces onto their owners (an owner appears several times: atomics)."""
    import torch
    from lammps_ani_b200 import ddstep
    x, sp, box = domain.make_water_box((5, 4, 4), seed=8)
    sub = domain.decompose(x, sp, box)[0]
    halo = domain.replica_halo(sub, box, (1, 1, 1), 0)
    il, nn, jl = domain.full_neighbor_list(sub.x, sub.nlocal)
    ani = ANI(ani2x_synth_path, 0, -1, True, True, True)
    dev = torch.device("cuda", 0)
    g = ddstep.GpuDomainStep(ani, sub, halo, dev, il, nn, jl)
    c = ddstep.DomainStep(sub, halo, torch.device("cpu"))
    rng = np.random.default_rng(1)
    dx = torch.tensor(rng.normal(scale=0.05, size=(sub.nlocal, 3)))
    g.x[:sub.nlocal] += dx.to(dev); c.x[:sub.nlocal] += dx
    g.forward_halo(); c.forward_halo()
    ani.synchronize()
    assert torch.equal(g.x.cpu(), c.x)
    ff = torch.tensor(rng.normal(size=(sub.ntotal, 3)))
    g.f.copy_(ff.to(dev)); c.f.copy_(ff)
    g.reverse_halo(); c.reverse_halo()
    ani.synchronize()
    assert (g.f[:sub.nlocal].cpu() - c.f[:sub.nlocal]).abs().max() < 1e-12
    # and the single-domain emulation used by bench.py's multi-GPU force check
    g.x[:sub.nlocal] += 0.01; c.x[:sub.nlocal] += 0.01
    g.forward_halo_local(); c.forward_halo()
    ani.synchronize()
    assert (g.x.cpu() - c.x).abs().max() < 1e-12


@pytest.mark.parametrize("repulsion", [False, True])
def test_overlapped_halo_step_matches_serial(ANI, ani2x_synth_path, ani1xnr_path, tmp_path, repulsion):
    """ani_compute_device_halo (exchange on a second stream under the interior atoms' AEV kernels; needs the device-built list's
    interior / boundary split) against the same step as three serial calls, with the host-installed list (always serial), and
    with the overlap switched off: forces within the fp32 summation-order noise, energy within 1e-9 relative.  One rank: every
    ghost is a periodic self image, so the exchange is device copies -- the stream / event choreography and the split of the AEV
    launches are what is exercised (bench.py's multi-GPU force check covers the NCCL leg)."""
    import torch
    from lammps_ani_b200 import aniw, ddstep, extract
    path = ani2x_synth_path
    if repulsion:
        path = str(tmp_path / "ani1x_nr_repulsion.aniw")
        aniw.save(extract.with_repulsion(aniw.load(ani1xnr_path)), path)
        x, sp, box = domain.make_combustion_box(3000, seed=5)
    else:
        x, sp, box = domain.make_water_box((8, 7, 7), seed=4)
    sub = domain.decompose(x, sp, box)[0]
    halo = domain.replica_halo(sub, box, (1, 1, 1), 0)
    il, nn, jl = domain.full_neighbor_list(sub.x, sub.nlocal)
    ani = ANI(path, 0, -1, True, True, True)
    dev = torch.device("cuda", 0)
    g = ddstep.GpuDomainStep(ani, sub, halo, dev, il, nn, jl)
    nl = sub.nlocal

    def run(fn):
        out = []
        for _ in range(3):                                   # (the third step runs the two-pass AEV configuration)
            e = fn(True)
            ani.synchronize()
            out.append((e, g.f[:nl].cpu().numpy().copy()))
        return out[-1]

    e_ref, f_ref = run(g.step_gpu_serial)                    # host list, three calls
    e_h, f_h = run(g.step_gpu)                               # host list, one call (serialised inside: no split known)
    g.use_device_list(7.1)
    e_o, f_o = run(g.step_gpu)                               # device list: overlapped
    ani.set_option("halo_overlap", 0)
    e_s, f_s = run(g.step_gpu)                               # device list, overlap off
    ani.set_option("halo_overlap", 1)
    for e, f in ((e_h, f_h), (e_o, f_o), (e_s, f_s)):
        assert abs(e - e_ref) <= 1e-9 * abs(e_ref)
        assert np.abs(f - f_ref).max() <= 2e-3               # kcal/mol/A: 1e-4 eV/A
    assert np.abs(f_ref).max() > 1.0
