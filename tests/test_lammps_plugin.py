"""End-to-end drop-in check: a stock serial LAMMPS (28 Mar 2023, built from the in-tree sources with PKG_PLUGIN; binary in
tools/_lmp/, not in git) loads lammps_ani_b200/ani_plugin.so and runs `pair_style ani` on the reference's 30-atom water
case (tests/lammps/in.ani mirrors /root/reference/tests/in.lammps).  Rows a1-a3 + b of SURVEY.md section 8: list
flattening on rebuild steps, ghost-force fold-back through comm->reverse_comm(Pair*), energy / force / virial hand-off."""
import os
import re
import subprocess

import numpy as np
import pytest

from lammps_ani_b200 import domain
from oracle.oracle import OracleANI

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LMP = os.path.join(ROOT, "tools", "_lmp", "lmp")
PLUGIN = os.path.join(ROOT, "lammps_ani_b200", "ani_plugin.so")
CASE = os.path.join(ROOT, "tests", "lammps")
EV2KCAL = 23.060548

needs_lmp = pytest.mark.skipif(not (os.path.exists(LMP) and os.path.exists(PLUGIN)),
                               reason="LAMMPS binary / plugin not built (tools/_lmp, lammps_ani_b200/ani_plugin.so)")


def run_lmp(tmp_path, model, *variables, data="water30.data", script="in.ani"):
    env = dict(os.environ, LD_LIBRARY_PATH=os.path.dirname(LMP) + ":" + os.environ.get("LD_LIBRARY_PATH", ""))
    cmd = [LMP, "-in", os.path.join(CASE, script), "-log", "none", "-var", "plugin", PLUGIN, "-var", "data_file",
           os.path.join(CASE, data), "-var", "ani_model_file", model]
    for k, v in variables:
        cmd += ["-var", k, str(v)]
    return subprocess.run(cmd, cwd=tmp_path, env=env, capture_output=True, text=True, timeout=600)


@needs_lmp
def test_plugin_registers_and_refuses_cpu(tmp_path, ani1xnr_path):
    import torch
    r = run_lmp(tmp_path, ani1xnr_path, ("steps", 0))
    assert "Loading plugin: ANI pair style" in r.stdout
    if not torch.cuda.is_available():
        assert "no CUDA device available" in r.stdout + r.stderr       # fails loudly: no CPU fallback


def _thermo(stdout):
    rows = []
    for line in stdout.splitlines():
        f = line.split()
        if len(f) == 10 and re.fullmatch(r"\d+", f[0]):
            rows.append([float(v) for v in f])
    return np.array(rows)


def _dump_forces(path):
    frames, cur = [], None
    with open(path) as fh:
        lines = fh.read().splitlines()
    i = 0
    while i < len(lines):
        if lines[i].startswith("ITEM: NUMBER OF ATOMS"):
            n = int(lines[i + 1])
        if lines[i].startswith("ITEM: ATOMS"):
            cur = np.array([[float(v) for v in l.split()] for l in lines[i + 1:i + 1 + n]])
            frames.append(cur[np.argsort(cur[:, 0])])
            i += n
        i += 1
    return frames


NKTV2P = 68568.415          # LAMMPS `units real`: pressure [atm] = nktv2p * energy [kcal/mol] / volume [A^3]  (update.cpp)


def _box_of(data_file):
    lo, hi = np.zeros(3), np.zeros(3)
    with open(os.path.join(CASE, data_file)) as fh:
        for line in fh:
            f = line.split()
            for d, tag in enumerate(("xlo", "ylo", "zlo")):
                if len(f) >= 4 and f[2] == tag:
                    lo[d], hi[d] = float(f[0]), float(f[1])
            if line.startswith("Atoms"):
                break
    return lo, hi


def check_frame_against_oracle(orc, fr, th_row, lo, hi, single, cutoff=7.1):
    """One dumped frame (id type x y z fx fy fz, sorted by id) + its thermo row (step pe ke etotal pxx pyy pzz pxy pxz pyz)
    against the oracle on the same coordinates: energy, forces with the ghost forces folded back, and the pair virial
    LAMMPS reports as c_press_pair (pair_ani.cpp:224-231: xx, yy, zz, xy, xz, yz)."""
    L = hi - lo
    x, sp = fr[:, 2:5], fr[:, 1].astype(np.int32) - 1
    xw = x - lo
    xw = xw - np.floor(xw / L) * L
    sub = domain.decompose(xw, sp, L, cut=cutoff)[0]
    il, nn, jl = domain.full_neighbor_list(sub.x, sub.nlocal, cut=cutoff)
    ref = orc.compute_full(sub.species, sub.x, il, nn, jl, sub.nlocal, single_coords=single)
    f_ref = domain.fold_ghost_forces(sub, ref["forces"])
    f_ref_by_id = np.empty_like(f_ref)
    f_ref_by_id[sub.global_id[:sub.nlocal]] = f_ref
    pe = th_row[1]
    vol = float(np.prod(L))
    W = ref["virial"]
    p_ref = np.array([W[0, 0], W[1, 1], W[2, 2], W[0, 1], W[0, 2], W[1, 2]]) * NKTV2P / vol
    p_lmp = th_row[4:10]
    pscale = max(1.0, np.abs(p_ref).max())
    if single:
        assert abs(pe - ref["energy"]) <= 1e-6 * abs(ref["energy"])
        assert np.abs(fr[:, 5:8] - f_ref_by_id).max() <= 1e-4 * EV2KCAL
        assert np.abs(p_lmp - p_ref).max() <= 2e-4 * pscale, (p_lmp, p_ref)
    else:
        assert abs(pe - ref["energy"]) <= 1e-10 * abs(ref["energy"])
        assert np.abs(fr[:, 5:8] - f_ref_by_id).max() <= 1e-8
        assert np.abs(p_lmp - p_ref).max() <= 1e-9 * pscale, (p_lmp, p_ref)
    return ref


@pytest.mark.gpu
@needs_lmp
@pytest.mark.parametrize("nbr,prec", [("full", "single"), ("half", "double"), ("full", "double"), ("device", "single"),
                                      ("device", "double")])
def test_lammps_run_matches_oracle(tmp_path, ani1xnr_path, nbr, prec):
    r = run_lmp(tmp_path, ani1xnr_path, ("ani_neighbor", nbr), ("ani_precision", prec), ("steps", 4))
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    th = _thermo(r.stdout)
    frames = _dump_forces(os.path.join(tmp_path, "dump.forces"))
    assert th.shape[0] == 5 and len(frames) == 5
    orc = OracleANI(ani1xnr_path)
    lo, hi = _box_of("water30.data")
    for step in (0, 4):                       # step 0: fresh list; step 4: after two rebuilds and four NVE updates
        check_frame_against_oracle(orc, frames[step], th[step], lo, hi, prec == "single")
    # NVE over 4 x 0.1 fs: total energy conserved to the integrator's accuracy
    assert abs(th[4, 3] - th[0, 3]) < 5e-3


@pytest.mark.gpu
@needs_lmp
@pytest.mark.parametrize("nbr,prec", [("full", "single"), ("device", "single"), ("half", "double")])
def test_alanine_dipeptide_config3(tmp_path, ani2x_synth_path, nbr, prec):
    """BASELINE.json config 3: the reference's solvated alanine dipeptide (examples/alanine-dipeptide/alanine-dipeptide.data,
    1 912 atoms: H 1 272, O 632, C 6, N 2 -- buckets of 6 and 2 atoms, 384 active AEV columns) through the LAMMPS plugin, all 8
    members, 4 NVE steps; ANI-2x geometry with seeded weights (the trained parameters are not available offline), the oracle
    evaluating the very same numbers."""
    if not os.path.exists(os.path.join(CASE, "alanine-dipeptide.data")):
        pytest.skip("fixture not generated (tools/make_golden.py)")
    r = run_lmp(tmp_path, ani2x_synth_path, ("ani_neighbor", nbr), ("ani_precision", prec), ("steps", 4), data="alanine-dipeptide.data")
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    th = _thermo(r.stdout)
    frames = _dump_forces(os.path.join(tmp_path, "dump.forces"))
    assert th.shape[0] == 5 and len(frames) == 5 and frames[0].shape[0] == 1912
    orc = OracleANI(ani2x_synth_path)
    lo, hi = _box_of("alanine-dipeptide.data")
    for step in (0, 4):
        check_frame_against_oracle(orc, frames[step], th[step], lo, hi, prec == "single")


@pytest.mark.gpu
@needs_lmp
@pytest.mark.parametrize("nbr,prec", [("full", "single"), ("half", "double")])
def test_restart_round_trip(tmp_path, ani1xnr_path, nbr, prec):
    """write_restart -> clear -> read_restart -> pair_coeff * * -> run 0 (examples/water/README.md:51-66; the restart record of
    pair_ani.cpp:408-455, exercised by LAMMPS's own test_pair_style.cpp:426-437): the restarted run reproduces energy, forces
    and pressure of the run that wrote the file, and the pair style comes back with the same settings."""
    r = run_lmp(tmp_path, ani1xnr_path, ("ani_neighbor", nbr), ("ani_precision", prec), ("steps", 2), script="in.restart")
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    th = _thermo(r.stdout)
    frames = _dump_forces(os.path.join(tmp_path, "dump.forces")) + _dump_forces(os.path.join(tmp_path, "dump.restart"))
    assert th.shape[0] == 4 and len(frames) == 4       # steps 0, 1, 2 of the first run + step 2 again after the restart
    a, b = th[2], th[3]
    tol = 1e-6 if prec == "single" else 1e-12
    assert abs(a[1] - b[1]) <= tol * abs(a[1])
    assert np.abs(a[4:10] - b[4:10]).max() <= max(tol, 1e-9) * max(1.0, np.abs(a[4:10]).max()) * 100
    ftol = 1e-4 * EV2KCAL if prec == "single" else 1e-8
    assert np.abs(frames[2][:, 5:8] - frames[3][:, 5:8]).max() <= ftol
    assert np.abs(frames[2][:, 2:5] - frames[3][:, 2:5]).max() <= 1e-9
    orc = OracleANI(ani1xnr_path)
    lo, hi = _box_of("water30.data")
    check_frame_against_oracle(orc, frames[3], th[3], lo, hi, prec == "single")


@pytest.mark.gpu
@needs_lmp
@pytest.mark.parametrize("prec", ["double", "single"])
def test_device_resident_integrator_matches_fix_nve(tmp_path, ani1xnr_path, prec):
    """`fix ani/md` (velocity Verlet on the GPU, x / v / f resident between rebuilds) against LAMMPS's own
    `fix nve` on the same input: 20 steps of 0.1 fs with a rebuild every 2 steps (so the upload-after-rebuild path runs ten
    times), forces / positions dumped every step, thermo (pe, ke from the downloaded v, pair pressure) every step."""
    base = (("ani_neighbor", "device"), ("ani_precision", prec), ("steps", 20))
    d0, d1 = tmp_path / "nve", tmp_path / "md"
    d0.mkdir(); d1.mkdir()
    r0 = run_lmp(d0, ani1xnr_path, *base, ("integrator", "nve"))
    r1 = run_lmp(d1, ani1xnr_path, *base, ("integrator", "ani/md"))
    assert r0.returncode == 0, r0.stdout[-2000:] + r0.stderr[-2000:]
    assert r1.returncode == 0, r1.stdout[-2000:] + r1.stderr[-2000:]
    t0, t1 = _thermo(r0.stdout), _thermo(r1.stdout)
    f0, f1 = _dump_forces(os.path.join(d0, "dump.forces")), _dump_forces(os.path.join(d1, "dump.forces"))
    assert t0.shape == t1.shape == (21, 10) and len(f0) == len(f1) == 21
    xtol, ftol, etol = (1e-10, 1e-8, 1e-11) if prec == "double" else (2e-6, 2e-4 * EV2KCAL, 2e-7)
    for k in range(21):
        assert np.abs(f0[k][:, 2:5] - f1[k][:, 2:5]).max() <= xtol, f"positions differ at step {k}"
        assert np.abs(f0[k][:, 5:8] - f1[k][:, 5:8]).max() <= ftol, f"forces differ at step {k}"
        assert abs(t0[k, 1] - t1[k, 1]) <= etol * abs(t0[k, 1]) and abs(t0[k, 2] - t1[k, 2]) <= 1e-6 * max(1.0, abs(t0[k, 2]))
    # NVE through LAMMPS: the conserved quantity of the device-resident run
    assert abs(t1[20, 3] - t1[0, 3]) < 2e-2
    orc = OracleANI(ani1xnr_path)
    lo, hi = _box_of("water30.data")
    check_frame_against_oracle(orc, f1[20], t1[20], lo, hi, prec == "single")


@pytest.mark.gpu
@needs_lmp
def test_device_resident_langevin_thermostat(tmp_path, ani1xnr_path):
    """fix ani/md langevin: the thermostat pulls a cold start towards the target temperature (uniform-noise Langevin as
    fix_langevin.cpp; the random stream is a counter-based hash, so only statistics can be compared with LAMMPS's)."""
    r = run_lmp(tmp_path, ani1xnr_path, ("ani_neighbor", "device"), ("ani_precision", "single"), ("steps", 400),
                ("integrator", "ani/md langevin 300.0 300.0 20.0 4711"), script="in.langevin")
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    rows = [[float(v) for v in l.split()] for l in r.stdout.splitlines() if len(l.split()) == 3 and l.split()[0].isdigit()]
    T = np.array([row[1] for row in rows])
    assert T[0] < 1e-6 and 120.0 < T[-5:].mean() < 600.0, T      # 30 atoms: large fluctuations around 300 K
