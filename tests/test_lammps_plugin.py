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


def run_lmp(tmp_path, model, *variables):
    env = dict(os.environ, LD_LIBRARY_PATH=os.path.dirname(LMP) + ":" + os.environ.get("LD_LIBRARY_PATH", ""))
    cmd = [LMP, "-in", os.path.join(CASE, "in.ani"), "-log", "none", "-var", "plugin", PLUGIN, "-var", "data_file",
           os.path.join(CASE, "water30.data"), "-var", "ani_model_file", model]
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
    single = prec == "single"
    orc = OracleANI(ani1xnr_path)
    for step in (0, 4):                       # step 0: fresh list; step 4: after two rebuilds and four NVE updates
        fr = frames[step]
        x, sp = fr[:, 2:5], fr[:, 1].astype(np.int32) - 1
        xw = x - np.floor((x + 4.0) / 8.0) * 8.0 + 4.0         # wrap into [0, 8)
        sub = domain.decompose(xw, sp, np.array([8.0, 8.0, 8.0]))[0]
        il, nn, jl = domain.full_neighbor_list(sub.x, sub.nlocal)
        ref = orc.compute_full(sub.species, sub.x, il, nn, jl, sub.nlocal, single_coords=single)
        f_ref = domain.fold_ghost_forces(sub, ref["forces"])
        f_ref_by_id = np.empty_like(f_ref)
        f_ref_by_id[sub.global_id[:sub.nlocal]] = f_ref
        pe = th[step, 1]
        if single:
            assert abs(pe - ref["energy"]) <= 1e-6 * abs(ref["energy"])
            assert np.abs(fr[:, 5:8] - f_ref_by_id).max() <= 1e-4 * EV2KCAL
        else:
            assert abs(pe - ref["energy"]) <= 1e-10 * abs(ref["energy"])
            assert np.abs(fr[:, 5:8] - f_ref_by_id).max() <= 1e-8
    # NVE over 4 x 0.1 fs: total energy conserved to the integrator's accuracy
    assert abs(th[4, 3] - th[0, 3]) < 5e-3
