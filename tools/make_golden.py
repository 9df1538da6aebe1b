#!/usr/bin/env python
"""Generate the committed fixtures under tests/golden/ (run in the authoring container only: it reads
/root/reference, which does not exist on the GPU box).

1. tests/golden/water30_reference.json -- the reference's own known-answer inputs and expected outputs,
   parsed from /root/reference/src/ani_csrc/test_model.cpp:84-137 (30-atom water, non-periodic half pair
   list, ANI-2x energy and forces) and from the YAML files under /root/reference/tests/lammps-unittest/
   (init_vdwl / init_forces / init_stress per pair_style line).  These absolute numbers need the ANI-2x
   weights, which are not available offline; they are carried so that the parity tests pin themselves
   the moment an `ani2x.aniw` is present (tests/test_reference_goldens.py).
2. tests/golden/water30_ani1xnr.json -- the same 30-atom inputs evaluated with the in-tree ANI-1xnr
   ensemble by the CPU oracle (periodic and non-periodic, full list), so the GPU box can check the CUDA
   path against real trained weights without /root/reference.
3. lammps_ani_b200/data/ani1x_nr.aniw -- extracted ANI-1xnr parameters (git-ignored, ships with gpurun).
4. tests/lammps/alanine-dipeptide.data -- BASELINE.json config 3's input, the reference's
   examples/alanine-dipeptide/alanine-dipeptide.data (1 912 atoms) as a test fixture: atom lines verbatim, the
   all-zero "xy xz yz" tilt line dropped so LAMMPS reads an orthogonal box (the reference's in.lammps keeps it
   triclinic with zero tilt, which is the same geometry).
"""
import json
import os
import re
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
REF = "/root/reference"


def parse_cpp_array(text, name):
    m = re.search(name + r"\s*=\s*\{(.*?)\};", text, flags=re.S)
    return [float(t) for t in re.findall(r"-?\d+\.?\d*(?:[eE][-+]?\d+)?", m.group(1))]


def parse_yaml_case(path):
    txt = open(path).read()
    out = {"file": os.path.relpath(path, REF)}
    out["pair_style"] = re.search(r"^pair_style:\s*(.*)$", txt, flags=re.M).group(1).strip()
    out["epsilon"] = float(re.search(r"^epsilon:\s*(\S+)", txt, flags=re.M).group(1))
    out["periodic"] = "boundary p p p" in txt
    for key in ("init_vdwl", "run_vdwl"):
        out[key] = float(re.search(rf"^{key}:\s*(\S+)", txt, flags=re.M).group(1))
    for key in ("init_forces", "run_forces"):
        block = re.search(rf"^{key}:\s*!\s*\|2?\n((?:\s+\d+\s+.*\n)+)", txt, flags=re.M).group(1)
        rows = [[float(t) for t in line.split()[1:4]] for line in block.strip().splitlines()]
        out[key] = rows
    for key in ("init_stress", "run_stress"):
        m = re.search(rf"^{key}:\s*!\s*\|-?\n\s+(.*)$", txt, flags=re.M)
        out[key] = [float(t) for t in m.group(1).split()] if m else None
    return out


def main():
    from lammps_ani_b200 import extract, aniw, domain
    from oracle.oracle import OracleANI

    gold = os.path.join(ROOT, "tests", "golden")
    os.makedirs(gold, exist_ok=True)
    cpp = open(os.path.join(REF, "src/ani_csrc/test_model.cpp")).read()
    body = cpp[cpp.index("int test_ani2x_withnbr"):]
    coords = parse_cpp_array(body, r"std::vector<double> coords")
    species = [int(v) for v in parse_cpp_array(body, r"std::vector<int64_t> species")]
    idx12 = [int(v) for v in parse_cpp_array(body, r"std::vector<int64_t> atom_index12")]
    forces = parse_cpp_array(body, r"std::vector<double> expected_force")
    e_ha = float(re.search(r"expected_energy\s*=\s*(-?\d+\.\d+)", body).group(1))
    ref = {
        "source": "src/ani_csrc/test_model.cpp:84-137 and tests/lammps-unittest/*/*.yaml (ANI-2x, 8 members)",
        "coords": coords, "species": species, "atom_index12": idx12,
        "test_model": {"expected_energy_hartree": e_ha, "expected_force_kcalmol_A": forces,
                       "threshold_fp64": 1e-8, "threshold_fp32": 3e-4},
        "box": [-4.0, 4.0], "yaml": [],
    }
    ydir = os.path.join(REF, "tests/lammps-unittest")
    for d in sorted(os.listdir(ydir)):
        p = os.path.join(ydir, d)
        if os.path.isdir(p):
            for fn in sorted(os.listdir(p)):
                if fn.endswith(".yaml"):
                    ref["yaml"].append(parse_yaml_case(os.path.join(p, fn)))
    with open(os.path.join(gold, "water30_reference.json"), "w") as f:
        json.dump(ref, f)
    print("reference goldens:", len(ref["yaml"]), "yaml cases")

    data_dir = os.path.join(ROOT, "lammps_ani_b200", "data")
    os.makedirs(data_dir, exist_ok=True)
    aniw_path = os.path.join(data_dir, "ani1x_nr.aniw")
    p = extract.from_neurochem_info(os.path.join(REF, "external/ani-1xnr/model/ani-1xnr.info"))
    aniw.save(p, aniw_path)
    print("wrote", aniw_path)

    x = np.array(coords).reshape(-1, 3)
    sp = np.array(species, dtype=np.int32)
    orc = OracleANI(aniw_path)
    out = {"weights": "ANI-1xnr (external/ani-1xnr/model/ani-1xnr.info), 8 members", "cases": []}
    # (a) the reference's own half list, non-periodic
    r = orc.compute_half(sp, x, np.array(idx12).reshape(2, -1), 30)
    out["cases"].append({"name": "nonperiodic_half_reference_list", "energy": r["energy"],
                         "forces": r["forces"].tolist(), "virial": r["virial"].tolist()})
    # (b) periodic 8 A box with 7.1 A halo (many self images), full list
    box = np.array([8.0, 8.0, 8.0])
    sub = domain.decompose(x + 4.0, sp, box, (1, 1, 1))[0]
    il, nn, jl = domain.full_neighbor_list(sub.x, sub.nlocal)
    r = orc.compute_full(sub.species, sub.x, il, nn, jl, sub.nlocal)
    out["cases"].append({"name": "periodic_full", "ntotal": int(sub.ntotal), "npairs": int(jl.size),
                         "energy": r["energy"], "forces_folded": domain.fold_ghost_forces(sub, r["forces"]).tolist(),
                         "virial": r["virial"].tolist(), "eatom": r["eatom"].tolist()})
    with open(os.path.join(gold, "water30_ani1xnr.json"), "w") as f:
        json.dump(out, f)
    print("oracle goldens written; E(nonperiodic) =", out["cases"][0]["energy"], " E(periodic) =", out["cases"][1]["energy"])


def test_model_header():
    """tools/test_model_golden.h: the vectors of the C++ known-answer driver tools/test_model.cpp (the reference's
    src/ani_csrc/test_model.cpp:84-137 inputs + ANI-2x expectations, and the oracle's ANI-1xnr results on the same inputs)."""
    gold = os.path.join(ROOT, "tests", "golden")
    ref = json.load(open(os.path.join(gold, "water30_reference.json")))
    # the driver constructs ANI with use_cuaev = false like test_model.cpp:83: the half list is consumed WITHOUT a radial
    # re-filter (SURVEY.md note N3), so the ANI-1xnr expectation is the oracle in that mode
    from oracle.oracle import OracleANI
    orc = OracleANI(os.path.join(ROOT, "lammps_ani_b200", "data", "ani1x_nr.aniw"))
    r = orc.compute_half(np.array(ref["species"], dtype=np.int32), np.array(ref["coords"]).reshape(-1, 3),
                         np.array(ref["atom_index12"]).reshape(2, -1), 30, radial_filter=False)
    case = {"forces": r["forces"].tolist(), "energy": r["energy"]}

    def arr(name, typ, vals, fmt):
        body = ", ".join(fmt % v for v in vals)
        return f"static const {typ} {name}[{len(vals)}] = {{{body}}};\n"
    out = ["// GENERATED by tools/make_golden.py -- do not edit.  Inputs and ANI-2x expectations: the reference's known-answer test",
           "// (src/ani_csrc/test_model.cpp:84-137); ANI-1xnr expectations: oracle/ani_oracle.c on the same inputs (fp64).",
           "#pragma once", "#include <cstdint>", ""]
    txt = "\n".join(out) + "\n"
    txt += arr("kCoords", "double", ref["coords"], "%.17g")
    txt += arr("kSpecies", "int", ref["species"], "%d")
    txt += arr("kAtomIndex12", "int64_t", ref["atom_index12"], "%d")
    txt += arr("kAni2xForce", "double", ref["test_model"]["expected_force_kcalmol_A"], "%.17g")
    txt += "static const double kAni2xEnergyHartree = %.17g;\n" % ref["test_model"]["expected_energy_hartree"]
    txt += arr("kAni1xnrForce", "double", np.array(case["forces"]).reshape(-1).tolist(), "%.17g")
    txt += "static const double kAni1xnrEnergyKcalmol = %.17g;\n" % case["energy"]
    with open(os.path.join(ROOT, "tools", "test_model_golden.h"), "w") as f:
        f.write(txt)
    print("wrote tools/test_model_golden.h")


def alanine_fixture():
    src = os.path.join(REF, "examples/alanine-dipeptide/alanine-dipeptide.data")
    dst = os.path.join(ROOT, "tests", "lammps", "alanine-dipeptide.data")
    lines = open(src).read().splitlines()
    out = ["# solvated alanine dipeptide, 1912 atoms: test fixture generated by tools/make_golden.py from the reference's "
           "examples/alanine-dipeptide/alanine-dipeptide.data"]
    for ln in lines[1:]:
        f = ln.split()
        if len(f) >= 6 and f[3:6] == ["xy", "xz", "yz"]:
            assert all(float(v) == 0.0 for v in f[:3])
            continue
        if "bond" in ln:
            continue
        out.append(ln.split("#")[0].rstrip() if len(f) >= 5 and f[0].isdigit() else ln)
    with open(dst, "w") as fh:
        fh.write("\n".join(out) + "\n")
    print("wrote", dst)


if __name__ == "__main__":
    main()
    alanine_fixture()
    test_model_header()
