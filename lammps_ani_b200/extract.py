"""One-time host-side parameter extraction -> `.aniw` (see aniw.py).

Replaces `torch::jit::load` + `module_to_dtype` + `select_models` of the reference runner
(/root/reference/src/ani_csrc/ani.cpp:35-97) and the exporter
/root/reference/models/ani_models.py:112-122.  Sources understood:

* a NeuroChem `.info` file (the in-tree ANI-1xnr ensemble,
  /root/reference/external/ani-1xnr/model/ani-1xnr.info) -- `from_neurochem_info`;
* a TorchScript archive exported by the reference (`ani2x.pt`, `ani1x_nr.pt`) -- `from_torchscript`
  walks the ZIP + pickle without importing torchani or libtorch's JIT;
* `synthetic_ani2x`: ANI-2x constants and layer shapes (published values) with
  seeded random weights, because the ANI-2x parameters are not available offline.

CLI:  python -m lammps_ani_b200.extract <source> <out.aniw> [--synthetic-ani2x SEED]
"""
from __future__ import annotations

import bz2
import math
import os
import re
import sys
from typing import Dict, List

import numpy as np

from .aniw import AniParams, save

# ----------------------------------------------------------------------------- NeuroChem


def _parse_list(text: str) -> List[str]:
    return [t.strip() for t in text.strip().lstrip("[").rstrip("]").split(",") if t.strip()]


def read_neurochem_consts(path: str) -> Dict:
    """`*.params` file: Rcr, Rca, EtaR, ShfR, Zeta, ShfZ, EtaA, ShfA, Atyp."""
    out: Dict = {}
    with open(path) as f:
        for line in f:
            if "=" not in line:
                continue
            key, val = [t.strip() for t in line.split("=", 1)]
            if key in ("Rcr", "Rca"):
                out[key] = float(val)
            elif key in ("EtaR", "ShfR", "Zeta", "ShfZ", "EtaA", "ShfA"):
                out[key] = np.array([float(t) for t in _parse_list(val)], dtype=np.float64)
            elif key == "Atyp":
                out["symbols"] = _parse_list(val)
    return out


def read_neurochem_sae(path: str, symbols: List[str]) -> np.ndarray:
    """`sae_linfit.dat`: lines "H,0=-0.578..." (symbol, index = energy)."""
    sae = np.zeros(len(symbols), dtype=np.float64)
    with open(path) as f:
        for line in f:
            line = line.split("#")[0].strip()
            if "=" not in line:
                continue
            lhs, rhs = line.split("=")
            sym, idx = lhs.split(",")
            if sym.strip() in symbols:
                sae[int(idx)] = float(rhs)
    return sae


def read_neurochem_network(nnf_path: str):
    """`ANN-X.nnf` = "<len>=\\n" + bz2(text); layers reference raw f32 `.wparam/.bparam` files
    holding W as (out, in) row-major."""
    raw = open(nnf_path, "rb").read()
    head_end = raw.index(b"=")
    payload = raw[head_end + 1:].lstrip(b"\n")
    text = bz2.decompress(payload).decode("ascii") if payload[:2] == b"BZ" else payload.decode("ascii")
    m = re.search(r"inputsize\s*=\s*(\d+)", text)
    in_size = int(m.group(1))
    layers = []
    d = os.path.dirname(nnf_path)
    for block in re.findall(r"layer\s*\[(.*?)\]\s*(?=layer|\$|$)", text, flags=re.S):
        nodes = int(re.search(r"nodes\s*=\s*(\d+)", block).group(1))
        act = int(re.search(r"activation\s*=\s*(\d+)", block).group(1))
        wf, wn = re.search(r"weights\s*=\s*FILE:([^\[]+)\[(\d+)\]", block).groups()
        bf, bn = re.search(r"biases\s*=\s*FILE:([^\[]+)\[(\d+)\]", block).groups()
        W = np.fromfile(os.path.join(d, wf), dtype="<f4")
        b = np.fromfile(os.path.join(d, bf), dtype="<f4")
        assert W.size == int(wn) and b.size == int(bn) == nodes, (nnf_path, W.size, wn, b.size, bn)
        W = W.reshape(nodes, in_size)
        layers.append((W, b, act))
        in_size = nodes
    for (_, _, act) in layers[:-1]:
        if act != 9:
            raise ValueError(f"{nnf_path}: hidden activation {act} is not CELU(0.1) (NeuroChem 9)")
    if layers[-1][2] != 6:
        raise ValueError(f"{nnf_path}: output activation {layers[-1][2]} is not linear (NeuroChem 6)")
    return [(W, b) for (W, b, _) in layers]


def from_neurochem_info(info_path: str) -> AniParams:
    """4-line info file: consts path, sae path, network-dir prefix, ensemble size."""
    base = os.path.dirname(os.path.abspath(info_path))
    lines = [l.strip() for l in open(info_path) if l.strip()]
    consts_rel, sae_rel, prefix_rel, size = lines[0], lines[1], lines[2], int(lines[3])
    c = read_neurochem_consts(os.path.join(base, consts_rel))
    symbols = c["symbols"]
    sae = read_neurochem_sae(os.path.join(base, sae_rel), symbols)
    weights = []
    dims = None
    for m in range(size):
        per_s = []
        for sym in symbols:
            per_s.append(read_neurochem_network(
                os.path.join(base, f"{prefix_rel}{m}", "networks", f"ANN-{sym}.nnf")))
        weights.append(per_s)
        d = [[net[0][0].shape[1]] + [W.shape[0] for (W, _) in net] for net in per_s]
        if dims is None:
            dims = d
        elif d != dims:
            raise ValueError("ensemble members have different architectures")
    p = AniParams(symbols, c["Rcr"], c["Rca"], float(c["EtaR"][0]), float(c["EtaA"][0]),
                  float(c["Zeta"][0]), c["ShfR"], c["ShfA"], c["ShfZ"], sae, dims, weights)
    p.validate()
    return p


# ----------------------------------------------------------------------------- synthetic ANI-2x

ANI2X_SYMBOLS = ["H", "C", "N", "O", "S", "F", "Cl"]
ANI2X_HIDDEN = {"H": [256, 192, 160], "C": [224, 192, 160], "N": [192, 160, 128], "O": [192, 160, 128],
                "S": [160, 128, 96], "F": [160, 128, 96], "Cl": [160, 128, 96]}
# published ANI-2x self energies (Hartree); only shift the total energy
ANI2X_SAE = [-0.5978583943827134, -38.08933878049795, -54.711968298621066, -75.19106774742086,
             -398.1577125334925, -99.80348506781634, -460.1681939421027]


def ani2x_constants() -> Dict:
    return dict(Rcr=5.1, Rca=3.5, EtaR=19.7, EtaA=12.5, Zeta=14.1,
                ShfR=0.8 + 0.26875 * np.arange(16), ShfA=0.8 + 0.3375 * np.arange(8),
                ShfZ=(np.arange(4) * 2 + 1) * math.pi / 8)


# GFN2-xTB repulsion parameters (Bannwarth, Ehlert, Grimme, JCTC 15 (2019) 1652, SI): exponent alpha and effective nuclear
# charge Y_eff per element; k_rep = 1 when both atoms are H / He, 3/2 otherwise.  Typed in from the publication (torchani's
# table is not in the tree): use them for tests and synthetic models; real `*_repulsion.pt` files carry their own tables.
XTB_ALPHA = {"H": 2.213717, "C": 1.247655, "N": 1.682689, "O": 2.165712, "F": 2.421394, "S": 1.214553, "Cl": 1.577144}
XTB_YEFF = {"H": 1.105388, "C": 4.231078, "N": 5.242592, "O": 5.784415, "F": 7.021486, "S": 14.995090, "Cl": 17.353134}


def xtb_repulsion_tables(symbols, cutoff: float = 5.2, cutoff_fn: int = 1) -> dict:
    """RepulsionXTB pair tables for `symbols` (torchani potentials.RepulsionXTB: y_ab = Y_a Y_b, sqrt_alpha_ab = sqrt(a_a a_b),
    k_rep_ab = 1 for H-H else 1.5; smooth cutoff function at 5.2 A)."""
    a = np.array([XTB_ALPHA[s] for s in symbols])
    y = np.array([XTB_YEFF[s] for s in symbols])
    light = np.array([s in ("H", "He") for s in symbols])
    return {"cutoff": float(cutoff), "cutoff_fn": int(cutoff_fn), "y_ab": np.outer(y, y), "sqrt_alpha_ab": np.sqrt(np.outer(a, a)),
            "k_rep_ab": np.where(np.outer(light, light), 1.0, 1.5)}


def with_repulsion(p: AniParams, **kw) -> AniParams:
    """The same model with the RepulsionXTB term switched on (what `ani2x_repulsion.pt` / `ani1x_nr_repulsion.pt` are to
    `ani2x.pt` / `ani1x_nr.pt`, examples/combustion/README.md:4,10)."""
    import copy
    q = copy.copy(p)
    q.use_repulsion = True
    q.repulsion = xtb_repulsion_tables(p.symbols, **kw)
    q.validate()
    return q


def synthetic_ani2x(seed: int = 2024, num_models: int = 8, weight_scale: float = 1.0) -> AniParams:
    """ANI-2x geometry (1008-long AEV, 7 species, 8 members) with seeded random weights.

    Weights ~ N(0, 1/fan_in) and small biases: throughput does not depend on the values, and the
    CPU oracle evaluates the very same numbers, so parity tests are meaningful.  The output layer
    is scaled so atomic energies are O(0.1 Ha) like the trained networks.  weight_scale multiplies the hidden-layer
    weights (tests of the fp16 operand-range guard)."""
    c = ani2x_constants()
    rng = np.random.default_rng(seed)
    L = 7 * 16 + 28 * 32
    dims = [[L] + ANI2X_HIDDEN[s] + [1] for s in ANI2X_SYMBOLS]
    weights = []
    for _m in range(num_models):
        per_s = []
        for d in dims:
            per_l = []
            for l in range(len(d) - 1):
                # AEV inputs are non-negative and sparse-ish: boost layer-1 scale so activations are O(1)
                scale = (2.0 if l == 0 else 1.4) / math.sqrt(d[l])
                if l == len(d) - 2:
                    scale = 0.3 / math.sqrt(d[l])
                if l < len(d) - 2:
                    scale *= weight_scale
                W = (rng.standard_normal((d[l + 1], d[l])) * scale).astype(np.float32)
                b = (rng.standard_normal(d[l + 1]) * 0.1).astype(np.float32)
                per_l.append((W, b))
            per_s.append(per_l)
        weights.append(per_s)
    p = AniParams(list(ANI2X_SYMBOLS), c["Rcr"], c["Rca"], c["EtaR"], c["EtaA"], c["Zeta"],
                  c["ShfR"], c["ShfA"], c["ShfZ"], np.array(ANI2X_SAE), dims, weights)
    p.validate()
    return p


# ----------------------------------------------------------------------------- TorchScript .pt


def from_torchscript(pt_path: str) -> AniParams:
    from .ptreader import read_torchscript_params  # kept separate: zip + restricted unpickler
    return read_torchscript_params(pt_path)


def extract(source: str, out_path: str) -> AniParams:
    if source.endswith(".info"):
        p = from_neurochem_info(source)
    elif source.endswith(".pt"):
        p = from_torchscript(source)
    else:
        raise ValueError(f"do not know how to read {source!r} (expected .info or .pt)")
    save(p, out_path)
    return p


def main(argv=None) -> int:
    argv = list(sys.argv[1:] if argv is None else argv)
    if len(argv) >= 2 and argv[0] == "--synthetic-ani2x":
        seed = int(argv[2]) if len(argv) > 2 else 2024
        save(synthetic_ani2x(seed), argv[1])
        print(f"wrote synthetic ANI-2x (seed {seed}) -> {argv[1]}")
        return 0
    if len(argv) != 2:
        print(__doc__)
        return 2
    p = extract(argv[0], argv[1])
    print(f"wrote {argv[1]}: species {p.symbols}, aev {p.aev_length}, members {p.num_models}")
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
