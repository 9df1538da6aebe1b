"""Read the parameters of a reference-exported TorchScript model (`ani2x.pt`, `ani1x_nr.pt`; written by
/root/reference/models/ani_models.py:112-122 `torch.jit.save`) WITHOUT libtorch, torchani or running any code from
the archive: the file is a ZIP_STORED archive with `<root>/data.pkl` (the module object graph, pickle protocol 2
with persistent ids for storages) and `<root>/data/<key>` raw little-endian storages.  A restricted unpickler rebuilds
only tensors (as numpy arrays) and generic attribute bags for the `__torch__.*` module classes.

The object graph is then searched by attribute NAME for what the hot path needs (lammps_ani.py:104-110, :219-257):
AEV constants (EtaR, ShfR, EtaA, Zeta, ShfA, ShfZ, Rcr, Rca), the per-member / per-species Linear layers (either an
`Ensemble` of `ANIModel`s with `<member>.<symbol>.<k>.weight [out,in]`, or the stacked `BmmEnsemble` form with
`[M, in, out]` weights), and the self energies.  torchani itself is not in /root/reference (missing submodule), so the
attribute names follow the public torchani layout and the reader reports exactly what it could not find.
"""
from __future__ import annotations

import io
import pickle
import zipfile
from typing import Any, Dict, List, Tuple

import numpy as np

_DTYPES = {
    "FloatStorage": np.float32, "DoubleStorage": np.float64, "HalfStorage": np.float16, "LongStorage": np.int64,
    "IntStorage": np.int32, "ShortStorage": np.int16, "CharStorage": np.int8, "ByteStorage": np.uint8, "BoolStorage": np.bool_,
}


class _Bag:
    """Stand-in for a `__torch__.*` scripted class instance: keeps the pickled attribute dict."""

    def __new__(cls, *a, **k):                      # pickle's NEWOBJ bypasses __init__
        o = object.__new__(cls)
        o.__dict__["_state"] = {}
        return o

    def __init__(self, *a, **k):
        pass

    def __setstate__(self, state):
        if isinstance(state, dict):
            self._state.update(state)
        elif isinstance(state, tuple) and len(state) == 2 and isinstance(state[1], dict):
            self._state.update(state[1] or {})
            if isinstance(state[0], dict):
                self._state.update(state[0])

    def __setattr__(self, k, v):
        self._state[k] = v


class _StorageType:
    def __init__(self, name):
        self.dtype = _DTYPES[name]


def _rebuild_tensor_v2(storage, offset, size, stride, *unused):
    arr, = storage
    size, stride = tuple(size), tuple(stride)
    if len(size) == 0:
        return np.array(arr[offset])
    item = arr.dtype.itemsize
    v = np.lib.stride_tricks.as_strided(arr[offset:], shape=size, strides=tuple(s * item for s in stride))
    return np.array(v)        # own, contiguous copy


def _rebuild_parameter(data, requires_grad=False, backward_hooks=None, *a):
    return data


class _Unpickler(pickle.Unpickler):
    def __init__(self, f, zf: zipfile.ZipFile, root: str):
        super().__init__(f)
        self.zf, self.root = zf, root

    def find_class(self, module, name):
        if module.startswith("__torch__"):
            return type(name, (_Bag,), {"__module__": module})
        if module == "torch._utils" and name in ("_rebuild_tensor_v2", "_rebuild_tensor"):
            return _rebuild_tensor_v2
        if module == "torch._utils" and name in ("_rebuild_parameter", "_rebuild_parameter_with_state"):
            return _rebuild_parameter
        if module == "torch" and name in _DTYPES:
            return _StorageType(name)
        if module == "collections" and name == "OrderedDict":
            import collections
            return collections.OrderedDict
        if module == "torch" and name in ("device", "Size"):
            return (lambda *a: a)
        if module == "torch.jit._pickle" or module == "torch._jit_internal":
            return (lambda *a: a[0] if a else None)        # build_intlist / build_tensorlist ... : keep the payload
        raise pickle.UnpicklingError(f"refusing to load {module}.{name} from a model archive")

    def persistent_load(self, pid):
        if not (isinstance(pid, tuple) and pid and pid[0] == "storage"):
            raise pickle.UnpicklingError(f"unexpected persistent id {pid!r}")
        _, stype, key, _location, numel = pid[:5]
        raw = self.zf.read(f"{self.root}/data/{key}")
        return (np.frombuffer(raw, dtype=stype.dtype, count=int(numel)),)


def load_object_graph(pt_path: str):
    with zipfile.ZipFile(pt_path) as zf:
        names = zf.namelist()
        pkl = [n for n in names if n.endswith("/data.pkl") and n.count("/") == 1]
        if not pkl:
            raise ValueError(f"{pt_path}: not a TorchScript archive (no <root>/data.pkl)")
        root = pkl[0].split("/")[0]
        return _Unpickler(io.BytesIO(zf.read(pkl[0])), zf, root).load()


def flatten(obj: Any, prefix: str = "", out: Dict[str, Any] = None, depth: int = 0) -> Dict[str, Any]:
    """{dotted attribute path: numpy array | python scalar | str | list of scalars}"""
    out = {} if out is None else out
    if depth > 24:
        return out
    if isinstance(obj, _Bag):
        for k, v in obj._state.items():
            flatten(v, f"{prefix}.{k}" if prefix else str(k), out, depth + 1)
    elif isinstance(obj, dict):
        for k, v in obj.items():
            flatten(v, f"{prefix}.{k}" if prefix else str(k), out, depth + 1)
    elif isinstance(obj, (list, tuple)) and obj and all(isinstance(x, (_Bag, dict, np.ndarray)) for x in obj):
        for k, v in enumerate(obj):
            flatten(v, f"{prefix}.{k}", out, depth + 1)
    elif isinstance(obj, (np.ndarray, int, float, bool, str, list, tuple)):
        out[prefix] = obj
    return out


# ------------------------------------------------------------------------------------------ parameter assembly
PERIODIC = ["H", "He", "Li", "Be", "B", "C", "N", "O", "F", "Ne", "Na", "Mg", "Al", "Si", "P", "S", "Cl"]


def _find(flat: Dict[str, Any], suffix: str):
    hits = [(k, v) for k, v in flat.items() if k == suffix or k.endswith("." + suffix)]
    hits.sort(key=lambda kv: len(kv[0]))
    return hits[0] if hits else (None, None)


def _scalar(v) -> float:
    return float(np.asarray(v).reshape(-1)[0])


def _layers_of(flat: Dict[str, Any], prefix: str) -> List[Tuple[np.ndarray, np.ndarray]]:
    """Linear layers under `prefix` in attribute order: [(weight, bias), ...]."""
    ws = sorted((k for k in flat if k.startswith(prefix) and k.endswith(".weight") and isinstance(flat[k], np.ndarray)),
                key=lambda k: [int(t) if t.isdigit() else t for t in k.split(".")])
    out = []
    for k in ws:
        b = flat.get(k[:-len("weight")] + "bias")
        out.append((np.asarray(flat[k]), None if b is None else np.asarray(b)))
    return out


def read_torchscript_params(pt_path: str):
    """-> aniw.AniParams.  Raises ValueError naming what is missing when the archive does not look like a LammpsANI export."""
    from .aniw import AniParams
    flat = flatten(load_object_graph(pt_path))
    missing = []
    consts = {}
    for name in ("EtaR", "ShfR", "EtaA", "Zeta", "ShfA", "ShfZ"):
        k, v = _find(flat, name)
        if v is None:
            missing.append(name)
        else:
            consts[name] = np.asarray(v, dtype=np.float64).reshape(-1)
    for name in ("Rcr", "Rca"):
        k, v = _find(flat, name)
        if v is None:
            missing.append(name)
        else:
            consts[name] = _scalar(v)
    _, sae = _find(flat, "self_energies")
    if sae is None:
        missing.append("energy_shifter.self_energies")
    if missing:
        raise ValueError(f"{pt_path}: cannot find {missing} in the TorchScript object graph "
                         f"({len(flat)} attributes; first few: {list(flat)[:8]})")
    sae = np.asarray(sae, dtype=np.float64).reshape(-1)
    S = sae.size
    _, syms = _find(flat, "species")
    if isinstance(syms, (list, tuple)) and len(syms) == S and all(isinstance(x, str) for x in syms):
        symbols = list(syms)
    else:
        _, znum = _find(flat, "atomic_numbers")
        z = np.asarray(znum).reshape(-1) if znum is not None else None
        symbols = [PERIODIC[int(q) - 1] for q in z] if z is not None and z.size == S else \
            (["H", "C", "N", "O", "S", "F", "Cl"][:S] if S in (4, 7) else [f"X{q}" for q in range(S)])
    nR, nA, nZ = consts["ShfR"].size, consts["ShfA"].size, consts["ShfZ"].size
    aev_len = S * nR + (S * (S + 1) // 2) * nA * nZ
    # ---- networks.  Form (b): stacked BmmEnsemble weights [M, in, out] / [M, out, in] per species
    nn_keys = [k for k in flat if k.endswith(".weight") and isinstance(flat[k], np.ndarray) and "neural_networks" in k]
    if not nn_keys:
        nn_keys = [k for k in flat if k.endswith(".weight") and isinstance(flat[k], np.ndarray) and not k.startswith("rep_calc.")]
    if not nn_keys:
        raise ValueError(f"{pt_path}: no Linear weights found")
    stacked = all(flat[k].ndim == 3 for k in nn_keys)
    weights: List[List[List[tuple]]] = []
    if stacked:
        # group by the path component that enumerates species (the first differing integer / symbol component)
        groups: Dict[str, List[str]] = {}
        for k in nn_keys:
            parts = k.split(".")
            groups.setdefault(".".join(parts[:-2]), []).append(k)
        # species order = the model's: an integer path component indexes `symbols`, a symbol component is looked up in it
        # (a ModuleDict keyed by element symbol must NOT be sorted alphabetically); otherwise keep the archive's own order
        def species_of(gk):
            for t in reversed(gk.split(".")):
                if t in symbols:
                    return symbols.index(t)
            return None
        by_symbol = {gk: species_of(gk) for gk in groups}
        if all(v is not None for v in by_symbol.values()) and len(set(by_symbol.values())) == len(groups):
            gkeys = sorted(groups, key=lambda gk: by_symbol[gk])
        else:
            def int_key(gk):
                ints = [int(t) for t in gk.split(".") if t.isdigit()]
                if not ints:
                    raise ValueError(f"{pt_path}: cannot tell which species the network group '{gk}' belongs to")
                return ints
            gkeys = sorted(groups, key=int_key)
        if len(gkeys) != S:
            raise ValueError(f"{pt_path}: expected {S} per-species network groups, found {len(gkeys)}: {gkeys}")
        M = flat[groups[gkeys[0]][0]].shape[0]
        weights = [[[] for _ in range(S)] for _ in range(M)]
        for s, g in enumerate(gkeys):
            prev = aev_len
            for W, b in _layers_of(flat, g + "."):
                Wm = W if W.shape[2] == prev and W.shape[1] != prev else np.transpose(W, (0, 2, 1)) if W.shape[1] == prev else W
                if Wm.shape[2] != prev:
                    raise ValueError(f"{pt_path}: layer shapes under {g} do not chain ({W.shape}, expected in={prev})")
                bb = np.zeros((M, Wm.shape[1]), np.float32) if b is None else np.asarray(b).reshape(M, -1)
                for m in range(M):
                    weights[m][s].append((np.ascontiguousarray(Wm[m], dtype=np.float32), np.ascontiguousarray(bb[m], dtype=np.float32)))
                prev = Wm.shape[1]
    else:
        # Form (a): <...>.<member>.<symbol or index>.<layer>.weight  [out, in]
        def key_ms(k):
            parts = k.split(".")
            i = parts.index("neural_networks") if "neural_networks" in parts else 0
            tail = parts[i + 1:-2]          # between the container and "<layer>.weight"
            return tail
        tails = {tuple(key_ms(k)) for k in nn_keys}
        members = sorted({t[0] for t in tails if len(t) >= 2}, key=lambda t: int(t) if t.isdigit() else t)
        if not members or any(len(t) < 2 for t in tails):
            members, single = ["_"], True       # a single ANIModel: <symbol>.<layer>.weight
        else:
            single = False
        for m in members:
            per_s = []
            for s, sym in enumerate(symbols):
                cands = [k for k in nn_keys if (single or key_ms(k)[0] == m) and key_ms(k)[-1] in (sym, str(s))]
                if not cands:
                    raise ValueError(f"{pt_path}: no network for species {sym} (member {m})")
                pref = cands[0][: cands[0].rfind(".", 0, cands[0].rfind(".weight"))] + "."
                layers = []
                for W, b in _layers_of(flat, pref):
                    bb = np.zeros(W.shape[0], np.float32) if b is None else np.asarray(b).reshape(-1)
                    layers.append((np.ascontiguousarray(W, dtype=np.float32), np.ascontiguousarray(bb, dtype=np.float32)))
                per_s.append(layers)
            weights.append(per_s)
    dims = [[aev_len] + [W.shape[0] for (W, _) in weights[0][s]] for s in range(S)]
    # The reference always stores a `use_repulsion` bool AND a dummy `rep_calc = RepulsionXTB(("H"))` submodule, also when the
    # term is off (models/lammps_ani.py:91-99): only the VALUE of the top-level attribute says whether it is used.
    rep = bool(flat.get("use_repulsion", False))
    repulsion = None
    if rep:
        # RepulsionXTB buffers (torchani potentials.RepulsionXTB; attribute names as in torchani 2.x -- not verifiable offline)
        tabs = {}
        for name in ("y_ab", "sqrt_alpha_ab", "k_rep_ab"):
            k, v = _find({kk: vv for kk, vv in flat.items() if kk.startswith("rep_calc.")}, name)
            if v is None or np.asarray(v).shape != (S, S):
                raise ValueError(f"{pt_path}: use_repulsion is set but rep_calc.{name} [{S},{S}] was not found")
            tabs[name] = np.asarray(v, dtype=np.float64)
        _, cut = _find({kk: vv for kk, vv in flat.items() if kk.startswith("rep_calc.")}, "cutoff")
        fn_names = " ".join(str(v) for kk, v in flat.items() if kk.startswith("rep_calc.") and isinstance(v, str)).lower()
        repulsion = dict(cutoff=_scalar(cut) if cut is not None else 5.2, cutoff_fn=0 if "cosine" in fn_names else 1, **tabs)
    p = AniParams(symbols, consts["Rcr"], consts["Rca"], float(consts["EtaR"][0]), float(consts["EtaA"][0]),
                  float(consts["Zeta"][0]), consts["ShfR"], consts["ShfA"], consts["ShfZ"], sae, dims, weights,
                  use_repulsion=bool(rep), repulsion=repulsion)
    p.validate()
    return p
