"""The `.aniw` flat parameter blob: what `ani_model_load` (include/ani_b200.h) reads.

The reference loads a TorchScript archive with `torch::jit::load` every time a pair style is
created (/root/reference/src/ani_csrc/ani.cpp:46).  Here the parameters are extracted ONCE on
the host (lammps_ani_b200/extract.py) into this little-endian blob, which both the CUDA library
(lammps_ani_b200/csrc/model_io.cpp) and the CPU oracle (oracle/ani_oracle.c) parse on their own.

Layout (all little endian, no padding):

    char[8]   magic  "ANIW0001"
    int32[8]  num_species, nShfR, nShfA, nShfZ, num_models, max_layers, use_repulsion, reserved
    f64[5]    Rcr, Rca, EtaR, EtaA, Zeta
    f64[nShfR] ShfR ; f64[nShfA] ShfA ; f64[nShfZ] ShfZ
    char[4]*num_species   element symbols, NUL padded ("H\\0\\0\\0", "Cl\\0\\0")
    f64[num_species]      self atomic energies (Hartree)
    per species s:  int32 nlayers ; int32[nlayers+1] dims   (dims[0] == aev_length)
    per model m, per species s, per layer l:
        f32[dims[l+1]*dims[l]]  W  (row-major [out][in], i.e. torch.nn.Linear.weight)
        f32[dims[l+1]]          b
    only when use_repulsion != 0 (the RepulsionXTB pair term, models/lammps_ani.py:186-192,301-330):
        f64 cutoff (Angstrom) ; int32 cutoff_fn (0 cosine, 1 smooth order 2) ; int32 reserved
        f64[S*S] y_ab ; f64[S*S] sqrt_alpha_ab ; f64[S*S] k_rep_ab      (row-major species x species, atomic units)

Hidden layers use CELU(alpha=0.1), the last layer is linear (NeuroChem activation 9 / 6,
/root/reference/external/ani-1xnr/model/ani-1xnr/train0/inputtrain.ipt).
"""
from __future__ import annotations

import struct
from dataclasses import dataclass, field
from typing import List

import numpy as np

MAGIC = b"ANIW0001"
CELU_ALPHA = 0.1
HARTREE2KCALMOL = 627.5094738898777  # /root/reference/src/ani_csrc/ani.h:9


@dataclass
class AniParams:
    symbols: List[str]
    Rcr: float
    Rca: float
    EtaR: float
    EtaA: float
    Zeta: float
    ShfR: np.ndarray
    ShfA: np.ndarray
    ShfZ: np.ndarray
    sae: np.ndarray                      # [S] f64, Hartree
    dims: List[List[int]]                # per species [in, h1, ..., 1]
    # weights[m][s][l] = (W [out,in] f32, b [out] f32)
    weights: List[List[List[tuple]]] = field(default_factory=list)
    use_repulsion: bool = False
    # RepulsionXTB tables when use_repulsion: dict(cutoff=float, cutoff_fn=0|1, y_ab=[S,S], sqrt_alpha_ab=[S,S], k_rep_ab=[S,S])
    repulsion: dict = None

    @property
    def num_species(self) -> int:
        return len(self.symbols)

    @property
    def num_models(self) -> int:
        return len(self.weights)

    @property
    def radial_length(self) -> int:
        return self.num_species * len(self.ShfR)

    @property
    def angular_sublength(self) -> int:
        return len(self.ShfA) * len(self.ShfZ)

    @property
    def aev_length(self) -> int:
        S = self.num_species
        return self.radial_length + (S * (S + 1) // 2) * self.angular_sublength

    def macs_per_atom(self, s: int) -> int:
        d = self.dims[s]
        return sum(d[i] * d[i + 1] for i in range(len(d) - 1))

    def validate(self) -> None:
        L = self.aev_length
        for s, d in enumerate(self.dims):
            if d[0] != L:
                raise ValueError(f"species {s}: first dim {d[0]} != aev length {L}")
            if d[-1] != 1:
                raise ValueError(f"species {s}: last layer must have one output")
        for m in range(self.num_models):
            for s, d in enumerate(self.dims):
                for l in range(len(d) - 1):
                    W, b = self.weights[m][s][l]
                    if W.shape != (d[l + 1], d[l]) or b.shape != (d[l + 1],):
                        raise ValueError(f"bad shape m={m} s={s} l={l}: {W.shape} {b.shape}")
        if self.use_repulsion:
            S = self.num_species
            r = self.repulsion
            if not r or any(np.asarray(r[k]).shape != (S, S) for k in ("y_ab", "sqrt_alpha_ab", "k_rep_ab")):
                raise ValueError("use_repulsion needs the RepulsionXTB tables y_ab, sqrt_alpha_ab, k_rep_ab of shape [S, S]")


def save(params: AniParams, path: str) -> None:
    params.validate()
    S = params.num_species
    max_layers = max(len(d) - 1 for d in params.dims)
    with open(path, "wb") as f:
        f.write(MAGIC)
        f.write(struct.pack("<8i", S, len(params.ShfR), len(params.ShfA), len(params.ShfZ),
                            params.num_models, max_layers, int(params.use_repulsion), 0))
        f.write(struct.pack("<5d", params.Rcr, params.Rca, params.EtaR, params.EtaA, params.Zeta))
        for arr in (params.ShfR, params.ShfA, params.ShfZ):
            f.write(np.asarray(arr, dtype="<f8").tobytes())
        for sym in params.symbols:
            f.write(sym.encode("ascii").ljust(4, b"\0")[:4])
        f.write(np.asarray(params.sae, dtype="<f8").tobytes())
        for d in params.dims:
            f.write(struct.pack("<i", len(d) - 1))
            f.write(struct.pack(f"<{len(d)}i", *d))
        for m in range(params.num_models):
            for s, d in enumerate(params.dims):
                for l in range(len(d) - 1):
                    W, b = params.weights[m][s][l]
                    f.write(np.ascontiguousarray(W, dtype="<f4").tobytes())
                    f.write(np.ascontiguousarray(b, dtype="<f4").tobytes())
        if params.use_repulsion:
            r = params.repulsion
            f.write(struct.pack("<dii", float(r["cutoff"]), int(r.get("cutoff_fn", 1)), 0))
            for k in ("y_ab", "sqrt_alpha_ab", "k_rep_ab"):
                f.write(np.ascontiguousarray(r[k], dtype="<f8").tobytes())


def load(path: str) -> AniParams:
    with open(path, "rb") as f:
        buf = f.read()
    if buf[:8] != MAGIC:
        raise ValueError(f"{path}: not an ANIW0001 file")
    off = 8
    S, nR, nA, nZ, M, _maxl, rep, _ = struct.unpack_from("<8i", buf, off); off += 32
    Rcr, Rca, EtaR, EtaA, Zeta = struct.unpack_from("<5d", buf, off); off += 40

    def f64(n):
        nonlocal off
        a = np.frombuffer(buf, dtype="<f8", count=n, offset=off).copy(); off += 8 * n
        return a

    ShfR, ShfA, ShfZ = f64(nR), f64(nA), f64(nZ)
    symbols = []
    for _s in range(S):
        symbols.append(buf[off:off + 4].rstrip(b"\0").decode("ascii")); off += 4
    sae = f64(S)
    dims = []
    for _s in range(S):
        (nl,) = struct.unpack_from("<i", buf, off); off += 4
        dims.append(list(struct.unpack_from(f"<{nl + 1}i", buf, off))); off += 4 * (nl + 1)
    weights = []
    for _m in range(M):
        per_s = []
        for d in dims:
            per_l = []
            for l in range(len(d) - 1):
                n = d[l + 1] * d[l]
                W = np.frombuffer(buf, dtype="<f4", count=n, offset=off).reshape(d[l + 1], d[l]).copy(); off += 4 * n
                b = np.frombuffer(buf, dtype="<f4", count=d[l + 1], offset=off).copy(); off += 4 * d[l + 1]
                per_l.append((W, b))
            per_s.append(per_l)
        weights.append(per_s)
    repulsion = None
    if rep:
        cutoff, fn, _pad = struct.unpack_from("<dii", buf, off); off += 16
        repulsion = {"cutoff": cutoff, "cutoff_fn": fn}
        for k in ("y_ab", "sqrt_alpha_ab", "k_rep_ab"):
            repulsion[k] = f64(S * S).reshape(S, S)
    if off != len(buf):
        raise ValueError(f"{path}: trailing bytes ({len(buf) - off})")
    p = AniParams(symbols, Rcr, Rca, EtaR, EtaA, Zeta, ShfR, ShfA, ShfZ, sae, dims, weights, bool(rep), repulsion)
    p.validate()
    return p
