"""ctypes front end of the CPU oracle (oracle/ani_oracle.c).  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import
this module.  Mirrors the call signature of the reference runner's two host overloads
(/root/reference/src/ani_csrc/ani.cpp:100-180 half list, :183-265 full list): species, coordinates,
neighbour data, nlocal -> energy (kcal/mol), forces [ntotal,3] (kcal/mol/A), atomic energies, virial.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from typing import Optional

import numpy as np

HARTREE2KCALMOL = 627.5094738898777  # /root/reference/src/ani_csrc/ani.h:9
_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def build(force: bool = False) -> str:
    so = os.path.join(_HERE, "libani_oracle.so")
    src = os.path.join(_HERE, "ani_oracle.c")
    if force or not os.path.exists(so) or (os.path.exists(src) and os.path.getmtime(src) > os.path.getmtime(so)):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return so


def lib():
    global _LIB
    if _LIB is None:
        L = C.CDLL(build())
        L.ani_oracle_model_load.restype = C.c_void_p
        L.ani_oracle_model_load.argtypes = [C.c_char_p]
        L.ani_oracle_model_free.argtypes = [C.c_void_p]
        for fn in ("ani_oracle_aev_length", "ani_oracle_num_models", "ani_oracle_num_species"):
            getattr(L, fn).argtypes = [C.c_void_p]
            getattr(L, fn).restype = C.c_int
        L.ani_oracle_compute.restype = C.c_int
        L.ani_oracle_compute.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                         C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                         C.c_void_p, C.c_void_p, C.c_int, C.c_int]
        L.ani_oracle_set_celu_exact.argtypes = [C.c_int]
        _LIB = L
    return _LIB


def set_celu_exact(on: bool) -> None:
    """False (default): CELU backward as torch autograd evaluates it (see dcelu in ani_oracle.c)."""
    lib().ani_oracle_set_celu_exact(int(on))


def csr_from_full(nlocal: int, ilist, numneigh_per_entry, jlist):
    """LAMMPS full list (as flattened by pair_ani.cpp:131-150) -> CSR over local atoms 0..nlocal-1."""
    ilist = np.asarray(ilist, dtype=np.int64)
    cnt_e = np.asarray(numneigh_per_entry, dtype=np.int64)
    jlist = np.asarray(jlist, dtype=np.int32)
    starts_e = np.concatenate([[0], np.cumsum(cnt_e)])
    cnt = np.zeros(nlocal, dtype=np.int64)
    cnt[ilist] = cnt_e
    off = np.concatenate([[0], np.cumsum(cnt)]).astype(np.int64)
    nbr = np.empty(int(off[-1]), dtype=np.int32)
    for e, i in enumerate(ilist):  # small/medium lists only; vectorised path below for identity ilist
        nbr[off[i]:off[i + 1]] = jlist[starts_e[e]:starts_e[e + 1]]
    return off, nbr


def csr_from_full_fast(nlocal: int, ilist, numneigh_per_entry, jlist):
    ilist = np.asarray(ilist)
    if len(ilist) == nlocal and np.array_equal(ilist, np.arange(nlocal)):
        off = np.concatenate([[0], np.cumsum(np.asarray(numneigh_per_entry, dtype=np.int64))]).astype(np.int64)
        return off, np.ascontiguousarray(jlist, dtype=np.int32)
    return csr_from_full(nlocal, ilist, numneigh_per_entry, jlist)


def csr_from_half(nlocal: int, atom_index12):
    """Half list int64[2][P] (pair_ani.cpp:143-146) -> CSR over local atoms (each pair seen from both ends)."""
    a = np.asarray(atom_index12, dtype=np.int64).reshape(2, -1)
    src = np.concatenate([a[0], a[1]])
    dst = np.concatenate([a[1], a[0]])
    keep = src < nlocal
    src, dst = src[keep], dst[keep]
    order = np.argsort(src, kind="stable")
    src, dst = src[order], dst[order]
    cnt = np.bincount(src, minlength=nlocal)
    off = np.concatenate([[0], np.cumsum(cnt)]).astype(np.int64)
    return off, dst.astype(np.int32)


class OracleANI:
    """CPU fp64 checker with the reference runner's shape: OracleANI(model).compute(...)."""

    def __init__(self, aniw_path: str, use_num_models: int = -1, nthreads: int = 0):
        self._L = lib()
        self._h = self._L.ani_oracle_model_load(aniw_path.encode())
        if not self._h:
            raise RuntimeError(f"oracle: cannot load {aniw_path}")
        self.aev_length = self._L.ani_oracle_aev_length(self._h)
        self.num_models = self._L.ani_oracle_num_models(self._h)
        self.num_species = self._L.ani_oracle_num_species(self._h)
        self.use_num_models = self.num_models if use_num_models in (-1, None) else int(use_num_models)
        self.nthreads = nthreads

    def __del__(self):
        try:
            if self._h:
                self._L.ani_oracle_model_free(self._h)
                self._h = None
        except Exception:
            pass

    def compute_csr(self, species, coordinates, nlocal: int, off, nbr, *, radial_filter: bool = True,
                    want_aev: bool = False, want_daev: bool = False, aev_only: bool = False,
                    single_coords: bool = False):
        species = np.ascontiguousarray(species, dtype=np.int32)
        x = np.ascontiguousarray(coordinates, dtype=np.float64).reshape(-1, 3)
        if single_coords:  # mirror `.to(dtype)` of the coordinates in single mode (ani.cpp:206-209)
            x = x.astype(np.float32).astype(np.float64)
        ntotal = x.shape[0]
        off = np.ascontiguousarray(off, dtype=np.int64)
        nbr = np.ascontiguousarray(nbr, dtype=np.int32)
        energy = C.c_double(0.0)
        forces = np.zeros((ntotal, 3), dtype=np.float64)
        eatom = np.zeros(nlocal, dtype=np.float64)
        virial = np.zeros(9, dtype=np.float64)
        aev = np.zeros((nlocal, self.aev_length), dtype=np.float64) if (want_aev or aev_only) else None
        daev = np.zeros((nlocal, self.aev_length), dtype=np.float64) if want_daev else None
        rc = self._L.ani_oracle_compute(
            self._h, self.use_num_models, int(radial_filter), nlocal, ntotal, species.ctypes.data, x.ctypes.data,
            off.ctypes.data, nbr.ctypes.data, C.addressof(energy), forces.ctypes.data, eatom.ctypes.data,
            virial.ctypes.data, aev.ctypes.data if aev is not None else None,
            daev.ctypes.data if daev is not None else None, int(aev_only), self.nthreads)
        if rc != 0:
            raise RuntimeError("oracle: bad species index")
        k = HARTREE2KCALMOL
        return dict(energy=energy.value * k, forces=forces * k, eatom=eatom * k, virial=virial.reshape(3, 3) * k,
                    aev=aev, daev=daev)

    def compute_full(self, species, coordinates, ilist, numneigh_per_entry, jlist, nlocal: int, **kw):
        off, nbr = csr_from_full_fast(nlocal, ilist, numneigh_per_entry, jlist)
        return self.compute_csr(species, coordinates, nlocal, off, nbr, **kw)

    def compute_half(self, species, coordinates, atom_index12, nlocal: int, **kw):
        off, nbr = csr_from_half(nlocal, atom_index12)
        return self.compute_csr(species, coordinates, nlocal, off, nbr, **kw)
