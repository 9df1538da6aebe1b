"""Python mirror of the reference's runner `class ANI` (src/ani_csrc/ani.h:11-84) over the C ABI.

    ANI(model_file, local_rank, use_num_models=-1, use_cuaev=True, use_fullnbr=True, use_single=True)
    .compute(species, coordinates, <neighbour data>, nlocal, ago, atomic, vflag)
        -> (energy, force[ntotal,3], atomic_energies | None, virial[3,3] | None)     (kcal/mol, kcal/mol/A)

Same argument meaning and error behaviour as the C++ class: neighbour data and species are only consumed when
`ago == 0` (cached on the device otherwise, ani.cpp:122-137,213-229); a load failure raises RuntimeError
(ani.cpp:93-96).  There is no CPU path: `local_rank == -1` ("cpu") raises.  The library is loaded with ctypes
from lammps_ani_b200/libani_b200.so, and importing this module fails loudly if it has not been built.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libani_b200.so")
HARTREE2KCALMOL = 627.5094738898777

MLP_FP32_SIMT, MLP_TF32X3, MLP_TF32, MLP_F16X3 = 0, 1, 2, 3

_lib = None


def load_library():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(f"{LIB_PATH} is missing: build it with `python -m lammps_ani_b200.build` "
                           "(there is no Python/CPU fallback)")
    L = C.CDLL(LIB_PATH)
    vp, i32, i64, dp = C.c_void_p, C.c_int, C.c_int64, C.POINTER(C.c_double)
    L.ani_last_error.restype = C.c_char_p
    L.ani_abi_version.restype = i32
    L.ani_model_load.argtypes = [C.c_char_p, C.POINTER(vp)]
    L.ani_model_free.argtypes = [vp]
    L.ani_model_info.argtypes = [vp, C.POINTER(i32), C.POINTER(i32), C.POINTER(i32), dp, dp]
    L.ani_ctx_create.argtypes = [vp, i32, i32, i32, i32, i32, C.POINTER(vp)]
    L.ani_ctx_destroy.argtypes = [vp]
    L.ani_ctx_set_mlp_mode.argtypes = [vp, i32]
    L.ani_ctx_set_option.argtypes = [vp, C.c_char_p, i32]
    L.ani_ctx_set_stream.argtypes = [vp, vp]
    L.ani_ctx_synchronize.argtypes = [vp]
    L.ani_ctx_reset_stream.argtypes = [vp]
    L.ani_ctx_enable_timing.argtypes = [vp, i32]
    L.ani_last_timing.argtypes = [vp, C.POINTER(C.c_float), C.POINTER(i32)]
    L.ani_set_neighbors_full.argtypes = [vp, i32, i32, vp, i32, vp, vp, vp, i64, i32]
    L.ani_set_neighbors_half.argtypes = [vp, i32, i32, vp, i64, vp, i32]
    L.ani_set_neighbors_table.argtypes = [vp, i32, i32, vp, i32, i32, vp, vp, vp, i64, i64, i32, i32]
    L.ani_device_count.argtypes = [C.POINTER(i32)]
    L.ani_build_neighbors.argtypes = [vp, i32, i32, vp, vp, i32, C.c_double, vp, vp, C.POINTER(i64)]
    L.ani_debug_get_neighbors.argtypes = [vp, vp, vp]
    L.ani_ctx_release_host_buffers.argtypes = [vp]
    L.ani_ctx_release_host_buffers.restype = i32
    L.ani_compute.argtypes = [vp, vp, vp, vp, vp, vp]
    L.ani_compute_device.argtypes = [vp, vp, i32, vp, i32, vp, vp, vp]
    L.ani_debug_get_aev.argtypes = [vp, vp, vp]
    L.ani_debug_tc_gemm.argtypes = [i32, i32, i32, i32, vp, vp, vp, i32]
    L.ani_debug_tc_gemm.restype = i32
    L.ani_md_nve_initial.argtypes = [vp, vp, vp, vp, vp, i32, C.c_double, C.c_double]
    L.ani_md_nve_final.argtypes = [vp, vp, vp, vp, i32, C.c_double]
    L.ani_md_langevin.argtypes = [vp, vp, vp, vp, vp, i32, C.c_double, C.c_double, C.c_uint64, i64]
    L.ani_md_sum_mv2.argtypes = [vp, vp, vp, i32, dp]
    L.ani_md_sum_mv2_static.argtypes = [vp, vp, vp, i32, vp]
    L.ani_md_max_disp2.argtypes = [vp, vp, vp, i32, dp]
    L.ani_ctx_pin_host.argtypes = [vp, vp, C.c_size_t]
    L.ani_md_ghost_x.argtypes = [vp, vp, vp, vp, i32, i32]
    L.ani_md_fold_ghost_f.argtypes = [vp, vp, vp, i32, i32]
    L.ani_halo_unique_id.argtypes = [C.c_char_p]
    L.ani_halo_plan_create.argtypes = [vp, C.c_char_p, i32, i32, i32, vp, vp, vp, vp, C.POINTER(vp)]
    L.ani_halo_plan_destroy.argtypes = [vp]
    L.ani_halo_plan_destroy.restype = None
    L.ani_halo_forward.argtypes = [vp, vp]
    L.ani_compute_device_halo.argtypes = [vp, vp, vp, vp, vp, vp, vp]
    L.ani_halo_reverse.argtypes = [vp, vp]
    L.ani_halo_pack_x.argtypes = [vp, vp, i32, vp, vp, i64, vp]
    L.ani_halo_unpack_add_f.argtypes = [vp, vp, i32, vp, vp, i64]
    for name in ("ani_model_load", "ani_model_info", "ani_ctx_create", "ani_ctx_set_mlp_mode", "ani_ctx_set_option", "ani_ctx_set_stream",
                 "ani_ctx_reset_stream", "ani_ctx_synchronize", "ani_ctx_enable_timing", "ani_last_timing", "ani_set_neighbors_full",
                 "ani_set_neighbors_half", "ani_set_neighbors_table", "ani_device_count", "ani_compute",
                 "ani_compute_device", "ani_debug_get_aev", "ani_halo_pack_x", "ani_halo_unpack_add_f", "ani_build_neighbors",
                 "ani_debug_get_neighbors", "ani_md_nve_initial", "ani_md_nve_final", "ani_md_langevin", "ani_md_sum_mv2", "ani_md_sum_mv2_static", "ani_md_max_disp2", "ani_ctx_pin_host",
                 "ani_md_ghost_x", "ani_md_fold_ghost_f", "ani_halo_unique_id", "ani_halo_plan_create", "ani_halo_forward",
                 "ani_halo_reverse", "ani_compute_device_halo"):
        getattr(L, name).restype = i32
    _lib = L
    return L


EXPORTED_SYMBOLS = [
    "ani_model_load", "ani_model_free", "ani_model_info", "ani_ctx_create", "ani_ctx_destroy", "ani_ctx_set_mlp_mode",
    "ani_ctx_set_stream", "ani_ctx_reset_stream", "ani_ctx_synchronize", "ani_set_neighbors_full", "ani_set_neighbors_half", "ani_compute",
    "ani_compute_device", "ani_halo_pack_x", "ani_halo_unpack_add_f", "ani_debug_get_aev", "ani_last_timing",
    "ani_ctx_enable_timing", "ani_last_error", "ani_abi_version", "ani_debug_tc_gemm", "ani_set_neighbors_table",
    "ani_device_count", "ani_build_neighbors", "ani_debug_get_neighbors", "ani_ctx_release_host_buffers", "ani_ctx_set_option",
    "ani_md_nve_initial", "ani_md_nve_final", "ani_md_langevin", "ani_md_sum_mv2", "ani_md_sum_mv2_static", "ani_md_max_disp2", "ani_ctx_pin_host", "ani_md_ghost_x", "ani_md_fold_ghost_f",
    "ani_dev_alloc", "ani_dev_free", "ani_dev_upload", "ani_dev_download", "ani_dev_zero",
    "ani_halo_unique_id", "ani_halo_plan_create", "ani_halo_plan_destroy", "ani_halo_forward", "ani_halo_reverse", "ani_compute_device_halo",
]


def debug_tc_gemm(A, B, three_pass=True, device: int = 0):
    """C = A @ B.T through the tcgen05 GEMM kernels (test hook): three_pass False = TF32, True = TF32x3, 2 = F16x3."""
    A = np.ascontiguousarray(A, dtype=np.float32)
    B = np.ascontiguousarray(B, dtype=np.float32)
    Cm = np.empty((A.shape[0], B.shape[0]), dtype=np.float32)
    _check(load_library().ani_debug_tc_gemm(device, A.shape[0], B.shape[0], A.shape[1], A.ctypes.data, B.ctypes.data,
                                            Cm.ctypes.data, int(three_pass)))
    return Cm


def _check(rc: int):
    if rc != 0:
        raise RuntimeError("libani_b200: " + load_library().ani_last_error().decode())


def _ptr(a):
    return None if a is None else a.ctypes.data


class ANI:
    def __init__(self, model_file: str, local_rank: int, use_num_models: int = -1, use_cuaev: bool = True,
                 use_fullnbr: bool = True, use_single: bool = True):
        L = load_library()
        self._L = L
        self._model = C.c_void_p()
        self._ctx = C.c_void_p()
        _check(L.ani_model_load(os.fsencode(model_file), C.byref(self._model)))
        ns, al, nm = C.c_int(), C.c_int(), C.c_int()
        rcr, rca = C.c_double(), C.c_double()
        _check(L.ani_model_info(self._model, C.byref(ns), C.byref(al), C.byref(nm), C.byref(rcr), C.byref(rca)))
        self.num_species, self.aev_length, self.num_models = ns.value, al.value, nm.value
        self.Rcr, self.Rca = rcr.value, rca.value
        self.use_fullnbr, self.use_cuaev, self.use_single = use_fullnbr, use_cuaev, use_single
        self.use_num_models = self.num_models if use_num_models == -1 else use_num_models
        try:
            _check(L.ani_ctx_create(self._model, local_rank, use_num_models, int(use_cuaev), int(use_fullnbr),
                                    int(use_single), C.byref(self._ctx)))
        except Exception:
            L.ani_model_free(self._model)
            self._model = C.c_void_p()
            raise
        self.nlocal = self.ntotal = 0

    def close(self):
        if getattr(self, "_ctx", None) and self._ctx:
            self._L.ani_ctx_destroy(self._ctx)
            self._ctx = C.c_void_p()
        if getattr(self, "_model", None) and self._model:
            self._L.ani_model_free(self._model)
            self._model = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- knobs
    def set_mlp_mode(self, mode: int):
        _check(self._L.ani_ctx_set_mlp_mode(self._ctx, mode))

    def set_option(self, name: str, value: int):
        """Per-context switch (include/ani_b200.h: "mlp_fused", "pin_user", "aev_two_pass", "gemm_ctas", "fused_prof", "halo_overlap")."""
        _check(self._L.ani_ctx_set_option(self._ctx, name.encode(), int(value)))

    def release_host_buffers(self):
        _check(self._L.ani_ctx_release_host_buffers(self._ctx))

    def set_stream(self, cuda_stream_ptr: Optional[int]):
        """Run on this cudaStream_t (0 = the legacy default stream); None = back to the context's own stream."""
        if cuda_stream_ptr is None:
            _check(self._L.ani_ctx_reset_stream(self._ctx))
        else:
            _check(self._L.ani_ctx_set_stream(self._ctx, C.c_void_p(cuda_stream_ptr)))

    def synchronize(self):
        _check(self._L.ani_ctx_synchronize(self._ctx))

    def enable_timing(self, on: bool = True):
        _check(self._L.ani_ctx_enable_timing(self._ctx, int(on)))

    def last_timing(self):
        ms = (C.c_float * 5)()
        n = C.c_int()
        _check(self._L.ani_last_timing(self._ctx, ms, C.byref(n)))
        return list(ms), n.value

    # ---- neighbour data (only when the list was rebuilt)
    def set_neighbors_full(self, species, ilist, numneigh, jlist, nlocal: int):
        species = np.ascontiguousarray(species, dtype=np.int32)
        ilist = np.ascontiguousarray(ilist, dtype=np.int32)
        numneigh = np.ascontiguousarray(numneigh, dtype=np.int32)
        jlist = np.ascontiguousarray(jlist, dtype=np.int32)
        self.nlocal, self.ntotal = int(nlocal), int(species.shape[0])
        _check(self._L.ani_set_neighbors_full(self._ctx, self.nlocal, self.ntotal, _ptr(species), int(ilist.shape[0]),
                                              _ptr(ilist), _ptr(numneigh), _ptr(jlist), int(jlist.shape[0]), 0))

    def build_neighbors(self, species, x, nlocal: int, cutoff: float = 7.1, lo=None, hi=None, x_dev_ptr: int = 0) -> int:
        """Full neighbour list built on the device and installed; returns the number of stored pairs.
        x: host float64 [ntotal, 3] (or pass x_dev_ptr, a device pointer to the same layout, plus lo / hi)."""
        species = np.ascontiguousarray(species, dtype=np.int32)
        self.nlocal, self.ntotal = int(nlocal), int(species.shape[0])
        if x_dev_ptr:
            xp, on_dev = C.c_void_p(x_dev_ptr), 1
        else:
            x = np.ascontiguousarray(x, dtype=np.float64).reshape(-1, 3)
            if lo is None:
                lo = x.min(0) - 1e-6
            if hi is None:
                hi = x.max(0) + 1e-6
            xp, on_dev = _ptr(x), 0
        lo = np.ascontiguousarray(lo, dtype=np.float64)
        hi = np.ascontiguousarray(hi, dtype=np.float64)
        n = C.c_int64()
        _check(self._L.ani_build_neighbors(self._ctx, self.nlocal, self.ntotal, _ptr(species), xp, on_dev, float(cutoff),
                                           _ptr(lo), _ptr(hi), C.byref(n)))
        self.npairs = int(n.value)
        return self.npairs

    def debug_neighbors(self):
        nn = np.empty(self.nlocal, dtype=np.int32)
        jl = np.empty(max(self.npairs, 1), dtype=np.int32)
        _check(self._L.ani_debug_get_neighbors(self._ctx, _ptr(nn), _ptr(jl)))
        return nn, jl[: self.npairs]

    def set_neighbors_half(self, species, atom_index12, nlocal: int):
        species = np.ascontiguousarray(species, dtype=np.int32)
        a = np.ascontiguousarray(atom_index12, dtype=np.int64).reshape(2, -1)
        self.nlocal, self.ntotal = int(nlocal), int(species.shape[0])
        _check(self._L.ani_set_neighbors_half(self._ctx, self.nlocal, self.ntotal, _ptr(species), int(a.shape[1]),
                                              _ptr(a), 0))

    def set_neighbors_table(self, types, ilist, numneigh, table, nlocal: int, layout_left: bool = False,
                            neighmask: int = 0x1FFFFFFF):
        """Padded [nrows][max_neighs] table as the KOKKOS package keeps it (pair_ani_kokkos.cpp:143-169); `types` are
        LAMMPS types (1-based).  layout_left: the table memory is column-major."""
        types = np.ascontiguousarray(types, dtype=np.int32)
        ilist = np.ascontiguousarray(ilist, dtype=np.int32)
        numneigh = np.ascontiguousarray(numneigh, dtype=np.int32)
        table = np.asarray(table, dtype=np.int32)
        nrows, maxn = table.shape
        mem = np.ascontiguousarray(table.T if layout_left else table)
        rs, cs = (1, nrows) if layout_left else (maxn, 1)
        self.nlocal, self.ntotal = int(nlocal), int(types.shape[0])
        _check(self._L.ani_set_neighbors_table(self._ctx, self.nlocal, self.ntotal, _ptr(types), 1, int(ilist.shape[0]),
                                               _ptr(ilist), _ptr(numneigh), _ptr(mem), rs, cs, neighmask, 0))

    # ---- one step, host arrays (ani.cpp:100-265)
    def compute(self, species, coordinates, *nbr, nlocal: int, ago: int = 0, atomic: bool = False, vflag: bool = False,
                out_force=None):
        """out_force: optional caller-owned float64 [ntotal, 3] array that receives the forces (the reference's
        `std::vector<double>& out_force`, ani.h); a buffer that stays at the same address from step to step is page-locked by
        the library and written by DMA."""
        if ago == 0:
            if self.use_fullnbr:
                ilist, jlist, numneigh = nbr      # argument order of the reference's full overload
                self.set_neighbors_full(species, ilist, numneigh, jlist, nlocal)
            else:
                (atom_index12,) = nbr
                self.set_neighbors_half(species, atom_index12, nlocal)
        x = np.ascontiguousarray(coordinates, dtype=np.float64).reshape(-1, 3)
        if x.shape[0] != self.ntotal:
            raise ValueError("coordinates do not match the cached neighbour data (ago != 0 with a different system)")
        energy = C.c_double()
        if out_force is not None:
            if out_force.dtype != np.float64 or out_force.shape != (self.ntotal, 3) or not out_force.flags.c_contiguous:
                raise ValueError("out_force must be a C-contiguous float64 array of shape (ntotal, 3)")
            force = out_force
        else:
            force = np.empty((self.ntotal, 3), dtype=np.float64)
        eatom = np.empty(self.nlocal, dtype=np.float64) if atomic else None
        virial = np.empty(9, dtype=np.float64) if vflag else None
        _check(self._L.ani_compute(self._ctx, _ptr(x), C.addressof(energy), _ptr(force), _ptr(eatom), _ptr(virial)))
        return energy.value, force, eatom, (virial.reshape(3, 3) if vflag else None)

    # ---- one step, device arrays (ani.cpp:268-316): raw CUDA pointers (e.g. torch.Tensor.data_ptr())
    def compute_device(self, x_ptr: int, x_is_float: bool, f_ptr: int, f_is_float: bool, eatom_ptr: int = 0,
                       want_energy: bool = True, want_virial: bool = False):
        energy = C.c_double()
        virial = np.empty(9, dtype=np.float64) if want_virial else None
        _check(self._L.ani_compute_device(self._ctx, C.c_void_p(x_ptr), int(x_is_float), C.c_void_p(f_ptr),
                                          int(f_is_float), C.c_void_p(eatom_ptr or 0),
                                          C.addressof(energy) if want_energy else None, _ptr(virial)))
        return (energy.value if want_energy else None), (virial.reshape(3, 3) if want_virial else None)

    def compute_device_halo(self, plan, x_ptr: int, f_ptr: int, eatom_ptr: int = 0, want_energy: bool = True, want_virial: bool = False):
        """ani_halo_forward + ani_compute_device + ani_halo_reverse in one call (x, f: device arrays of doubles); the library
        overlaps the exchange with the interior atoms' AEV kernels when the list was built on the device."""
        energy = C.c_double()
        virial = np.empty(9, dtype=np.float64) if want_virial else None
        _check(self._L.ani_compute_device_halo(self._ctx, plan, C.c_void_p(x_ptr), C.c_void_p(f_ptr), C.c_void_p(eatom_ptr or 0),
                                               C.addressof(energy) if want_energy else None, _ptr(virial)))
        return (energy.value if want_energy else None), (virial.reshape(3, 3) if want_virial else None)

    # ---- device-resident integration (fix nve / fix langevin on the GPU; include/ani_b200.h ani_md_*): raw device pointers
    def md_nve_initial(self, x_ptr, v_ptr, f_ptr, mass_ptr, n, dtv, dtf):
        _check(self._L.ani_md_nve_initial(self._ctx, C.c_void_p(x_ptr), C.c_void_p(v_ptr), C.c_void_p(f_ptr), C.c_void_p(mass_ptr),
                                          int(n), float(dtv), float(dtf)))

    def md_nve_final(self, v_ptr, f_ptr, mass_ptr, n, dtf):
        _check(self._L.ani_md_nve_final(self._ctx, C.c_void_p(v_ptr), C.c_void_p(f_ptr), C.c_void_p(mass_ptr), int(n), float(dtf)))

    def md_langevin(self, v_ptr, f_ptr, mass_ptr, tag_ptr, n, g1, g2, seed, step):
        _check(self._L.ani_md_langevin(self._ctx, C.c_void_p(v_ptr), C.c_void_p(f_ptr), C.c_void_p(mass_ptr), C.c_void_p(tag_ptr or 0),
                                       int(n), float(g1), float(g2), int(seed), int(step)))

    def md_sum_mv2(self, v_ptr, mass_ptr, n) -> float:
        out = C.c_double()
        _check(self._L.ani_md_sum_mv2(self._ctx, C.c_void_p(v_ptr), C.c_void_p(mass_ptr), int(n), C.byref(out)))
        return out.value

    def md_sum_mv2_static(self, v_ptr, mass_ptr, n, out_ptr) -> None:
        """Allocation-free variant (include/ani_b200.h): the sum lands in the DEVICE double at out_ptr, in stream order."""
        _check(self._L.ani_md_sum_mv2_static(self._ctx, C.c_void_p(v_ptr), C.c_void_p(mass_ptr), int(n), C.c_void_p(out_ptr)))

    # ---- ghost exchange inside the library (NCCL send / recv groups on the context's stream)
    @staticmethod
    def halo_unique_id() -> bytes:
        buf = C.create_string_buffer(128)
        _check(load_library().ani_halo_unique_id(buf))
        return buf.raw

    def halo_plan_create(self, unique_id: Optional[bytes], nranks: int, rank: int, nlocal: int, send_index, send_shift, send_counts,
                         recv_counts):
        si = np.ascontiguousarray(send_index, dtype=np.int64)
        sh = np.ascontiguousarray(send_shift, dtype=np.float64).reshape(-1, 3)
        sc = np.ascontiguousarray(send_counts, dtype=np.int32)
        rc = np.ascontiguousarray(recv_counts, dtype=np.int32)
        plan = C.c_void_p()
        _check(self._L.ani_halo_plan_create(self._ctx, unique_id, int(nranks), int(rank), int(nlocal), _ptr(si), _ptr(sh), _ptr(sc), _ptr(rc),
                                            C.byref(plan)))
        return plan

    def halo_plan_destroy(self, plan):
        self._L.ani_halo_plan_destroy(plan)

    def halo_forward(self, plan, x_ptr):
        _check(self._L.ani_halo_forward(plan, C.c_void_p(x_ptr)))

    def halo_reverse(self, plan, f_ptr):
        _check(self._L.ani_halo_reverse(plan, C.c_void_p(f_ptr)))

    def halo_pack_x(self, x_ptr, is_float, index_ptr, shift_ptr, n, out_ptr):
        _check(self._L.ani_halo_pack_x(self._ctx, C.c_void_p(x_ptr), int(is_float), C.c_void_p(index_ptr),
                                       C.c_void_p(shift_ptr or 0), int(n), C.c_void_p(out_ptr)))

    def halo_unpack_add_f(self, f_ptr, is_float, index_ptr, buf_ptr, n):
        _check(self._L.ani_halo_unpack_add_f(self._ctx, C.c_void_p(f_ptr), int(is_float), C.c_void_p(index_ptr),
                                             C.c_void_p(buf_ptr), int(n)))

    def debug_aev(self, want_daev: bool = True):
        aev = np.empty((self.nlocal, self.aev_length), dtype=np.float64)
        daev = np.empty_like(aev) if want_daev else None
        _check(self._L.ani_debug_get_aev(self._ctx, _ptr(aev), _ptr(daev)))
        return aev, daev
