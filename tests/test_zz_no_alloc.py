"""ani_md_sum_mv2_static: the allocation-free variant of ani_md_sum_mv2 (static device scratch, include/ani_b200.h).

Collected last on purpose.  Everything goes through the C ABI: the caller's buffers come
from ani_dev_alloc BEFORE the calls under test, the new entry point itself allocates nothing and does not synchronise, and its
result is compared with the existing entry point's on two calls in a row.  The second call is smaller than the first, so the
shared scratch still holds the first call's partial sums: a result that picked up stale partials would show.
"""
import ctypes as C

import numpy as np
import pytest

from lammps_ani_b200 import runner

pytestmark = pytest.mark.gpu

REL_TOL = 1e-12          # fp64 sums of the same per-block partials; only the order of the last <= 1024 additions differs


def _bind(L):
    vp = C.c_void_p
    L.ani_dev_alloc.argtypes = [vp, C.c_size_t, C.POINTER(vp)]
    L.ani_dev_free.argtypes = [vp, vp]
    L.ani_dev_upload.argtypes = [vp, vp, vp, C.c_size_t]
    L.ani_dev_download.argtypes = [vp, vp, vp, C.c_size_t]
    for n in ("ani_dev_alloc", "ani_dev_free", "ani_dev_upload", "ani_dev_download"):
        getattr(L, n).restype = C.c_int


def test_sum_mv2_static_matches_allocating_entry_point(ani2x_synth_path):
    ani = runner.ANI(ani2x_synth_path, 0, -1, True, True, True)
    L, ctx = ani._L, ani._ctx
    _bind(L)
    rng = np.random.default_rng(11)
    nmax = 300_000
    v = np.ascontiguousarray(rng.normal(0.0, 0.02, size=(nmax, 3)))
    mass = np.ascontiguousarray(rng.choice([1.008, 12.0107, 14.0067, 15.999], size=nmax))

    def alloc(nbytes):
        p = C.c_void_p()
        assert L.ani_dev_alloc(ctx, nbytes, C.byref(p)) == 0, L.ani_last_error()
        return p

    d_v, d_m, d_out = alloc(v.nbytes), alloc(mass.nbytes), alloc(8)
    try:
        assert L.ani_dev_upload(ctx, d_v, v.ctypes.data_as(C.c_void_p), v.nbytes) == 0
        assert L.ani_dev_upload(ctx, d_m, mass.ctypes.data_as(C.c_void_p), mass.nbytes) == 0
        for n in (nmax, 70_001):                       # two calls in a row; the second leaves most of the scratch stale
            ref = ani.md_sum_mv2(d_v.value, d_m.value, n)
            ani.md_sum_mv2_static(d_v.value, d_m.value, n, d_out.value)
            got = C.c_double(-1.0)
            assert L.ani_dev_download(ctx, C.byref(got), d_out, 8) == 0, L.ani_last_error()
            host = float(np.sum(mass[:n] * np.sum(v[:n] * v[:n], axis=1)))
            assert abs(got.value - ref) <= REL_TOL * abs(ref), (n, got.value, ref)
            assert abs(got.value - host) <= 1e-10 * abs(host), (n, got.value, host)
        # n = 0 is a valid call and gives 0; a size beyond the stated maximum is refused with a message, not truncated
        ani.md_sum_mv2_static(d_v.value, d_m.value, 0, d_out.value)
        got = C.c_double(-1.0)
        assert L.ani_dev_download(ctx, C.byref(got), d_out, 8) == 0
        assert got.value == 0.0
        with pytest.raises(RuntimeError, match="ani_md_sum_mv2_static"):
            ani.md_sum_mv2_static(d_v.value, d_m.value, (1 << 24) + 1, d_out.value)
    finally:
        for p in (d_v, d_m, d_out):
            L.ani_dev_free(ctx, p)
