"""NVE energy-drift check of the device-resident step (BASELINE.json north_star: every tensor-core MLP variant needs its
own stated tolerance plus an NVE energy-drift check; reference protocol tests/in.lammps:55-63: short NVE run, dt 0.1 fs).

Real ANI-1xnr weights (in tree), periodic water box, velocity Verlet on the device (lammps_ani_b200/nve.py), fixed 7.1 A
neighbour list (1.9 A skin; atoms move < 0.2 A during the run).  The conserved quantity is pe + ke; its excursion is
compared between the strict fp32 FFMA MLP and the tcgen05 f16x3 / tf32x3 modes, and bounded absolutely.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
if not torch.cuda.is_available():
    pytest.skip("needs a CUDA device", allow_module_level=True)

from lammps_ani_b200 import ddstep, domain, nve, runner  # noqa: E402

NSTEPS, DT = 150, 0.1
# kcal/mol per atom.  DRIFT: max |E(t) - E(0)| over the run, dominated by the velocity-Verlet discretisation error at this dt and
# the same in every mode (measured on B200: 1.1e-4 in all four).  DEV: max |(E - E0)_mode(t) - (E - E0)_fp32-FFMA(t)|, what the MLP
# arithmetic adds on top (measured on B200: f16x3 3.7e-6, tf32x3 6.3e-6, tf32 single pass 3.0e-4; KE swing 0.85).
DRIFT_TOL = 4e-4
DEV_TOL = {"simt": 0.0, "f16x3": 2e-5, "tf32x3": 2e-5, "tf32": 4e-4}
_cache = {}


def run_mode(path, mode, rebuild_every=0):
    x, sp, box = domain.make_water_box((6, 6, 6), seed=21, jitter=0.1)
    sub = domain.decompose(x, sp, box, (1, 1, 1))[0]
    halo = domain.replica_halo(sub, box, (1, 1, 1), 0)
    il, nn, jl = domain.full_neighbor_list(sub.x, sub.nlocal)
    ani = runner.ANI(path, 0, -1, True, True, True)
    ani.set_mlp_mode({"simt": runner.MLP_FP32_SIMT, "tf32x3": runner.MLP_TF32X3, "tf32": runner.MLP_TF32,
                      "f16x3": runner.MLP_F16X3}[mode])
    dev = torch.device("cuda:0")
    with torch.cuda.device(dev):
        step = ddstep.GpuDomainStep(ani, sub, halo, dev, il, nn, jl)
        masses = np.where(sub.species[: sub.nlocal] == 0, nve.MASS["H"], nve.MASS["O"])
        md = nve.DeviceNVE(step, masses, DT, temperature=300.0, seed=7, rebuild_every=rebuild_every)
        traj = md.run(NSTEPS, energy_every=5)
        torch.cuda.synchronize()
    e = np.array([pe + ke for _, pe, ke in traj])
    ke = np.array([k for _, _, k in traj])
    return sub.nlocal, e, ke


@pytest.mark.parametrize("mode", ["simt", "f16x3", "tf32x3", "tf32"])
def test_nve_energy_drift(ani1xnr_path, mode):
    if "simt" not in _cache:
        _cache["simt"] = run_mode(ani1xnr_path, "simt")
    n, e, ke = _cache["simt"] if mode == "simt" else run_mode(ani1xnr_path, mode)
    e_ref = _cache["simt"][1]
    drift = np.abs(e - e[0]).max() / n
    dev = np.abs((e - e[0]) - (e_ref - e_ref[0])).max() / n      # a constant offset between modes is not drift
    swing = (ke.max() - ke.min()) / n
    print(f"NVE {mode}: {n} atoms, {NSTEPS} steps of {DT} fs: max |E-E0| = {drift:.3e}, max |dE-dE_fp32| = {dev:.3e}, "
          f"KE swing {swing:.3e} kcal/mol/atom, E0 = {e[0]:.6f}")
    assert np.isfinite(e).all()
    assert swing > 50 * drift, "the exchange between potential and kinetic energy must dwarf the drift"
    assert drift <= DRIFT_TOL
    assert dev <= DEV_TOL[mode]


def test_nve_with_device_neighbor_rebuilds(ani1xnr_path):
    """Same run with the list rebuilt on the device every 25 steps (rows f2 + f3 together): the conserved quantity must not
    notice (the list radius exceeds the cutoff by 1.9 A, so the pair SET inside the cutoff is the same either way)."""
    if "simt" not in _cache:
        _cache["simt"] = run_mode(ani1xnr_path, "simt")
    n, e, ke = run_mode(ani1xnr_path, "simt", rebuild_every=25)
    e_ref = _cache["simt"][1]
    dev = np.abs((e - e[0]) - (e_ref - e_ref[0])).max() / n
    print(f"NVE with rebuilds: max |dE-dE_fixed-list| = {dev:.3e} kcal/mol/atom")
    assert dev <= 5e-6


def test_max_disp2_and_pinned_host_arrays(ani2x_synth_path):
    """ani_md_max_disp2 = Neighbor::check_distance's max |x - x0|^2 on device arrays; ani_ctx_pin_host page-locks a caller array
    (twice is accepted), ani_dev_upload / download through it give the same bytes, release drops the registration."""
    import ctypes as C
    ani = runner.ANI(ani2x_synth_path, 0, -1, True, True, True)
    L = ani._L
    rng = np.random.default_rng(5)
    n = 100_003
    x0 = rng.uniform(0, 50, size=(n, 3))
    x = x0 + rng.normal(0, 0.05, size=(n, 3))
    x[77_001] += np.array([0.9, -0.8, 0.7])
    dev = torch.device("cuda:0")
    tx, tx0 = torch.from_numpy(x).to(dev), torch.from_numpy(x0).to(dev)
    out = C.c_double()
    assert L.ani_md_max_disp2(ani._ctx, C.c_void_p(tx.data_ptr()), C.c_void_p(tx0.data_ptr()), n, C.byref(out)) == 0
    ref = float(((x - x0) ** 2).sum(axis=1).max())
    assert abs(out.value - ref) <= 1e-14 * ref
    assert L.ani_md_max_disp2(ani._ctx, C.c_void_p(tx0.data_ptr()), C.c_void_p(tx0.data_ptr()), n, C.byref(out)) == 0 and out.value == 0.0
    # page-locked caller array: upload + download round trip
    L.ani_dev_upload.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t]
    L.ani_dev_download.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t]
    host = np.ascontiguousarray(x.copy())
    back = np.zeros_like(host)
    for arr in (host, host, back):                      # the second registration of `host` is accepted as is
        assert L.ani_ctx_pin_host(ani._ctx, arr.ctypes.data_as(C.c_void_p), arr.nbytes) == 0, L.ani_last_error()
    buf = torch.empty(n * 3, dtype=torch.float64, device=dev)
    assert L.ani_dev_upload(ani._ctx, C.c_void_p(buf.data_ptr()), host.ctypes.data_as(C.c_void_p), host.nbytes) == 0
    assert L.ani_dev_download(ani._ctx, back.ctypes.data_as(C.c_void_p), C.c_void_p(buf.data_ptr()), back.nbytes) == 0
    assert np.array_equal(back, host)
    ani.release_host_buffers()
