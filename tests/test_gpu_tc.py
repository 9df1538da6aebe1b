"""GPU tests of the tcgen05 / TMEM / TMA GEMM path (B200 only)."""
import numpy as np
import pytest

from lammps_ani_b200 import domain
from oracle.oracle import OracleANI

pytestmark = pytest.mark.gpu
EV2KCAL = 23.060548


@pytest.fixture(scope="module")
def runner():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from lammps_ani_b200 import build, runner
    build.build()
    return runner


@pytest.mark.parametrize("R,N,K", [(128, 160, 32), (128, 256, 1008), (300, 192, 256), (1000, 1008, 2048), (77, 144, 96)])
def test_tc_gemm_tf32x3_matches_fp64(runner, R, N, K):
    rng = np.random.default_rng(R + N + K)
    A = rng.standard_normal((R, K)).astype(np.float32)
    B = rng.standard_normal((N, K)).astype(np.float32)
    ref = A.astype(np.float64) @ B.astype(np.float64).T
    scale = np.sqrt(K)
    C3 = runner.debug_tc_gemm(A, B, three_pass=True)
    err3 = np.abs(C3 - ref).max() / scale
    C1 = runner.debug_tc_gemm(A, B, three_pass=False)
    err1 = np.abs(C1 - ref).max() / scale
    fp32 = np.abs((A @ B.T) - ref).max() / scale
    print(f"\nR={R} N={N} K={K}: tf32x3 err {err3:.2e}, tf32 err {err1:.2e}, numpy fp32 err {fp32:.2e}")
    # tcgen05 accumulates in fp32 with truncation, so the split result drifts ~K/8 * 2^-24 relative: measured
    # 2e-6 (K=32) .. 3e-5 (K=1008) of sqrt(K), i.e. 30-700x better than one TF32 pass, ~10x worse than an fp32 FFMA chain
    assert err3 < 1e-4
    assert err1 < 5e-3          # plain TF32
    assert err3 < err1 / 10


@pytest.mark.parametrize("model", ["ani2x_synth", "ani1xnr"])
def test_full_path_tf32x3_parity(runner, ani2x_synth_path, ani1xnr_path, model):
    path = ani2x_synth_path if model == "ani2x_synth" else ani1xnr_path
    x, sp, box = domain.make_water_box((8, 9, 10), seed=7)
    sub = domain.decompose(x, sp, box)[0]
    il, nn, jl = domain.full_neighbor_list(sub.x, sub.nlocal)
    ref = OracleANI(path).compute_full(sub.species, sub.x, il, nn, jl, sub.nlocal, single_coords=True, want_daev=True)
    ani = runner.ANI(path, 0, -1, True, True, True)
    ani.set_mlp_mode(runner.MLP_TF32X3)
    e, f, eatom, virial = ani.compute(sub.species, sub.x, il, jl, nn, nlocal=sub.nlocal, ago=0, atomic=True, vflag=True)
    _, daev = ani.debug_aev()
    print("daev max err", np.abs(daev - ref["daev"]).max(), "force max err", np.abs(f - ref["forces"]).max(),
          "E rel err", abs(e - ref["energy"]) / abs(ref["energy"]))
    assert abs(e - ref["energy"]) <= 1e-6 * abs(ref["energy"])
    assert np.abs(f - ref["forces"]).max() <= 1e-4 * EV2KCAL
    assert np.abs(eatom - ref["eatom"]).max() <= 1e-5 * np.abs(ref["eatom"]).max()
    # the virial sums same-sign pair terms, so the tensor core's truncating accumulation shows up as a ~2-3e-5 relative bias
    assert np.abs(virial - ref["virial"]).max() <= 1e-4 * max(1.0, np.abs(ref["virial"]).max())


def test_full_path_tf32_single_pass_has_its_own_tolerance(runner, ani2x_synth_path):
    """ANI_MLP_TF32 (the reference's LAMMPS_ANI_ALLOW_TF32=1): stated tolerance 5e-3 eV/A on forces, 1e-5 relative on E."""
    x, sp, box = domain.make_water_box((6, 6, 6), seed=2)
    sub = domain.decompose(x, sp, box)[0]
    il, nn, jl = domain.full_neighbor_list(sub.x, sub.nlocal)
    ref = OracleANI(ani2x_synth_path).compute_full(sub.species, sub.x, il, nn, jl, sub.nlocal, single_coords=True)
    ani = runner.ANI(ani2x_synth_path, 0, -1, True, True, True)
    ani.set_mlp_mode(runner.MLP_TF32)
    e, f, _, _ = ani.compute(sub.species, sub.x, il, jl, nn, nlocal=sub.nlocal, ago=0)
    print("tf32: force max err", np.abs(f - ref["forces"]).max(), "E rel", abs(e - ref["energy"]) / abs(ref["energy"]))
    assert abs(e - ref["energy"]) <= 1e-5 * abs(ref["energy"])
    assert np.abs(f - ref["forces"]).max() <= 5e-3 * EV2KCAL


@pytest.mark.parametrize("R,N,K", [(128, 160, 32), (128, 256, 1008), (300, 192, 256), (1000, 1008, 2048), (77, 96, 160),
                                   (5000, 224, 448)])
def test_f16x3_gemm_matches_fp64(runner, R, N, K):
    """fp16 hi/lo planes, three kind::f16 MMAs per k-step, TMA-store epilogue (ragged rows, ragged N tile, ragged last
    k-block).  Same error budget as TF32x3."""
    rng = np.random.default_rng(R + N + K)
    A = (rng.standard_normal((R, K)) * np.exp(rng.uniform(-4, 2, size=(R, 1)))).astype(np.float32)   # rows of mixed magnitude
    B = (rng.standard_normal((N, K)) * 0.05).astype(np.float32)
    ref = A.astype(np.float64) @ B.astype(np.float64).T
    C = runner.debug_tc_gemm(A, B, three_pass=2)
    rowscale = np.sqrt((A.astype(np.float64) ** 2).sum(1, keepdims=True)) * 0.05
    err = np.abs(C - ref) / rowscale
    fp32 = np.abs((A @ B.T) - ref) / rowscale
    print(f"\nR={R} N={N} K={K}: f16x3 err {err.max():.2e}  numpy fp32 err {fp32.max():.2e}")
    assert err.max() < 1e-4


@pytest.mark.parametrize("model", ["ani2x_synth", "ani1xnr"])
def test_full_path_f16x3_parity(runner, ani2x_synth_path, ani1xnr_path, model):
    path = ani2x_synth_path if model == "ani2x_synth" else ani1xnr_path
    x, sp, box = domain.make_water_box((8, 9, 10), seed=7)
    sub = domain.decompose(x, sp, box)[0]
    il, nn, jl = domain.full_neighbor_list(sub.x, sub.nlocal)
    ref = OracleANI(path).compute_full(sub.species, sub.x, il, nn, jl, sub.nlocal, single_coords=True, want_aev=True, want_daev=True)
    ani = runner.ANI(path, 0, -1, True, True, True)
    ani.set_mlp_mode(runner.MLP_F16X3)
    e, f, eatom, virial = ani.compute(sub.species, sub.x, il, jl, nn, nlocal=sub.nlocal, ago=0, atomic=True, vflag=True)
    aev, daev = ani.debug_aev()
    print("f16x3: aev max err", np.abs(aev - ref["aev"]).max(), "daev max err", np.abs(daev - ref["daev"]).max(),
          "force max err", np.abs(f - ref["forces"]).max(), "E rel err", abs(e - ref["energy"]) / abs(ref["energy"]))
    assert np.abs(aev - ref["aev"]).max() <= 2e-5 * max(1.0, np.abs(ref["aev"]).max())
    assert abs(e - ref["energy"]) <= 1e-6 * abs(ref["energy"])
    assert np.abs(f - ref["forces"]).max() <= 1e-4 * EV2KCAL
    assert np.abs(eatom - ref["eatom"]).max() <= 1e-5 * np.abs(ref["eatom"]).max()
    assert np.abs(virial - ref["virial"]).max() <= 1e-4 * max(1.0, np.abs(ref["virial"]).max())
