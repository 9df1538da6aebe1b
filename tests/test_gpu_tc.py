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


@pytest.mark.parametrize("fused", [1, 0])
@pytest.mark.parametrize("model", ["ani2x_synth", "ani1xnr"])
def test_full_path_f16x3_parity(runner, ani2x_synth_path, ani1xnr_path, model, fused):
    """fused 1: the whole network of a bucket in one kernel (mlp_f16_fused.cuh, the default); 0: layer by layer (mlp_f16.cuh)."""
    path = ani2x_synth_path if model == "ani2x_synth" else ani1xnr_path
    x, sp, box = domain.make_water_box((8, 9, 10), seed=7)
    sub = domain.decompose(x, sp, box)[0]
    il, nn, jl = domain.full_neighbor_list(sub.x, sub.nlocal)
    ref = OracleANI(path).compute_full(sub.species, sub.x, il, nn, jl, sub.nlocal, single_coords=True, want_aev=True, want_daev=True)
    ani = runner.ANI(path, 0, -1, True, True, True)
    ani.set_mlp_mode(runner.MLP_F16X3)
    ani.set_option("mlp_fused", fused)
    e, f, eatom, virial = ani.compute(sub.species, sub.x, il, jl, nn, nlocal=sub.nlocal, ago=0, atomic=True, vflag=True)
    aev, daev = ani.debug_aev()
    print(f"f16x3 fused={fused}: aev max err", np.abs(aev - ref["aev"]).max(), "daev max err", np.abs(daev - ref["daev"]).max(),
          "force max err", np.abs(f - ref["forces"]).max(), "E rel err", abs(e - ref["energy"]) / abs(ref["energy"]))
    assert np.abs(aev - ref["aev"]).max() <= 2e-5 * max(1.0, np.abs(ref["aev"]).max())
    assert abs(e - ref["energy"]) <= 1e-6 * abs(ref["energy"])
    assert np.abs(f - ref["forces"]).max() <= 1e-4 * EV2KCAL
    assert np.abs(eatom - ref["eatom"]).max() <= 1e-5 * np.abs(ref["eatom"]).max()
    assert np.abs(virial - ref["virial"]).max() <= 1e-4 * max(1.0, np.abs(ref["virial"]).max())


def _mixed_box(cells, seed, species_pool, nswap):
    rng = np.random.default_rng(seed)
    x, sp, box = domain.make_water_box(cells, seed=seed)
    sp = sp.copy()
    pick = rng.choice(len(sp), size=nswap, replace=False)
    sp[pick] = rng.choice(species_pool, size=nswap)
    sub = domain.decompose(x, sp, box)[0]
    il, nn, jl = domain.full_neighbor_list(sub.x, sub.nlocal)
    return sub, il, nn, jl


@pytest.mark.parametrize("fused", [1, 0])
def test_f16x3_mixed_species_wide_aev(runner, ani2x_synth_path, fused):
    """H, C, N, O present: 384 active AEV columns (> 256), so the fused kernel hands g1 to a separate dE/dAEV GEMM; tiny
    C / N buckets (a single ragged row tile)."""
    sub, il, nn, jl = _mixed_box((6, 6, 5), 21, [1, 2], 40)
    ref = OracleANI(ani2x_synth_path).compute_full(sub.species, sub.x, il, nn, jl, sub.nlocal, single_coords=True, want_daev=True)
    ani = runner.ANI(ani2x_synth_path, 0, -1, True, True, True)
    ani.set_option("mlp_fused", fused)
    e, f, eatom, virial = ani.compute(sub.species, sub.x, il, jl, nn, nlocal=sub.nlocal, ago=0, atomic=True, vflag=True)
    _, daev = ani.debug_aev()
    print(f"mixed fused={fused}: daev max err", np.abs(daev - ref["daev"]).max(), "force max err", np.abs(f - ref["forces"]).max(),
          "E rel err", abs(e - ref["energy"]) / abs(ref["energy"]))
    assert abs(e - ref["energy"]) <= 1e-6 * abs(ref["energy"])
    assert np.abs(f - ref["forces"]).max() <= 1e-4 * EV2KCAL
    assert np.abs(eatom - ref["eatom"]).max() <= 1e-5 * np.abs(ref["eatom"]).max()
    assert np.abs(virial - ref["virial"]).max() <= 1e-4 * max(1.0, np.abs(ref["virial"]).max())


def test_fused_matches_layerwise_many_items_per_cta(runner, ani2x_synth_path):
    """12 288 atoms: 512 + 256 work items over 148 CTAs, so every CTA pipelines several items back to back (ring phases wrap,
    TMEM regions are reused across items); repeated steps must reproduce, and the fused kernel must agree with the
    layer-by-layer kernels to fp32 noise."""
    x, sp, box = domain.make_water_box((16, 16, 16), seed=4)
    sub = domain.decompose(x, sp, box)[0]
    il, nn, jl = domain.full_neighbor_list(sub.x, sub.nlocal)
    out = {}
    for fused in (0, 1):
        ani = runner.ANI(ani2x_synth_path, 0, -1, True, True, True)
        ani.set_option("mlp_fused", fused)
        e, f, eatom, _ = ani.compute(sub.species, sub.x, il, jl, nn, nlocal=sub.nlocal, ago=0, atomic=True)
        for _ in range(3):
            e2, f2, eatom2, _ = ani.compute(sub.species, sub.x, nlocal=sub.nlocal, ago=1, atomic=True)
            assert abs(e2 - e) <= 1e-9 * abs(e) and np.abs(f2 - f).max() <= 1e-3   # fp32 atomics: summation order only
        out[fused] = (e, f, eatom)
    (e0, f0, a0), (e1, f1, a1) = out[0], out[1]
    print("fused vs layerwise: E rel", abs(e1 - e0) / abs(e0), "F max", np.abs(f1 - f0).max(), "eatom max", np.abs(a1 - a0).max())
    assert abs(e1 - e0) <= 2e-7 * abs(e0)
    assert np.abs(f1 - f0).max() <= 0.5e-4 * EV2KCAL
    assert np.abs(a1 - a0).max() <= 2e-6 * np.abs(a0).max()


@pytest.mark.parametrize("fused", [1])
def test_f16_operand_range_guard(runner, tmp_path, fused):
    """Weights x64 on a compressed box: activations leave the fp16 range of the f16x3 planes (|16 h| >= 65504).  The library
    must say so (status 3 -> RuntimeError) instead of returning inf / nan forces; a tf32x3 context evaluates the same
    system."""
    from lammps_ani_b200 import aniw, extract
    p = str(tmp_path / "big.aniw")
    big = extract.synthetic_ani2x(seed=7, weight_scale=64.0)
    aniw.save(big, p)
    x, sp, box = domain.make_water_box((5, 5, 5), seed=3)
    x = x * 0.6
    box = np.asarray(box) * 0.6
    sub = domain.decompose(x, sp, box)[0]
    il, nn, jl = domain.full_neighbor_list(sub.x, sub.nlocal)
    ani = runner.ANI(p, 0, -1, True, True, True)
    ani.set_option("mlp_fused", fused)
    with pytest.raises(RuntimeError, match="fp16 operand range"):
        ani.compute(sub.species, sub.x, il, jl, nn, nlocal=sub.nlocal, ago=0)
    # The strict fp32 MLP evaluates the same system, but with weights x64 in three layers the angular force terms themselves can
    # exceed the fixed-point range of the fp32 AEV backward kernel: if so that is reported too (same status, its own message) ...
    ani2 = runner.ANI(p, 0, -1, True, True, True)
    ani2.set_mlp_mode(runner.MLP_FP32_SIMT)
    try:
        e, f, _, _ = ani2.compute(sub.species, sub.x, il, jl, nn, nlocal=sub.nlocal, ago=0)
        assert np.isfinite(e) and np.isfinite(f).all()
    except RuntimeError as err:
        assert "fixed-point range" in str(err)
    # ... and the double-precision context (no fp16 planes, no fixed point) handles it
    ani3 = runner.ANI(p, 0, -1, True, True, False)
    e, f, _, _ = ani3.compute(sub.species, sub.x, il, jl, nn, nlocal=sub.nlocal, ago=0)
    assert np.isfinite(e) and np.isfinite(f).all()


def test_fixed_point_range_guard(runner, ani2x_synth_path):
    """the fp32 AEV backward kernel sums angular gradient terms in 2^-36 fixed point, which holds |term| < ~5000.
    Two atoms 1e-6 A apart give terms of order dE/dcos / r far beyond that: the library must report status 3 instead of
    returning wrapped-around forces; the next (sane) step on the same context is clean."""
    x, sp, box = domain.make_water_box((5, 5, 5), seed=3)
    sub = domain.decompose(x, sp, box)[0]
    good = sub.x.copy()
    bad = good - (good[0] - 0.25)                 # atom 0 at (0.25, 0.25, 0.25): 2^-20 A is exact in fp32 there
    bad[1] = bad[0] + np.array([2.0 ** -20, 0.0, 0.0])
    ani = runner.ANI(ani2x_synth_path, 0, -1, True, True, True)
    ani.set_mlp_mode(runner.MLP_FP32_SIMT)        # no fp16 planes involved: only the fixed-point guard can fire
    il, nn, jl = domain.full_neighbor_list(bad, sub.nlocal)
    with pytest.raises(RuntimeError, match="fixed-point range"):
        ani.compute(sub.species, bad, il, jl, nn, nlocal=sub.nlocal, ago=0)
    il, nn, jl = domain.full_neighbor_list(good, sub.nlocal)
    e, f, _, _ = ani.compute(sub.species, good, il, jl, nn, nlocal=sub.nlocal, ago=0)
    assert np.isfinite(e) and np.isfinite(f).all()
