"""Device neighbour-list build against the host list (scipy cKDTree = the LAMMPS full list of
pair_ani.cpp:107-151 as a set): integer work, so per-atom neighbour sets must be identical; then the step on the device-built
list must match the step on the host-built list."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
if not torch.cuda.is_available():
    pytest.skip("needs a CUDA device", allow_module_level=True)

from lammps_ani_b200 import domain, runner  # noqa: E402


def rows(nn, jl):
    off = np.concatenate([[0], np.cumsum(nn)])
    return [np.sort(jl[off[i]:off[i + 1]]) for i in range(len(nn))]


@pytest.mark.parametrize("cells,cut", [((5, 6, 7), 7.1), ((3, 3, 3), 7.1), ((6, 5, 4), 5.2), ((1, 1, 2), 7.1)])
def test_device_list_equals_host_list(ani2x_synth_path, cells, cut):
    x, sp, box = domain.make_water_box(cells, seed=31)
    sub = domain.decompose(x, sp, box)[0]                      # periodic images become ghosts (halo 7.1 A)
    il, nn, jl = domain.full_neighbor_list(sub.x, sub.nlocal, cut)
    ani = runner.ANI(ani2x_synth_path, 0, -1, True, True, True)
    P = ani.build_neighbors(sub.species, sub.x, sub.nlocal, cut)
    assert P == int(nn.sum()) == jl.size
    nn_d, jl_d = ani.debug_neighbors()
    assert np.array_equal(nn_d, nn)
    for a, b in zip(rows(nn_d, jl_d), rows(nn, jl)):
        assert np.array_equal(a, b)
    # a deliberately tight bounding box (atoms outside get clamped into the border cells) must give the same list
    lo, hi = sub.x.min(0) + 3.0, sub.x.max(0) - 3.0
    if np.all(hi > lo):
        assert ani.build_neighbors(sub.species, sub.x, sub.nlocal, cut, lo=lo, hi=hi) == P
        nn_c, jl_c = ani.debug_neighbors()
        assert np.array_equal(nn_c, nn)
        for a, b in zip(rows(nn_c, jl_c), rows(nn, jl)):
            assert np.array_equal(a, b)


def test_step_on_device_built_list(ani2x_synth_path):
    x, sp, box = domain.make_water_box((6, 6, 6), seed=5)
    sub = domain.decompose(x, sp, box)[0]
    il, nn, jl = domain.full_neighbor_list(sub.x, sub.nlocal)
    ani = runner.ANI(ani2x_synth_path, 0, -1, True, True, True)
    e0, f0, _, _ = ani.compute(sub.species, sub.x, il, jl, nn, nlocal=sub.nlocal, ago=0)
    xd = torch.tensor(sub.x, dtype=torch.float64, device="cuda:0")
    torch.cuda.synchronize()
    ani.build_neighbors(sub.species, None, sub.nlocal, 7.1, lo=sub.x.min(0) - 1e-6, hi=sub.x.max(0) + 1e-6, x_dev_ptr=xd.data_ptr())
    e1, f1, _, _ = ani.compute(None, sub.x, nlocal=sub.nlocal, ago=1)
    assert abs(e1 - e0) <= 1e-7 * abs(e0)
    assert np.abs(f1 - f0).max() <= 1e-4        # same sets, different order inside a row: fp32 summation order only


def test_rebuild_after_caller_array_grew_in_place(ani2x_synth_path):
    """What `pair_style ani ... device` does under LAMMPS: atom->x is page-locked in place by ani_compute once it is stable; at a
    later rebuild the same allocation holds MORE atoms (ghosts grew, realloc extended in place), so the host copy of the
    rebuild spans memory beyond the registered range.  The library must drop the stale page-lock first (it used to fail with
    `invalid argument`), and ani_ctx_release_host_buffers must be harmless at any time."""
    x, sp, box = domain.make_water_box((5, 5, 5), seed=77)
    sub = domain.decompose(x, sp, box)[0]
    nt = sub.x.shape[0]
    n1 = sub.nlocal + (nt - sub.nlocal) // 2                        # first: only half of the ghosts are there
    arena = np.zeros((nt + 64, 3))                                  # the caller's allocation, larger than its contents
    arena[:nt] = sub.x
    f = np.zeros((nt + 64, 3))
    ani = runner.ANI(ani2x_synth_path, 0, -1, True, True, True)
    ani.build_neighbors(sub.species[:n1], arena[:n1], sub.nlocal, 7.1)
    for _ in range(4):                                              # arena[:n1] becomes page-locked in place
        ani.compute(None, arena[:n1], nlocal=sub.nlocal, ago=1, out_force=f[:n1])
    P = ani.build_neighbors(sub.species, arena[:nt], sub.nlocal, 7.1)   # same pointer, more bytes
    e, fo, _, _ = ani.compute(None, arena[:nt], nlocal=sub.nlocal, ago=1, out_force=f[:nt])
    fresh = runner.ANI(ani2x_synth_path, 0, -1, True, True, True)
    assert fresh.build_neighbors(sub.species, sub.x, sub.nlocal, 7.1) == P
    e0, f0, _, _ = fresh.compute(None, sub.x, nlocal=sub.nlocal, ago=1)
    assert abs(e - e0) <= 1e-9 * abs(e0) and np.abs(fo - f0).max() <= 1e-4
    assert ani._L.ani_ctx_release_host_buffers(ani._ctx) == 0
    e2, f2, _, _ = ani.compute(None, arena[:nt], nlocal=sub.nlocal, ago=1, out_force=f[:nt])
    assert abs(e2 - e0) <= 1e-9 * abs(e0) and np.abs(f2 - f0).max() <= 1e-4
