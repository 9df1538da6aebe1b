"""The TorchScript `.pt` reader (lammps_ani_b200/ptreader.py): zip + restricted unpickler, no libtorch on the read side.
The reference's own exports (ani2x.pt, ani1x_nr.pt; models/ani_models.py:112-122) need torchani, which is not available
here, so the archives under test are written by torch.jit.save from mock modules that carry the ANI-1xnr parameters
(in tree, NeuroChem format) under the public torchani attribute names, in both network layouts."""
import numpy as np
import pytest
import torch

from lammps_ani_b200 import aniw, extract, ptreader


class _AEV(torch.nn.Module):
    def __init__(self, p):
        super().__init__()
        self.Rcr: float = float(p.Rcr)
        self.Rca: float = float(p.Rca)
        for name in ("ShfR", "ShfA", "ShfZ"):
            self.register_buffer(name, torch.tensor(getattr(p, name), dtype=torch.float32))
        for name in ("EtaR", "EtaA", "Zeta"):
            self.register_buffer(name, torch.tensor([getattr(p, name)], dtype=torch.float32))

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return x * self.Rcr + self.Rca + self.EtaR.sum()


class _Shifter(torch.nn.Module):
    def __init__(self, p):
        super().__init__()
        self.register_buffer("self_energies", torch.tensor(p.sae, dtype=torch.float64))

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return x + self.self_energies.sum()


def _net(layers):
    mods = []
    for k, (W, b) in enumerate(layers):
        lin = torch.nn.Linear(W.shape[1], W.shape[0])
        with torch.no_grad():
            lin.weight.copy_(torch.tensor(W)); lin.bias.copy_(torch.tensor(b))
        mods.append(lin)
        if k + 1 < len(layers):
            mods.append(torch.nn.CELU(0.1))
    return torch.nn.Sequential(*mods)


class _Stacked(torch.nn.Module):
    """BmmEnsemble-like: one module per species with [M, in, out] weights and [M, 1, out] biases."""
    def __init__(self, p, s):
        super().__init__()
        self.layers = torch.nn.ModuleList()
        for l in range(len(p.dims[s]) - 1):
            m = torch.nn.Module()
            m.weight = torch.nn.Parameter(torch.tensor(np.stack([p.weights[k][s][l][0].T for k in range(p.num_models)])))
            m.bias = torch.nn.Parameter(torch.tensor(np.stack([p.weights[k][s][l][1][None] for k in range(p.num_models)])))
            self.layers.append(m)


class _Rep(torch.nn.Module):
    """Stand-in for RepulsionXTB: the dummy RepulsionXTB(("H")) every exported LammpsANI carries (models/lammps_ani.py:91-99),
    or real tables when the term is on."""
    def __init__(self, tabs=None):
        super().__init__()
        for name in ("y_ab", "sqrt_alpha_ab", "k_rep_ab"):
            self.register_buffer(name, torch.ones(1, 1) if tabs is None else torch.tensor(tabs[name]))
        self.cutoff: float = 5.2


class _Top(torch.nn.Module):
    def __init__(self, p, stacked, use_repulsion=False):
        super().__init__()
        self.aev_computer = _AEV(p)
        self.energy_shifter = _Shifter(p)
        self.use_repulsion: bool = use_repulsion
        self.rep_calc = _Rep(extract.xtb_repulsion_tables(p.symbols) if use_repulsion else None)
        if stacked == "dict":      # BmmEnsemble container keyed by element symbol: archive order is NOT alphabetical order
            self.neural_networks = torch.nn.ModuleDict({sym: _Stacked(p, s) for s, sym in enumerate(p.symbols)})
        elif stacked:
            self.neural_networks = torch.nn.ModuleList([_Stacked(p, s) for s in range(p.num_species)])
        else:
            self.neural_networks = torch.nn.ModuleList(
                [torch.nn.ModuleDict({sym: _net(p.weights[m][s]) for s, sym in enumerate(p.symbols)}) for m in range(p.num_models)])

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return self.energy_shifter(self.aev_computer(x))


@pytest.mark.parametrize("stacked", [False, True, "dict"])
def test_pt_archive_round_trip(ani1xnr_path, tmp_path, stacked):
    p = aniw.load(ani1xnr_path)
    pt = str(tmp_path / "mock.pt")
    torch.jit.save(torch.jit.script(_Top(p, stacked)), pt)
    q = ptreader.read_torchscript_params(pt)
    assert q.symbols == p.symbols and q.dims == p.dims and q.num_models == p.num_models
    assert (q.Rcr, q.Rca) == (p.Rcr, p.Rca)
    assert not q.use_repulsion          # the dummy rep_calc submodule does not switch the term on
    for name in ("ShfR", "ShfA", "ShfZ"):
        assert np.allclose(getattr(q, name), getattr(p, name), rtol=1e-6)
    assert np.allclose(q.sae, p.sae)
    for m in range(p.num_models):
        for s in range(p.num_species):
            for l in range(len(p.dims[s]) - 1):
                assert np.array_equal(q.weights[m][s][l][0], p.weights[m][s][l][0])
                assert np.array_equal(q.weights[m][s][l][1], p.weights[m][s][l][1])
    # and through the extractor's CLI path to an .aniw blob the C++ loader accepts
    out = str(tmp_path / "mock.pt.aniw")
    extract.extract(pt, out)
    r = aniw.load(out)
    assert r.dims == p.dims and np.array_equal(r.weights[3][1][2][0], p.weights[3][1][2][0])


def test_archive_code_is_never_executed(tmp_path):
    """The unpickler refuses anything that is not a tensor rebuild or a __torch__ attribute bag."""
    import io, pickle, zipfile
    evil = pickle.dumps(eval)            # a GLOBAL builtins.eval opcode
    z = str(tmp_path / "evil.pt")
    with zipfile.ZipFile(z, "w") as zf:
        zf.writestr("evil/data.pkl", evil)
    with pytest.raises(pickle.UnpicklingError):
        ptreader.load_object_graph(z)


def test_use_repulsion_is_read_from_the_attribute_value(ani1xnr_path, tmp_path):
    p = aniw.load(ani1xnr_path)
    pt = str(tmp_path / "rep.pt")
    torch.jit.save(torch.jit.script(_Top(p, True, use_repulsion=True)), pt)
    q = ptreader.read_torchscript_params(pt)
    assert q.use_repulsion and q.repulsion["cutoff"] == 5.2
    assert np.allclose(q.repulsion["y_ab"], extract.xtb_repulsion_tables(p.symbols)["y_ab"])
    out = str(tmp_path / "rep.aniw")
    aniw.save(q, out)
    r = aniw.load(out)
    assert r.use_repulsion and np.array_equal(r.repulsion["k_rep_ab"], q.repulsion["k_rep_ab"])
