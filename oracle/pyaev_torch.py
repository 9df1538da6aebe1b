"""Second, independent CPU restatement: the pyaev pair/triple formulation in torch fp64 + autograd.

TEST INFRASTRUCTURE ONLY (same rules as oracle/ani_oracle.c).  Where ani_oracle.c is a per-centre
loop with hand-derived gradients, this file mirrors what the reference actually executes on its
`pyaev` path: a HALF pair list (`atom_index12`, `diff_vector`, `distances` built exactly like
/root/reference/src/ani_csrc/ani.cpp:143-144), radial terms scattered to both ends of every pair,
angular terms over "triples" (a centre with two of its <= Rca pairs), per-species networks with the
ensemble mean, the energy shifter over LOCAL atoms, and then
`torch.autograd.grad(E, [coordinates, diff_vector])` with
`virial = -sym(dEdR^T @ diff_vector)` (/root/reference/models/lammps_ani.py:195-215).
The two restatements share no code, so their agreement (tests/test_oracle.py) pins the hand-written
derivatives; the torchani source itself is not in the tree.
"""
from __future__ import annotations

import math

import numpy as np
import torch

HARTREE2KCALMOL = 627.5094738898777


def _triples(atom_index12: np.ndarray, natoms: int):
    """All (centre, pair_a, pair_b) with pair_a < pair_b both incident to `centre`; returns centre index,
    the two pair indices and the signs that orient each pair's diff_vector away from the centre."""
    P = atom_index12.shape[1]
    centre = np.concatenate([atom_index12[0], atom_index12[1]])
    pair = np.concatenate([np.arange(P), np.arange(P)])
    # diff = x[idx0] - x[idx1]; vector centre->other is -diff when centre==idx0, +diff when centre==idx1
    sign = np.concatenate([-np.ones(P), np.ones(P)])
    order = np.argsort(centre, kind="stable")
    centre, pair, sign = centre[order], pair[order], sign[order]
    counts = np.bincount(centre, minlength=natoms)
    starts = np.concatenate([[0], np.cumsum(counts)])
    c_out, a_out, b_out, sa_out, sb_out = [], [], [], [], []
    for i in np.nonzero(counts >= 2)[0]:
        s, n = starts[i], counts[i]
        ia, ib = np.triu_indices(n, k=1)
        c_out.append(np.full(ia.size, i)); a_out.append(pair[s + ia]); b_out.append(pair[s + ib])
        sa_out.append(sign[s + ia]); sb_out.append(sign[s + ib])
    if not c_out:
        z = np.zeros(0, dtype=np.int64)
        return z, z, z, np.zeros(0), np.zeros(0)
    return (np.concatenate(c_out), np.concatenate(a_out), np.concatenate(b_out),
            np.concatenate(sa_out), np.concatenate(sb_out))


def pyaev_forward(params, species, coordinates, atom_index12, nlocal, *, num_models=None, radial_filter=True,
                  want_aev=False):
    """species [T] (real species for ghosts), coordinates [T,3] f64, atom_index12 int64[2,P] (half list).
    Returns dict(energy, forces [T,3], eatom [nlocal], virial [3,3]) in kcal/mol units."""
    dt = torch.float64
    S = params.num_species
    nR, nA, nZ = len(params.ShfR), len(params.ShfA), len(params.ShfZ)
    M = params.num_models if num_models in (None, -1) else num_models
    T = len(species)
    sp = torch.as_tensor(np.asarray(species), dtype=torch.long)
    x = torch.tensor(np.asarray(coordinates, dtype=np.float64), dtype=dt, requires_grad=True)
    idx = torch.as_tensor(np.asarray(atom_index12, dtype=np.int64))
    diff = x.index_select(0, idx[0]) - x.index_select(0, idx[1])          # ani.cpp:143
    dist = diff.norm(2, -1)                                               # ani.cpp:144
    ShfR = torch.tensor(params.ShfR, dtype=dt); ShfA = torch.tensor(params.ShfA, dtype=dt)
    ShfZ = torch.tensor(params.ShfZ, dtype=dt)

    # ---- radial: one term vector per pair, scattered to both atoms under the OTHER atom's species
    if radial_filter:
        keep = (dist <= params.Rcr).nonzero().flatten()
    else:
        keep = torch.arange(dist.numel())
    d_r = dist.index_select(0, keep); i_r = idx.index_select(1, keep)
    fc = 0.5 * torch.cos(d_r * (math.pi / params.Rcr)) + 0.5
    rad = 0.25 * torch.exp(-params.EtaR * (d_r.unsqueeze(-1) - ShfR) ** 2) * fc.unsqueeze(-1)     # [P,nR]
    radial = torch.zeros(T * S, nR, dtype=dt)
    radial = radial.index_add(0, i_r[0] * S + sp[i_r[1]], rad)
    radial = radial.index_add(0, i_r[1] * S + sp[i_r[0]], rad)
    radial = radial.reshape(T, S * nR)

    # ---- angular: pairs within Rca, all triples around each centre
    close = (dist <= params.Rca).nonzero().flatten()
    i_a = idx.index_select(1, close)
    v_a = diff.index_select(0, close)
    c, pa, pb, sa, sb = _triples(i_a.numpy(), T)
    c_t = torch.as_tensor(c); pa_t = torch.as_tensor(pa); pb_t = torch.as_tensor(pb)
    va = v_a.index_select(0, pa_t) * torch.as_tensor(sa, dtype=dt).unsqueeze(-1)
    vb = v_a.index_select(0, pb_t) * torch.as_tensor(sb, dtype=dt).unsqueeze(-1)
    # species of the non-centre atom of each pair
    other_a = torch.where(torch.as_tensor(sa) < 0, i_a[1][pa_t], i_a[0][pa_t])
    other_b = torch.where(torch.as_tensor(sb) < 0, i_a[1][pb_t], i_a[0][pb_t])
    ra, rb = va.norm(2, -1), vb.norm(2, -1)
    cosang = (va * vb).sum(-1) / torch.clamp(ra * rb, min=1e-10)
    ang = torch.acos(0.95 * cosang)
    fcab = (0.5 * torch.cos(ra * (math.pi / params.Rca)) + 0.5) * (0.5 * torch.cos(rb * (math.pi / params.Rca)) + 0.5)
    f1 = ((1 + torch.cos(ang.view(-1, 1, 1) - ShfZ.view(1, 1, -1))) / 2) ** params.Zeta             # [t,1,nZ]
    f2 = torch.exp(-params.EtaA * (((ra + rb) / 2).view(-1, 1, 1) - ShfA.view(1, -1, 1)) ** 2)       # [t,nA,1]
    terms = (2 * f1 * f2 * fcab.view(-1, 1, 1)).reshape(-1, nA * nZ)
    triu = torch.zeros(S, S, dtype=torch.long)
    iu = torch.triu_indices(S, S)
    triu[iu[0], iu[1]] = torch.arange(iu.shape[1]); triu[iu[1], iu[0]] = torch.arange(iu.shape[1])
    npairs = S * (S + 1) // 2
    angular = torch.zeros(T * npairs, nA * nZ, dtype=dt)
    angular = angular.index_add(0, c_t * npairs + triu[sp[other_a], sp[other_b]], terms)
    aev = torch.cat([radial, angular.reshape(T, npairs * nA * nZ)], -1)

    # ---- networks on LOCAL atoms only (ghosts are padding -1: ani.cpp:226-228)
    eatom = torch.zeros(nlocal, dtype=dt)
    for s in range(S):
        rows = (sp[:nlocal] == s).nonzero().flatten()
        if rows.numel() == 0:
            continue
        acc = torch.zeros(rows.numel(), dtype=dt)
        for m in range(M):
            h = aev.index_select(0, rows)
            net = params.weights[m][s]
            for l, (W, b) in enumerate(net):
                h = torch.nn.functional.linear(h, torch.tensor(W, dtype=dt), torch.tensor(b, dtype=dt))
                if l < len(net) - 1:
                    h = torch.nn.functional.celu(h, alpha=0.1)
            acc = acc + h.squeeze(-1)
        eatom = eatom.index_add(0, rows, acc / M + params.sae[s])
    energy = eatom.sum()
    dEdx, dEdR = torch.autograd.grad([energy], [x, diff])
    virial = dEdR.t() @ diff.detach()
    virial = -(virial.t() + virial) / 2
    k = HARTREE2KCALMOL
    out = dict(energy=float(energy.detach()) * k, forces=(-dEdx).numpy() * k, eatom=eatom.detach().numpy() * k,
               virial=virial.numpy() * k)
    if want_aev:
        out["aev"] = aev.detach().numpy()[:nlocal]
    return out
