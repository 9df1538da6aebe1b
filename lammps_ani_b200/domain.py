"""Host-side stand-in for the LAMMPS machinery on either side of the pair style:
synthetic water boxes, periodic ghost atoms, full/half neighbour lists at cutoff+skin, and the brick
domain decomposition with its halo send/receive plan.  Used by bench.py and the tests to feed the
C ABI exactly what `PairANI::compute` hands to `ANI::compute`
(/root/reference/src/pair_ani.cpp:98-151): x[ntotal][3], type-1 species, ilist/numneigh/jlist.

Conventions copied from LAMMPS: local atoms first (0..nlocal-1), ghosts after; a ghost is a copy of an
owned atom displaced by a periodic image and/or owned by another rank; full list = every j within the
list cutoff of local i (j local or ghost); half list with newton off = local-local pairs once, plus every
local-ghost pair stored on the local atom (external/lammps/src/npair_half_bin_newtoff.cpp behaviour).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Optional, Tuple

import numpy as np

WATER_MOLECULES_PER_A3 = 0.0334   # ~0.998 g/cc; the reference's 20k box: 7128 molecules / 60^3 A^3 = 0.0330
LIST_CUTOFF = 7.1                 # pair cutoff 5.1 + skin 2.0 (/root/reference/tests/in.lammps:19)


def make_water_box(cells: Tuple[int, int, int], seed: int = 12345, jitter: float = 0.25):
    """Waters on a jittered simple-cubic lattice with uniformly random orientations
    (the fallback generator of SURVEY.md 8d; geometry r_OH 0.9572 A, HOH 104.52 deg).
    Returns (positions [N,3] f64, species [N] int32 with ANI-2x indices H=0,O=3, box [3])."""
    rng = np.random.default_rng(seed)
    a = (1.0 / WATER_MOLECULES_PER_A3) ** (1.0 / 3.0)
    nx, ny, nz = cells
    gx, gy, gz = np.meshgrid(np.arange(nx), np.arange(ny), np.arange(nz), indexing="ij")
    o = (np.stack([gx, gy, gz], -1).reshape(-1, 3) + 0.5) * a
    nmol = o.shape[0]
    o = o + rng.uniform(-jitter, jitter, size=o.shape)
    # random rotations from unit quaternions
    q = rng.standard_normal((nmol, 4))
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    w, x, y, z = q.T
    R = np.stack([
        np.stack([1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)], -1),
        np.stack([2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)], -1),
        np.stack([2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)], -1)], 1)
    r_oh, ang = 0.9572, np.deg2rad(104.52)
    h1 = np.array([r_oh * np.sin(ang / 2), 0.0, r_oh * np.cos(ang / 2)])
    h2 = np.array([-r_oh * np.sin(ang / 2), 0.0, r_oh * np.cos(ang / 2)])
    H1 = o + R @ h1
    H2 = o + R @ h2
    pos = np.stack([o, H1, H2], 1).reshape(-1, 3)
    species = np.tile(np.array([3, 0, 0], dtype=np.int32), nmol)
    box = np.array([nx, ny, nz], dtype=np.float64) * a
    pos = np.mod(pos, box)
    return pos, species, box


def _random_rotations(rng, n):
    q = rng.standard_normal((n, 4))
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    w, x, y, z = q.T
    return np.stack([
        np.stack([1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)], -1),
        np.stack([2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)], -1),
        np.stack([2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)], -1)], 1)


def make_combustion_box(n_ch4: int = 33334, seed: int = 12345, density_g_cc: float = 0.25):
    """BASELINE.json config 5: the reference's combustion system -- n CH4 + 2n O2 at 0.25 g/cc (33 334 + 66 668 molecules =
    300 006 atoms in a 260.715 A box: examples/combustion/prepare_system/system-300k.inp; its .data file is a missing blob).
    packmol's random packing (tolerance 2.0 A) is replaced by a seeded jittered lattice: molecules on a random subset of the sites
    of a simple-cubic lattice, random orientations, +-0.5 A jitter, so that atoms of different molecules stay >= 2 A apart.
    Returns (positions [N,3], species [N] int32 in ANI-1x order H=0, C=1, N=2, O=3, box [3])."""
    rng = np.random.default_rng(seed)
    nmol = 3 * n_ch4
    mass = n_ch4 * (12.0107 + 4 * 1.008) + 2 * n_ch4 * 2 * 15.999
    L = (mass / (density_g_cc * 0.602214076)) ** (1.0 / 3.0)
    n = int(np.ceil(nmol ** (1.0 / 3.0)))
    a = L / n
    if a < 5.3:
        raise ValueError("density too high for the lattice placement (sites closer than 5.3 A)")
    sites = rng.permutation(n ** 3)[:nmol]
    c = (np.stack(np.unravel_index(sites, (n, n, n)), -1) + 0.5) * a + rng.uniform(-0.5, 0.5, size=(nmol, 3))
    is_ch4 = np.zeros(nmol, dtype=bool)
    is_ch4[rng.permutation(nmol)[:n_ch4]] = True
    R = _random_rotations(rng, nmol)
    t = 1.09 / np.sqrt(3.0)
    ch4 = np.array([[0, 0, 0], [t, t, t], [t, -t, -t], [-t, t, -t], [-t, -t, t]], dtype=np.float64)
    o2 = np.array([[0, 0, 0.604], [0, 0, -0.604]], dtype=np.float64)
    pos, spc = [], []
    cm = c[is_ch4]; Rm = R[is_ch4]
    pos.append((cm[:, None, :] + np.einsum("nij,aj->nai", Rm, ch4)).reshape(-1, 3))
    spc.append(np.tile(np.array([1, 0, 0, 0, 0], dtype=np.int32), cm.shape[0]))
    co = c[~is_ch4]; Ro = R[~is_ch4]
    pos.append((co[:, None, :] + np.einsum("nij,aj->nai", Ro, o2)).reshape(-1, 3))
    spc.append(np.tile(np.array([3, 3], dtype=np.int32), co.shape[0]))
    box = np.array([L, L, L])
    return np.mod(np.concatenate(pos), box), np.concatenate(spc), box


def cells_for_atoms(natoms: int) -> Tuple[int, int, int]:
    """Near-cubic lattice whose atom count is close to `natoms` (3 atoms per cell)."""
    nmol = max(1, natoms // 3)
    n = max(1, round(nmol ** (1.0 / 3.0)))
    best = None
    for nx in range(max(1, n - 3), n + 4):
        for ny in range(max(1, n - 3), n + 4):
            nz = max(1, round(nmol / (nx * ny)))
            err = abs(nx * ny * nz - nmol) + 0.01 * (abs(nx - ny) + abs(ny - nz))
            if best is None or err < best[0]:
                best = (err, (nx, ny, nz))
    return best[1]


@dataclass
class Subdomain:
    """What one MPI rank of LAMMPS holds for the pair style."""
    nlocal: int
    x: np.ndarray                 # [ntotal,3] f64, locals then ghosts
    species: np.ndarray           # [ntotal] int32 (type-1)
    ghost_owner_rank: np.ndarray  # [nghost] rank owning the real atom
    ghost_owner_index: np.ndarray  # [nghost] local index on the owner
    ghost_shift: np.ndarray       # [nghost,3] periodic image shift (in box lengths) that produced the ghost
    global_id: np.ndarray         # [ntotal] id in the global box
    lo: np.ndarray
    hi: np.ndarray

    @property
    def ntotal(self) -> int:
        return self.x.shape[0]

    @property
    def nghost(self) -> int:
        return self.ntotal - self.nlocal


def _ghost_images(pos, box, lo, hi, cut, periodic=True):
    """All images (including shift 0 for atoms outside [lo,hi)) of `pos` inside the halo shell of the brick
    [lo-cut, hi+cut) minus the brick itself.  Returns (index into pos, shifted coordinates)."""
    idx_all, x_all, s_all = [], [], []
    rng = [-1, 0, 1] if periodic else [0]
    # enough images when box >= cut; for tiny boxes (8 A box, 7.1 A halo) allow +-ceil(cut/box) images
    reach = [int(np.ceil(cut / b)) if periodic else 0 for b in box]
    for sx in range(-reach[0], reach[0] + 1):
        for sy in range(-reach[1], reach[1] + 1):
            for sz in range(-reach[2], reach[2] + 1):
                shift = np.array([sx, sy, sz]) * box
                y = pos + shift
                inside_halo = np.all((y >= lo - cut) & (y < hi + cut), axis=1)
                inside_brick = np.all((y >= lo) & (y < hi), axis=1)
                sel = np.nonzero(inside_halo & ~inside_brick)[0]
                if sel.size:
                    idx_all.append(sel)
                    x_all.append(y[sel])
                    s_all.append(np.tile(np.array([sx, sy, sz], dtype=np.int64), (sel.size, 1)))
    if not idx_all:
        return np.zeros(0, dtype=np.int64), np.zeros((0, 3)), np.zeros((0, 3), dtype=np.int64)
    return np.concatenate(idx_all), np.concatenate(x_all), np.concatenate(s_all)


def decompose(pos, species, box, grid=(1, 1, 1), cut: float = LIST_CUTOFF, periodic: bool = True) -> List[Subdomain]:
    """Brick decomposition of a periodic box onto px*py*pz ranks (rank = ix*py*pz + iy*pz + iz) with the
    cut-wide halo LAMMPS' Comm::borders would create (comm_brick.cpp:771-)."""
    pos = np.asarray(pos, dtype=np.float64)
    grid = np.asarray(grid)
    cell = box / grid
    owner_cell = np.minimum((pos / cell).astype(np.int64), grid - 1)
    owner_rank = (owner_cell[:, 0] * grid[1] + owner_cell[:, 1]) * grid[2] + owner_cell[:, 2]
    nranks = int(np.prod(grid))
    local_ids = [np.nonzero(owner_rank == r)[0] for r in range(nranks)]
    local_index = np.empty(pos.shape[0], dtype=np.int64)
    for ids in local_ids:
        local_index[ids] = np.arange(ids.size)
    out = []
    for r in range(nranks):
        ix, rem = divmod(r, grid[1] * grid[2])
        iy, iz = divmod(rem, grid[2])
        lo = np.array([ix, iy, iz]) * cell
        hi = lo + cell
        ids = local_ids[r]
        gidx, gx, gshift = _ghost_images(pos, box, lo, hi, cut, periodic)
        x = np.concatenate([pos[ids], gx])
        sp = np.concatenate([species[ids], species[gidx]]).astype(np.int32)
        out.append(Subdomain(nlocal=ids.size, x=x, species=sp, ghost_owner_rank=owner_rank[gidx],
                             ghost_owner_index=local_index[gidx], ghost_shift=gshift,
                             global_id=np.concatenate([ids, gidx]),
                             lo=lo, hi=hi))
    return out


def full_neighbor_list(x, nlocal: int, cut: float = LIST_CUTOFF):
    """LAMMPS-style full list for local atoms over local+ghost coordinates.
    Returns (ilist [nlocal] i32, numneigh [nlocal] i32, jlist [P] i32 flattened in ilist order)."""
    from scipy.spatial import cKDTree
    x = np.asarray(x, dtype=np.float64)
    tree = cKDTree(x)
    pairs = tree.query_pairs(cut, output_type="ndarray")  # i<j, all atoms
    a = np.concatenate([pairs[:, 0], pairs[:, 1]])
    b = np.concatenate([pairs[:, 1], pairs[:, 0]])
    keep = a < nlocal
    a, b = a[keep], b[keep]
    order = np.lexsort((b, a))
    a, b = a[order], b[order]
    numneigh = np.bincount(a, minlength=nlocal).astype(np.int32)
    return np.arange(nlocal, dtype=np.int32), numneigh, b.astype(np.int32)


def half_neighbor_list(x, nlocal: int, cut: float = LIST_CUTOFF):
    """Half list with newton off: int64[2][P] as `atom_index12` (pair_ani.cpp:143-146): local-local pairs
    once (i<j), local-ghost pairs stored with the local atom first."""
    from scipy.spatial import cKDTree
    x = np.asarray(x, dtype=np.float64)
    pairs = cKDTree(x).query_pairs(cut, output_type="ndarray")
    keep = pairs[:, 0] < nlocal            # i<j and locals come first => i local covers all needed pairs
    p = pairs[keep]
    order = np.lexsort((p[:, 1], p[:, 0]))
    p = p[order]
    return np.ascontiguousarray(p.T.astype(np.int64))


def fold_ghost_forces(sub: Subdomain, forces: np.ndarray) -> np.ndarray:
    """Single-rank reverse communication (pair_ani.cpp:197-201,461-484): add each ghost's force onto its owner."""
    f = np.array(forces[:sub.nlocal], dtype=np.float64, copy=True)
    np.add.at(f, sub.ghost_owner_index, forces[sub.nlocal:])
    return f


@dataclass
class HaloPlan:
    """Per-rank send/recv description for the forward (positions) and reverse (forces) exchanges.
    send_index[r]: local atom indices whose coordinates (plus send_shift[r]) rank r needs as ghosts;
    recv_slice[r]: where, inside this rank's ghost block, the atoms received from rank r land."""
    send_index: List[np.ndarray]
    send_shift: List[np.ndarray]
    recv_start: List[int]
    recv_count: List[int]


def halo_plans(subs: List[Subdomain], pos_global: np.ndarray) -> List[HaloPlan]:
    """Build all ranks' plans at once (host setup; in LAMMPS this is Comm::borders)."""
    n = len(subs)
    # order every rank's ghosts by owner rank so each peer's block is contiguous
    for s in subs:
        order = np.argsort(s.ghost_owner_rank, kind="stable")
        nl = s.nlocal
        s.x[nl:] = s.x[nl:][order]
        s.species[nl:] = s.species[nl:][order]
        s.global_id[nl:] = s.global_id[nl:][order]
        s.ghost_owner_rank = s.ghost_owner_rank[order]
        s.ghost_owner_index = s.ghost_owner_index[order]
        s.ghost_shift = s.ghost_shift[order]
    plans = [HaloPlan([None] * n, [None] * n, [0] * n, [0] * n) for _ in range(n)]
    for r, s in enumerate(subs):
        counts = np.bincount(s.ghost_owner_rank, minlength=n)
        starts = np.concatenate([[0], np.cumsum(counts)])
        for o in range(n):
            plans[r].recv_start[o] = int(starts[o])
            plans[r].recv_count[o] = int(counts[o])
            sel = slice(int(starts[o]), int(starts[o + 1]))
            oi = s.ghost_owner_index[sel]
            shift = s.x[s.nlocal:][sel] - subs[o].x[oi]      # image displacement seen by the receiver
            plans[o].send_index[r] = oi.astype(np.int64)
            plans[o].send_shift[r] = shift
    return plans


# ----------------------------------------------------------------------------------------------------------
# Weak scaling the way the reference does it: LAMMPS `replicate px py pz` of one periodic base box
# (/root/reference/examples/benchmark/in.lammps:31, submit_scaling.py:13-22), one replica per rank.  Every rank
# then holds the SAME local box; only the owners of its ghosts differ (the neighbour replica in direction s).

DIRECTIONS = [(sx, sy, sz) for sx in (-1, 0, 1) for sy in (-1, 0, 1) for sz in (-1, 0, 1) if (sx, sy, sz) != (0, 0, 0)]


def grid_for_ranks(n: int) -> Tuple[int, int, int]:
    """1 -> 1x1x1, 2 -> 2x1x1, 4 -> 2x2x1, 8 -> 2x2x2 (submit_scaling.py:13-22); other counts: near-cubic factors."""
    table = {1: (1, 1, 1), 2: (2, 1, 1), 4: (2, 2, 1), 8: (2, 2, 2)}
    if n in table:
        return table[n]
    g = [1, 1, 1]
    k, f = n, 2
    while k > 1:
        while k % f == 0:
            g[int(np.argmin(g))] *= f
            k //= f
        f += 1
    return tuple(sorted(g, reverse=True))


def rank_coords(rank: int, grid) -> Tuple[int, int, int]:
    ix, rem = divmod(rank, grid[1] * grid[2])
    iy, iz = divmod(rem, grid[2])
    return ix, iy, iz


def coords_rank(q, grid) -> int:
    return ((q[0] % grid[0]) * grid[1] + (q[1] % grid[1])) * grid[2] + (q[2] % grid[2])


@dataclass
class ReplicaHalo:
    """Halo exchange description of one rank in a replicated run.  Ghost atoms are re-ordered so that the block
    received from each source rank is contiguous; `send_index`/`send_shift` are already concatenated in
    destination-rank order, ready for one all_to_all_single."""
    send_index: np.ndarray        # [nsend] local atom index
    send_shift: np.ndarray        # [nsend,3] displacement to add (receiver's frame), Angstrom
    send_splits: List[int]        # per destination rank
    recv_splits: List[int]        # per source rank (sum == nghost)


def replica_halo(sub: Subdomain, box, grid, rank: int) -> ReplicaHalo:
    """Re-orders sub's ghosts in place (x, species, owner index, shift) and returns the exchange plan."""
    n = int(np.prod(grid))
    q = rank_coords(rank, grid)
    if np.abs(sub.ghost_shift).max(initial=0) > 1:
        raise ValueError("replicated runs need a base box at least as wide as the halo")
    dir_id = {d: k for k, d in enumerate(DIRECTIONS)}
    gdir = np.array([dir_id[tuple(s)] for s in sub.ghost_shift], dtype=np.int64) if sub.nghost else np.zeros(0, np.int64)
    src_of_dir = np.array([coords_rank((q[0] + d[0], q[1] + d[1], q[2] + d[2]), grid) for d in DIRECTIONS])
    dst_of_dir = np.array([coords_rank((q[0] - d[0], q[1] - d[1], q[2] - d[2]), grid) for d in DIRECTIONS])
    gsrc = src_of_dir[gdir] if sub.nghost else np.zeros(0, np.int64)
    order = np.lexsort((np.arange(sub.nghost), gdir, gsrc))      # by source rank, then direction, stable
    nl = sub.nlocal
    sub.x[nl:] = sub.x[nl:][order]
    sub.species[nl:] = sub.species[nl:][order]
    sub.global_id[nl:] = sub.global_id[nl:][order]
    sub.ghost_owner_index = sub.ghost_owner_index[order]
    sub.ghost_shift = sub.ghost_shift[order]
    sub.ghost_owner_rank = gsrc[order]
    gdir = gdir[order]
    recv_splits = np.bincount(sub.ghost_owner_rank, minlength=n).astype(int).tolist()
    # what I send "with shift d" is what I myself receive with shift d (identical replicas), to rank q - d
    send_idx, send_shift, send_dst = [], [], []
    for k, d in enumerate(DIRECTIONS):
        sel = np.nonzero(gdir == k)[0]
        send_idx.append(sub.ghost_owner_index[sel])
        send_shift.append(np.tile(np.array(d, dtype=np.float64) * np.asarray(box), (sel.size, 1)))
        send_dst.append(np.full(sel.size, dst_of_dir[k]))
    send_idx = np.concatenate(send_idx) if send_idx else np.zeros(0, np.int64)
    send_shift = np.concatenate(send_shift) if send_shift else np.zeros((0, 3))
    send_dst = np.concatenate(send_dst) if send_dst else np.zeros(0, np.int64)
    sorder = np.argsort(send_dst, kind="stable")                 # by destination rank, direction order preserved
    return ReplicaHalo(send_index=send_idx[sorder].astype(np.int64), send_shift=send_shift[sorder],
                       send_splits=np.bincount(send_dst, minlength=n).astype(int).tolist(), recv_splits=recv_splits)
