"""world_size-2 test (gloo, CPU) of the domain-decomposed step: forward halo (ghost positions), pair compute on each
subdomain, reverse halo (ghost forces folded onto their owners) -- the exchange logic the GPU path runs over NCCL
(lammps_ani_b200/ddstep.py; reference: pair_ani.cpp:197-210,461-484, comm_brick.cpp:538-593,1023-1053).  The CPU oracle
is the checker here (tests/ may use it); pack / unpack are DomainStep's torch index ops."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from lammps_ani_b200 import domain
from lammps_ani_b200.ddstep import DomainStep

class GlooDomainStep(DomainStep):
    """torch.distributed transport for the CPU test: one all_to_all_single each way, the same destination-ordered send layout and
    source-ordered ghost layout the library's NCCL plan uses (csrc/halo.cuh)."""

    def exchange_forward(self, ghosts, packed):
        dist.all_to_all_single(ghosts, packed, output_split_sizes=self.recv_rows, input_split_sizes=self.send_rows)

    def exchange_reverse(self, packed, ghosts):
        dist.all_to_all_single(packed, ghosts, output_split_sizes=self.send_rows, input_split_sizes=self.recv_rows)


GRID = (2, 1, 1)
CELLS = (6, 5, 5)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _displacement(gid):
    g = np.asarray(gid, dtype=np.float64)[:, None]
    return 0.05 * np.sin(0.37 * g + np.array([0.0, 1.3, 2.1]))


def _halo_from_plan(plan):
    n = len(plan.send_index)
    return domain.ReplicaHalo(send_index=np.concatenate(plan.send_index).astype(np.int64),
                              send_shift=np.concatenate(plan.send_shift).reshape(-1, 3),
                              send_splits=[int(len(plan.send_index[r])) for r in range(n)],
                              recv_splits=[int(c) for c in plan.recv_count])


def _worker(rank, world, port, model, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle.oracle import OracleANI
        x, sp, box = domain.make_water_box(CELLS, seed=21)
        subs = domain.decompose(x, sp, box, GRID)
        plans = domain.halo_plans(subs, x)
        sub, halo = subs[rank], _halo_from_plan(plans[rank])
        step = GlooDomainStep(sub, halo, torch.device("cpu"), world=world)
        # move the LOCAL atoms only: the ghosts are stale until the forward halo refreshes them
        nl = sub.nlocal
        step.x[:nl] += torch.tensor(_displacement(sub.global_id[:nl]))
        il, nn, jl = domain.full_neighbor_list(sub.x, nl)          # list of the undisplaced positions (ago != 0 semantics)
        orc = OracleANI(model, use_num_models=2)

        def compute(s):
            r = orc.compute_full(sub.species, s.x.numpy(), il, nn, jl, nl)
            s.f += torch.tensor(r["forces"])
            return r["energy"]

        e_local = step.step(compute)
        # ghosts after the forward halo = displaced owner + the image shift they had before
        xg_expected = sub.x[nl:] + _displacement(sub.global_id[nl:])
        assert np.abs(step.x[nl:].numpy() - xg_expected).max() < 1e-12
        e = torch.tensor([e_local], dtype=torch.float64)
        dist.all_reduce(e)
        np.savez(os.path.join(out_dir, f"rank{rank}.npz"), gid=sub.global_id[:nl], f=step.f[:nl].numpy(), e=e.numpy())
    finally:
        dist.destroy_process_group()


def test_two_rank_halo_exchange_matches_single_domain(ani1xnr_path, tmp_path):
    from oracle.oracle import OracleANI
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), ani1xnr_path, str(tmp_path)), nprocs=world, join=True)
    # single-domain reference on the same displaced coordinates
    x, sp, box = domain.make_water_box(CELLS, seed=21)
    sub = domain.decompose(x, sp, box)[0]
    il, nn, jl = domain.full_neighbor_list(sub.x, sub.nlocal)
    xd = sub.x + _displacement(sub.global_id)
    ref = OracleANI(ani1xnr_path, use_num_models=2).compute_full(sub.species, xd, il, nn, jl, sub.nlocal)
    f_ref = domain.fold_ghost_forces(sub, ref["forces"])
    f_ref_by_gid = np.empty_like(f_ref)
    f_ref_by_gid[sub.global_id[:sub.nlocal]] = f_ref
    seen = 0
    for r in range(world):
        d = np.load(tmp_path / f"rank{r}.npz")
        assert np.abs(d["f"] - f_ref_by_gid[d["gid"]]).max() < 1e-8
        assert abs(d["e"][0] - ref["energy"]) < 1e-9 * abs(ref["energy"])
        seen += d["gid"].size
    assert seen == sub.nlocal


def test_replica_halo_single_rank_is_periodic_self_exchange():
    """The weak-scaling layout of bench.py at N=1: every ghost is a periodic self image; pack + fold round trip."""
    x, sp, box = domain.make_water_box((4, 4, 4), seed=3)
    sub = domain.decompose(x, sp, box)[0]
    halo = domain.replica_halo(sub, box, (1, 1, 1), 0)
    step = DomainStep(sub, halo, torch.device("cpu"))
    step.x[:sub.nlocal] += 0.01
    step.forward_halo()
    assert np.abs(step.x[sub.nlocal:].numpy() - (sub.x[sub.nlocal:] + 0.01)).max() < 1e-12
    step.f.zero_()
    step.f[sub.nlocal:] = 1.0
    step.reverse_halo()
    counts = np.bincount(sub.ghost_owner_index, minlength=sub.nlocal)
    assert np.array_equal(step.f[:sub.nlocal, 0].numpy(), counts.astype(np.float64))
