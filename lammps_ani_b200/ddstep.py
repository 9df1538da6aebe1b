"""One domain-decomposed timestep of the hot path, device resident: forward halo (ghost positions) ->
pair compute -> reverse halo (ghost forces folded onto their owners).

This is the B200 mapping of what surrounds `PairANI::compute` in a multi-rank LAMMPS run
(/root/reference/src/pair_ani.cpp:197-210 manual `comm->reverse_comm(this)`;
external/lammps/src/comm_brick.cpp:538-593 forward_comm, :1023-1053 reverse_comm(Pair*)).  The exchange itself lives in
the library (csrc/halo.cuh: ani_halo_plan_create / ani_halo_forward / ani_halo_reverse -- one pack kernel, one
ncclSend / ncclRecv group straight into the ghost slab, one fold kernel); this module only builds the plan and calls it.
`DomainStep` holds the exchange-independent logic (buffers, step order) with pluggable transport, so the world_size-2 gloo
tests on CPU (tests/test_domain_gloo.py supplies a torch.distributed transport there) check the same plan layout the GPU
path hands to the library; torch.distributed is used here only as the control plane that distributes the ncclUniqueId.
"""
from __future__ import annotations

from typing import Callable, List, Optional

import numpy as np
import torch

from . import domain


class DomainStep:
    """Step order and buffers shared by the GPU product path and the CPU (gloo) tests.  Subclasses provide the transport:
    `exchange_forward(ghost_rows_out, packed_in)` and `exchange_reverse(packed_out, ghost_rows_in)`."""

    def __init__(self, sub: domain.Subdomain, halo: domain.ReplicaHalo, device, dtype=torch.float64, world: int = 1):
        self.sub, self.halo, self.device, self.dtype = sub, halo, device, dtype
        self.nlocal, self.ntotal = sub.nlocal, sub.ntotal
        self.world = world
        self.x = torch.tensor(sub.x, dtype=dtype, device=device)          # [ntotal,3], ghosts refreshed every step
        self.f = torch.zeros((self.ntotal, 3), dtype=dtype, device=device)
        self.send_index = torch.tensor(halo.send_index, dtype=torch.int64, device=device)
        self.send_shift = torch.tensor(halo.send_shift, dtype=dtype, device=device)
        self.nsend = int(self.send_index.numel())
        self.send_buf = torch.empty((self.nsend, 3), dtype=dtype, device=device)
        self.recv_back = torch.empty((self.nsend, 3), dtype=dtype, device=device)
        self.send_rows = list(halo.send_splits)
        self.recv_rows = list(halo.recv_splits)

    # -- primitives on torch tensors (CPU tests); the GPU subclass hands the whole exchange to the library instead
    def pack_x(self, out: torch.Tensor):
        torch.add(self.x.index_select(0, self.send_index), self.send_shift, out=out)

    def unpack_add_f(self, buf: torch.Tensor):
        self.f.index_add_(0, self.send_index, buf.clone())   # buf may be a view of f's ghost rows (world == 1)

    def exchange_forward(self, ghosts: torch.Tensor, packed: torch.Tensor):
        raise NotImplementedError("multi-rank transport is provided by a subclass")

    def exchange_reverse(self, packed: torch.Tensor, ghosts: torch.Tensor):
        raise NotImplementedError("multi-rank transport is provided by a subclass")

    def forward_halo(self):
        """Ghost positions: x[nlocal:] <- owners' x (+ periodic shift)."""
        ghosts = self.x[self.nlocal:]
        if self.world == 1:
            self.pack_x(ghosts)                      # all 26 directions are periodic self images
            return
        self.pack_x(self.send_buf)
        self.exchange_forward(ghosts, self.send_buf)

    def reverse_halo(self):
        """Ghost forces: f[owner] += f[ghost]  (newton off: the pair style folds its own ghost forces)."""
        ghosts = self.f[self.nlocal:]
        if self.world == 1:
            self.unpack_add_f(ghosts)
            return
        self.exchange_reverse(self.recv_back, ghosts.contiguous())
        self.unpack_add_f(self.recv_back)

    def step(self, compute: Callable[["DomainStep"], Optional[float]]):
        """compute(self) must ADD the forces on local+ghost atoms into self.f and may return the local energy."""
        self.forward_halo()
        self.f.zero_()                               # Verlet::force_clear
        e = compute(self)
        self.reverse_halo()
        return e


class GpuDomainStep(DomainStep):
    """Product wiring: the library's halo plan for the exchange (NCCL groups on the context's stream), ani_compute_device for
    the pair compute, everything on torch's current CUDA stream so that nothing synchronises with the host."""

    def __init__(self, ani, sub, halo, device, ilist, numneigh, jlist, group=None, atom_index12=None):
        import torch.distributed as dist
        world = dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1
        rank = dist.get_rank(group) if world > 1 else 0
        super().__init__(sub, halo, device, torch.float64, world)
        self.ani = ani
        ani.set_stream(torch.cuda.current_stream(device).cuda_stream)
        if atom_index12 is not None:                 # half-list context (`ani_neighbor half`, pair_ani.cpp:141-150)
            ani.set_neighbors_half(sub.species, atom_index12, sub.nlocal)
        else:
            ani.set_neighbors_full(sub.species, ilist, numneigh, jlist, sub.nlocal)
        uid = None
        if world > 1:                                # control plane only: hand rank 0's ncclUniqueId to everybody
            box = [ani.halo_unique_id() if rank == 0 else None]
            dist.broadcast_object_list(box, src=0, group=group)
            uid = box[0]
        self.plan = ani.halo_plan_create(uid, world, rank, sub.nlocal, halo.send_index, halo.send_shift, halo.send_splits, halo.recv_splits)
        # every ghost's owner INSIDE this subdomain's own replica and its displacement from it: what a single periodic domain
        # (N = 1) would use -- identical replicas make that the exact answer of the N-rank exchange (bench.py's force check)
        nl = self.nlocal
        own = np.asarray(sub.ghost_owner_index, dtype=np.int64)
        self.ghost_owner = torch.tensor(own, dtype=torch.int32, device=device)
        self.ghost_shift = torch.tensor(sub.x[nl:] - sub.x[own], dtype=torch.float64, device=device).contiguous()

    def forward_halo(self):
        self.ani.halo_forward(self.plan, self.x.data_ptr())

    def reverse_halo(self):
        self.ani.halo_reverse(self.plan, self.f.data_ptr())

    def forward_halo_local(self):
        from .runner import _check
        import ctypes as C
        _check(self.ani._L.ani_md_ghost_x(self.ani._ctx, C.c_void_p(self.x.data_ptr()), C.c_void_p(self.ghost_owner.data_ptr()),
                                         C.c_void_p(self.ghost_shift.data_ptr()), self.nlocal, self.ntotal - self.nlocal))

    def reverse_halo_local(self):
        from .runner import _check
        import ctypes as C
        _check(self.ani._L.ani_md_fold_ghost_f(self.ani._ctx, C.c_void_p(self.f.data_ptr()), C.c_void_p(self.ghost_owner.data_ptr()),
                                              self.nlocal, self.ntotal - self.nlocal))

    def step_local_periodic(self, compute):
        """The same step with the ghosts served from this rank's own atoms (no communication): the N = 1 periodic result."""
        self.forward_halo_local()
        self.f.zero_()
        e = compute(self)
        self.reverse_halo_local()
        return e

    def pair_compute(self, want_energy: bool = False):
        e, _ = self.ani.compute_device(self.x.data_ptr(), False, self.f.data_ptr(), False, 0, want_energy=want_energy)
        return e

    def use_device_list(self, cutoff: float):
        """Replace the host-built list by one built on the device from the current positions (ani_build_neighbors).  Besides
        being what an MD run does on rebuild steps, a device-built list carries the interior / boundary split of the local
        atoms that lets the library overlap the ghost exchange with the interior atoms' AEV kernels."""
        self.forward_halo()
        lo = (self.x.min(0).values - 1e-6).cpu().numpy()
        hi = (self.x.max(0).values + 1e-6).cpu().numpy()
        return self.ani.build_neighbors(self.sub.species, None, self.nlocal, cutoff, lo=lo, hi=hi, x_dev_ptr=self.x.data_ptr())

    def step_gpu(self, want_energy: bool = False):
        """One step through ani_compute_device_halo: ghost positions in, forces, ghost forces back -- one library call."""
        self.f.zero_()                               # Verlet::force_clear
        e, _ = self.ani.compute_device_halo(self.plan, self.x.data_ptr(), self.f.data_ptr(), want_energy=want_energy)
        return e

    def step_gpu_serial(self, want_energy: bool = False):
        """The same step as three library calls (exchange, compute, exchange) on one stream: the reference for the fused call."""
        return self.step(lambda s: s.pair_compute(want_energy))
