"""One domain-decomposed timestep of the hot path, device resident: forward halo (ghost positions) ->
pair compute -> reverse halo (ghost forces folded onto their owners).

This is the B200 mapping of what surrounds `PairANI::compute` in a multi-rank LAMMPS run
(/root/reference/src/pair_ani.cpp:197-210 manual `comm->reverse_comm(this)`;
external/lammps/src/comm_brick.cpp:538-593 forward_comm, :1023-1053 reverse_comm(Pair*)): instead of six
host-staged MPI swaps, each rank packs all 26 directions with one kernel (ani_halo_pack_x), moves them with ONE
all-to-all over NCCL/NVLink, and folds the returned ghost forces with one atomic-add kernel
(ani_halo_unpack_add_f).  torch / torch.distributed are only plumbing here: device buffers, the stream, NCCL.

`compute` and the pack / unpack operations are injected, so the world_size-2 gloo tests on CPU exercise exactly
this exchange logic with torch index ops and the oracle as the checker; the product wiring is `GpuDomainStep`.
"""
from __future__ import annotations

from typing import Callable, List, Optional

import numpy as np
import torch
import torch.distributed as dist

from . import domain


class DomainStep:
    """Exchange logic shared by the GPU product path and the CPU (gloo) tests."""

    def __init__(self, sub: domain.Subdomain, halo: domain.ReplicaHalo, device, dtype=torch.float64, group=None):
        self.sub, self.halo, self.device, self.dtype, self.group = sub, halo, device, dtype, group
        self.nlocal, self.ntotal = sub.nlocal, sub.ntotal
        self.world = dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1
        self.x = torch.tensor(sub.x, dtype=dtype, device=device)          # [ntotal,3], ghosts refreshed every step
        self.f = torch.zeros((self.ntotal, 3), dtype=dtype, device=device)
        self.send_index = torch.tensor(halo.send_index, dtype=torch.int64, device=device)
        self.send_shift = torch.tensor(halo.send_shift, dtype=dtype, device=device)
        self.nsend = int(self.send_index.numel())
        self.send_buf = torch.empty((self.nsend, 3), dtype=dtype, device=device)
        self.recv_back = torch.empty((self.nsend, 3), dtype=dtype, device=device)
        self.send_rows = list(halo.send_splits)
        self.recv_rows = list(halo.recv_splits)
        # every ghost's owner INSIDE this subdomain's own replica and its displacement from it: what a single periodic domain
        # (N = 1) would use -- identical replicas make that the exact answer of the N-rank exchange (bench.py's force check)
        nl = self.nlocal
        own = np.asarray(sub.ghost_owner_index, dtype=np.int64)
        self.ghost_owner = torch.tensor(own, dtype=torch.int32, device=device)
        self.ghost_shift = torch.tensor(sub.x[nl:] - sub.x[own], dtype=dtype, device=device).contiguous()

    # -- overridable primitives (GPU subclass swaps in the library kernels)
    def pack_x(self, out: torch.Tensor):
        torch.add(self.x.index_select(0, self.send_index), self.send_shift, out=out)

    def unpack_add_f(self, buf: torch.Tensor):
        self.f.index_add_(0, self.send_index, buf.clone())   # buf may be a view of f's ghost rows (world == 1)

    def forward_halo(self):
        """Ghost positions: x[nlocal:] <- owners' x (+ periodic shift)."""
        ghosts = self.x[self.nlocal:]
        if self.world == 1:
            self.pack_x(ghosts)                      # all 26 directions are periodic self images
            return
        self.pack_x(self.send_buf)
        dist.all_to_all_single(ghosts, self.send_buf, output_split_sizes=self.recv_rows,
                               input_split_sizes=self.send_rows, group=self.group)

    def reverse_halo(self):
        """Ghost forces: f[owner] += f[ghost]  (newton off: the pair style folds its own ghost forces)."""
        ghosts = self.f[self.nlocal:]
        if self.world == 1:
            self.unpack_add_f(ghosts)
            return
        dist.all_to_all_single(self.recv_back, ghosts.contiguous(), output_split_sizes=self.send_rows,
                               input_split_sizes=self.recv_rows, group=self.group)
        self.unpack_add_f(self.recv_back)

    def forward_halo_local(self):
        self.x[self.nlocal:] = self.x[self.ghost_owner.long()] + self.ghost_shift

    def reverse_halo_local(self):
        self.f.index_add_(0, self.ghost_owner.long(), self.f[self.nlocal:].clone())

    def step_local_periodic(self, compute: Callable[["DomainStep"], Optional[float]]):
        """The same step with the ghosts served from this rank's own atoms (no communication): the N = 1 periodic result."""
        self.forward_halo_local()
        self.f.zero_()
        e = compute(self)
        self.reverse_halo_local()
        return e

    def step(self, compute: Callable[["DomainStep"], Optional[float]]):
        """compute(self) must ADD the forces on local+ghost atoms into self.f and may return the local energy."""
        self.forward_halo()
        self.f.zero_()                               # Verlet::force_clear
        e = compute(self)
        self.reverse_halo()
        return e


class GpuDomainStep(DomainStep):
    """Product wiring: library kernels for pack / unpack, ani_compute_device for the pair compute, everything on
    torch's current CUDA stream so NCCL and the kernels stay ordered without host synchronisation."""

    def __init__(self, ani, sub, halo, device, ilist, numneigh, jlist, group=None, atom_index12=None):
        super().__init__(sub, halo, device, torch.float64, group)
        self.ani = ani
        ani.set_stream(torch.cuda.current_stream(device).cuda_stream)
        if atom_index12 is not None:                 # half-list context (`ani_neighbor half`, pair_ani.cpp:141-150)
            ani.set_neighbors_half(sub.species, atom_index12, sub.nlocal)
        else:
            ani.set_neighbors_full(sub.species, ilist, numneigh, jlist, sub.nlocal)

    def pack_x(self, out):
        self.ani.halo_pack_x(self.x.data_ptr(), False, self.send_index.data_ptr(), self.send_shift.data_ptr(), self.nsend,
                             out.data_ptr())

    def unpack_add_f(self, buf):
        assert buf.is_contiguous()
        self.ani.halo_unpack_add_f(self.f.data_ptr(), False, self.send_index.data_ptr(), buf.data_ptr(), self.nsend)

    def forward_halo_local(self):
        _md = self.ani._L.ani_md_ghost_x
        from .runner import _check
        import ctypes as C
        _check(_md(self.ani._ctx, C.c_void_p(self.x.data_ptr()), C.c_void_p(self.ghost_owner.data_ptr()),
                   C.c_void_p(self.ghost_shift.data_ptr()), self.nlocal, self.ntotal - self.nlocal))

    def reverse_halo_local(self):
        from .runner import _check
        import ctypes as C
        _check(self.ani._L.ani_md_fold_ghost_f(self.ani._ctx, C.c_void_p(self.f.data_ptr()), C.c_void_p(self.ghost_owner.data_ptr()),
                                              self.nlocal, self.ntotal - self.nlocal))

    def pair_compute(self, want_energy: bool = False):
        e, _ = self.ani.compute_device(self.x.data_ptr(), False, self.f.data_ptr(), False, 0, want_energy=want_energy)
        return e

    def step_gpu(self, want_energy: bool = False):
        return self.step(lambda s: s.pair_compute(want_energy))
