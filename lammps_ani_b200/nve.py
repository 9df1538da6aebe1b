"""Device-resident velocity-Verlet (`fix nve`) around the pair compute: positions, velocities and forces never leave the GPU.

SURVEY.md row f3: what `modify->initial_integrate` / `final_integrate` do around `PairANI::compute` in the reference's
run loop (external/lammps/src/verlet.cpp:253-256,347-349; fix_nve.cpp initial_integrate / final_integrate), in LAMMPS
`units real` (A, fs, g/mol, kcal/mol).  The two half kicks and the drift are elementwise updates on the same CUDA stream as
the library's kernels (torch is the plumbing: buffers and the stream); the force evaluation is `GpuDomainStep.step_gpu`
(ghost-position halo -> ani_compute_device -> ghost-force fold).  Used by the NVE energy-drift checks that the reduced
precision MLP modes must pass (tests/test_gpu_nve.py) and by `bench.py --md`.
"""
from __future__ import annotations

import numpy as np
import torch

FTM2V = 1.0 / 48.88821291 / 48.88821291     # LAMMPS units real: (kcal/mol/A)/(g/mol) -> A/fs^2   (update.cpp, force->ftm2v)
MVV2E = 48.88821291 * 48.88821291           # g/mol (A/fs)^2 -> kcal/mol                          (force->mvv2e)
BOLTZ = 0.0019872067                        # kcal/mol/K
MASS = {"H": 1.008, "C": 12.011, "N": 14.007, "O": 15.999, "S": 32.06, "F": 18.998, "Cl": 35.45}


class DeviceNVE:
    def __init__(self, step, masses: np.ndarray, dt_fs: float, temperature: float = 300.0, seed: int = 12345,
                 rebuild_every: int = 0, list_cutoff: float = 7.1):
        """step: a ddstep.GpuDomainStep (owns x [ntotal,3] f64 and f on the device); masses: [nlocal] g/mol."""
        self.step = step
        self.dt = float(dt_fs)
        self.rebuild_every, self.list_cutoff = int(rebuild_every), float(list_cutoff)
        self.rebuilds = 0
        dev = step.x.device
        self.nlocal = step.nlocal
        self.m = torch.tensor(masses, dtype=torch.float64, device=dev).reshape(-1, 1)
        g = torch.Generator(device="cpu").manual_seed(seed)
        v = torch.randn((self.nlocal, 3), generator=g, dtype=torch.float64).to(dev)
        v *= torch.sqrt(BOLTZ * temperature / (self.m * MVV2E))          # Maxwell-Boltzmann, `velocity ... dist gaussian`
        v -= (v * self.m).sum(0, keepdim=True) / self.m.sum()            # `mom yes`
        self.v = v
        self.dtf_over_m = 0.5 * self.dt * FTM2V / self.m
        self.pe = float(self.step.step_gpu(True))                        # forces at t = 0

    def rebuild(self):
        """`neigh_modify every N`: refresh the ghosts, then build the full list on the device from the device coordinates
        (ani_build_neighbors).  Atoms do not migrate between subdomains in these short runs."""
        s = self.step
        s.forward_halo()
        xmin, xmax = s.x.min(0).values.cpu().numpy(), s.x.max(0).values.cpu().numpy()
        s.ani.build_neighbors(s.sub.species, None, self.nlocal, self.list_cutoff, lo=xmin - 1e-6, hi=xmax + 1e-6,
                              x_dev_ptr=s.x.data_ptr())
        self.rebuilds += 1

    def kinetic(self) -> float:
        return float(0.5 * MVV2E * (self.m * self.v * self.v).sum())

    def run(self, nsteps: int, energy_every: int = 1):
        """Returns [(step, pe, ke)] in kcal/mol; step 0 included."""
        s = self.step
        out = [(0, self.pe, self.kinetic())]
        for k in range(1, nsteps + 1):
            f = s.f[: self.nlocal]
            self.v.addcmul_(f, self.dtf_over_m)                          # FixNVE::initial_integrate
            s.x[: self.nlocal].add_(self.v, alpha=self.dt)
            if self.rebuild_every and k % self.rebuild_every == 0:
                self.rebuild()
            want = (k % energy_every == 0) or k == nsteps
            e = s.step_gpu(want)                                         # ghosts refreshed inside (forward halo)
            self.v.addcmul_(s.f[: self.nlocal], self.dtf_over_m)         # FixNVE::final_integrate
            if want:
                self.pe = float(e)
                out.append((k, self.pe, self.kinetic()))
        return out
