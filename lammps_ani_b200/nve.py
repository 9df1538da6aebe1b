"""Device-resident MD driver (`fix nve` [+ `fix langevin`]) around the pair compute: positions, velocities and forces never
leave the GPU.

SURVEY.md row f3: what `modify->initial_integrate` / `post_force` / `final_integrate` do around `PairANI::compute` in the
reference's run loop (external/lammps/src/verlet.cpp:253-256,347-349; fix_nve.cpp, fix_langevin.cpp), in LAMMPS `units real`
(A, fs, g/mol, kcal/mol).  The integration kernels are the library's (csrc/md.cuh through ani_md_nve_initial / ani_md_langevin /
ani_md_nve_final); torch only owns the buffers and the stream.  The force evaluation is `GpuDomainStep.step_gpu` (ghost-position
halo -> ani_compute_device -> ghost-force fold), the neighbour list is rebuilt on the device every `rebuild_every` steps
(ani_build_neighbors).  Used by the NVE energy-drift checks that the reduced-precision MLP modes must pass
(tests/test_gpu_nve.py) and by `bench.py`'s MD leg (`value_md`), which follows the reference's benchmark protocol
(examples/benchmark/in.lammps:24-26,53-59: neigh_modify every 10, langevin 300 K damp 100 fs + nve, dt 0.5 fs).
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
                 rebuild_every: int = 0, list_cutoff: float = 7.1, langevin_damp_fs: float = 0.0):
        """step: a ddstep.GpuDomainStep (owns x [ntotal,3] f64 and f on the device); masses: [nlocal] g/mol.
        langevin_damp_fs > 0 adds `fix langevin T T damp seed` (uniform noise, as fix_langevin.cpp's default)."""
        self.step = step
        self.ani = step.ani
        self.dt = float(dt_fs)
        self.rebuild_every, self.list_cutoff = int(rebuild_every), float(list_cutoff)
        self.rebuilds = 0
        self.nstep = 0
        dev = step.x.device
        self.nlocal = step.nlocal
        self.m = torch.tensor(masses, dtype=torch.float64, device=dev).reshape(-1).contiguous()
        g = torch.Generator(device="cpu").manual_seed(seed)
        v = torch.randn((self.nlocal, 3), generator=g, dtype=torch.float64).to(dev)
        v *= torch.sqrt(BOLTZ * temperature / (self.m.reshape(-1, 1) * MVV2E))     # Maxwell-Boltzmann, `velocity ... dist gaussian`
        v -= (v * self.m.reshape(-1, 1)).sum(0, keepdim=True) / self.m.sum()       # `mom yes`
        self.v = v.contiguous()
        self.dtf = 0.5 * self.dt * FTM2V                                            # FixNVE::init: dtf = 0.5 * dt * ftm2v
        self.seed = int(seed)
        self.damp = float(langevin_damp_fs)
        if self.damp > 0.0:                                                         # FixLangevin::init
            self.g1 = 1.0 / self.damp / FTM2V
            self.g2 = float(np.sqrt(24.0 * BOLTZ / self.damp / self.dt / MVV2E) / FTM2V * np.sqrt(temperature))
        self.pe = float(self.step.step_gpu(True))                                   # forces at t = 0
        self._post_force()

    def _post_force(self):
        if self.damp > 0.0:
            s = self.step
            self.ani.md_langevin(self.v.data_ptr(), s.f.data_ptr(), self.m.data_ptr(), 0, self.nlocal, self.g1, self.g2, self.seed,
                                 self.nstep)

    def rebuild(self):
        """`neigh_modify every N`: refresh the ghosts, then build the full list on the device from the device coordinates
        (ani_build_neighbors).  Atoms do not migrate between subdomains in these short runs."""
        s = self.step
        s.forward_halo()
        if not hasattr(self, "_bb"):            # a box that holds every atom with room to move: atoms outside it are only clamped
            xmin, xmax = s.x.min(0).values.cpu().numpy(), s.x.max(0).values.cpu().numpy()
            self._bb = (xmin - 2.0, xmax + 2.0)
        s.ani.build_neighbors(s.sub.species, None, self.nlocal, self.list_cutoff, lo=self._bb[0], hi=self._bb[1],
                              x_dev_ptr=s.x.data_ptr())
        self.rebuilds += 1

    def kinetic(self) -> float:
        return 0.5 * MVV2E * self.ani.md_sum_mv2(self.v.data_ptr(), self.m.data_ptr(), self.nlocal)

    def advance(self, want_energy: bool = False):
        """One timestep of the run loop (verlet.cpp:229-360): initial_integrate, [rebuild | forward halo], force, post_force,
        final_integrate.  Returns the potential energy when asked for it."""
        s = self.step
        self.nstep += 1
        self.ani.md_nve_initial(s.x.data_ptr(), self.v.data_ptr(), s.f.data_ptr(), self.m.data_ptr(), self.nlocal, self.dt, self.dtf)
        if self.rebuild_every and self.nstep % self.rebuild_every == 0:
            self.rebuild()
        e = s.step_gpu(want_energy)                                      # ghosts refreshed inside (forward halo), f cleared
        self._post_force()
        self.ani.md_nve_final(self.v.data_ptr(), s.f.data_ptr(), self.m.data_ptr(), self.nlocal, self.dtf)
        return e

    def run(self, nsteps: int, energy_every: int = 1):
        """Returns [(step, pe, ke)] in kcal/mol; step 0 included."""
        out = [(0, self.pe, self.kinetic())]
        for k in range(1, nsteps + 1):
            want = (k % energy_every == 0) or k == nsteps
            e = self.advance(want)
            if want:
                self.pe = float(e)
                out.append((k, self.pe, self.kinetic()))
        return out
