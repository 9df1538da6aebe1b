"""Probe: do two half-size contexts on two streams of ONE GPU overlap (AEV kernels of one under the HBM-bound MLP kernels of
the other)?  Prints ms per pair of half steps against ms of one full step."""
import os, sys, time
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lammps_ani_b200 import aniw, domain, extract, runner

def system(n):
    cells = domain.cells_for_atoms(n)
    x, sp, box = domain.make_water_box(cells, seed=12345)
    sub = domain.decompose(x, sp, box, (1, 1, 1))[0]
    il, nn, jl = domain.full_neighbor_list(sub.x, sub.nlocal)
    return sub, il, nn, jl

path = "/tmp/ani2x_probe.aniw"
aniw.save(extract.synthetic_ani2x(seed=2024), path)
dev = torch.device("cuda:0")
def make(n, stream):
    sub, il, nn, jl = system(n)
    ani = runner.ANI(path, 0, -1, True, True, True)
    ani.set_stream(stream.cuda_stream)
    ani.set_neighbors_full(sub.species, il, nn, jl, sub.nlocal)
    x = torch.tensor(sub.x, dtype=torch.float64, device=dev)
    f = torch.zeros_like(x)
    return ani, x, f, sub.nlocal

def run(ctxs, steps):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps):
        for ani, x, f, n in ctxs:
            ani.compute_device(x.data_ptr(), False, f.data_ptr(), False, 0, want_energy=False)
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) * 1e3 / steps

s1, s2 = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
full = [make(300000, s1)]
run(full, 3); print("one 300k context        : %.3f ms per step" % run(full, 10))
del full
halves = [make(150000, s1), make(150000, s2)]
run(halves, 3); print("two 150k contexts, 2 streams: %.3f ms per pair" % run(halves, 10))
same = [halves[0], (halves[1][0], halves[1][1], halves[1][2], halves[1][3])]
halves[1][0].set_stream(s1.cuda_stream)
run(halves, 3); print("two 150k contexts, 1 stream : %.3f ms per pair" % run(halves, 10))
