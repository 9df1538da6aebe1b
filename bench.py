#!/usr/bin/env python
"""bench.py -- ANI-2x water atom-timesteps/s on N B200 (BASELINE.json metric), one process per GPU.

    python bench.py --gpus 1 --steps 20 --warmup 5
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus N --steps K --warmup W
    python bench.py --impl reference ...      # CPU arm: the oracle port on the host cores

Workload (config.workload): the reference's `examples/water` case -- a ~300 k-atom periodic water box per GPU,
ANI-2x geometry (7 species, 1008-long AEV), all 8 ensemble members, fp32, full neighbour list at 5.1+2.0 A --
weak-scaled the way the reference does it (`replicate`: one identical replica per rank, brick grid 1x1x1 /
2x1x1 / 2x2x1 / 2x2x2).  ANI-2x weights are not available offline: seeded random weights of the ANI-2x shapes
(throughput does not depend on the values).  A step = forward halo (ghost positions) -> AEV -> ensemble MLP
fwd+bwd -> force/virial scatter -> reverse halo (ghost forces).

Three numbers per run (config.workload says which is which):
  value     the pair path on static coordinates with the list installed once (the hot path of BASELINE.json's north_star);
  value_md  the reference's benchmark protocol (examples/benchmark/in.lammps:24-26,53-59) with everything on the device: fix nve +
            fix langevin (library kernels), moving coordinates, the neighbour list rebuilt on the device every 10 steps INSIDE the
            timed region -- the quantity LAMMPS's "Loop time" measures;
  e2e       the reference-facing host call (ani_compute, host buffers, H2D + D2H inside the timed region).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

A100_PUBLISHED_M8 = 1.125e6   # 300 k water, Kokkos, 8 models, 1xA100 (examples/water/README.md:29-32)
DT_FS = 0.5


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--atoms", type=int, default=int(os.environ.get("ANI_BENCH_ATOMS", 300000)), help="atoms per GPU")
    ap.add_argument("--models", type=int, default=8)
    ap.add_argument("--mlp-mode", default=os.environ.get("ANI_B200_MLP_MODE", ""), help="simt | tf32x3 | tf32 | f16x3 (default: the library default, f16x3)")
    ap.add_argument("--cpu-sample-atoms", type=int, default=0)
    ap.add_argument("--no-cpu-baseline", action="store_true", default=bool(os.environ.get("ANI_BENCH_NO_CPU")))
    ap.add_argument("--no-md", action="store_true", help="skip the MD leg (value_md)")
    ap.add_argument("--workload", default="water", choices=["water", "combustion"],
                    help="water: BASELINE.json configs 2/4 (ANI-2x shapes); combustion: config 5 (CH4 + 2 O2 at 0.25 g/cc, ANI-1xnr "
                         "weights + RepulsionXTB, half neighbour list)")
    ap.add_argument("--nbr", default="", choices=["", "full", "half"], help="neighbour list kind (default: full for water, half for combustion)")
    ap.add_argument("--no-repulsion", action="store_true", help="combustion workload without the RepulsionXTB term")
    return ap.parse_args()


def model_path(num_models_total=8):
    from lammps_ani_b200 import aniw, extract
    d = os.path.join(ROOT, "lammps_ani_b200", "data")
    os.makedirs(d, exist_ok=True)
    p = os.path.join(d, "ani2x_synthetic_seed2024.aniw")
    if not os.path.exists(p):
        tmp = p + f".tmp{os.getpid()}"
        aniw.save(extract.synthetic_ani2x(seed=2024, num_models=num_models_total), tmp)
        os.replace(tmp, p)
    return p


def combustion_model_path(repulsion=True):
    """The in-tree ANI-1xnr ensemble (real trained weights), optionally with the RepulsionXTB tables switched on (the
    `ani1x_nr_repulsion.pt` of examples/combustion/README.md)."""
    from lammps_ani_b200 import aniw, extract
    d = os.path.join(ROOT, "lammps_ani_b200", "data")
    base = os.path.join(d, "ani1x_nr.aniw")
    if not os.path.exists(base):
        raise RuntimeError(f"{base} is missing (python __graft_entry__.py extracts it where /root/reference exists)")
    if not repulsion:
        return base
    p = os.path.join(d, "ani1x_nr_repulsion.aniw")
    if not os.path.exists(p):
        tmp = p + f".tmp{os.getpid()}"
        aniw.save(extract.with_repulsion(aniw.load(base)), tmp)
        os.replace(tmp, p)
    return p


def build_system(atoms_per_gpu, grid, rank, workload="water"):
    from lammps_ani_b200 import domain
    if workload == "combustion":
        n = max(1, atoms_per_gpu // 9)
        x, sp, box = domain.make_combustion_box(n, seed=12345)
        cells = (n, 2 * n, 0)
    else:
        cells = domain.cells_for_atoms(atoms_per_gpu)
        x, sp, box = domain.make_water_box(cells, seed=12345)
    sub = domain.decompose(x, sp, box, (1, 1, 1))[0]
    halo = domain.replica_halo(sub, box, grid, rank)
    il, nn, jl = domain.full_neighbor_list(sub.x, sub.nlocal)
    return cells, box, sub, halo, il, nn, jl


def active_aev_length(params, species_all):
    """AEV columns that can be non-zero: those of the species (pairs) present in the subdomain (library: Engine::set_active)."""
    if os.environ.get("ANI_B200_DENSE_AEV", "0") not in ("", "0"):
        return params.aev_length
    sp = len(set(int(s) for s in species_all))
    return sp * (params.radial_length // params.num_species) + sp * (sp + 1) // 2 * params.angular_sublength


def domain_list_cutoff():
    from lammps_ani_b200 import domain
    return domain.LIST_CUTOFF


def flops_per_step(params, species_local, M, in_len=None):
    """fwd 2*MACs + dE/dAEV backward 2*MACs per member.  in_len: contraction length of the first layer
    (None: the full AEV as the reference executes it; the library contracts over the active columns only)."""
    counts = np.bincount(species_local, minlength=params.num_species)
    macs = 0
    for s in range(params.num_species):
        d = list(params.dims[s])
        if in_len is not None:
            d[0] = in_len
        macs += int(counts[s]) * sum(d[i] * d[i + 1] for i in range(len(d) - 1))
    return 4.0 * M * macs


class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, pw = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for l in self.lines:
            parts = [p.strip() for p in l.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0])); mx.append(float(parts[1])); pw.append(float(parts[2]))
            except ValueError:
                continue
            for n, v in zip(names, parts[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": sorted(reasons)}


def cpu_baseline(model, sub, il, nn, jl, M, sample_atoms, repeats=1):
    """Oracle port (oracle/ani_oracle.c, OpenMP over centres) on the first `sample_atoms` local atoms."""
    from oracle.oracle import OracleANI
    ncores = os.cpu_count() or 1
    orc = OracleANI(model, use_num_models=M, nthreads=ncores)
    S = min(sample_atoms, sub.nlocal)
    off = np.concatenate([[0], np.cumsum(nn[:S].astype(np.int64))])
    nbr = jl[:off[-1]]
    best = None
    for _ in range(repeats):
        t0 = time.perf_counter()
        orc.compute_csr(sub.species, sub.x, S, off, nbr)
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    return S / best, S, ncores, best


def kernel_source_hash():
    """Short hash of the MLP / AEV kernel sources: names the ncu capture a traffic figure may be quoted from (the GPU box has no
    .git, so the commit id cannot be used)."""
    import hashlib
    h = hashlib.sha1()
    for f in ("mlp_f16_fused.cuh", "mlp_f16.cuh", "aev2.cuh"):
        with open(os.path.join(ROOT, "lammps_ani_b200", "csrc", f), "rb") as fh:
            h.update(fh.read())
    return h.hexdigest()[:10]


def measured_traffic(nlocal):
    """DRAM bytes per step of the MLP stage from an `ncu --set full` capture of THIS kernel source and THIS workload size:
    profiles/*_<source hash>_<nlocal>_raw.csv (written by tools/ncu_traffic.py from the .ncu-rep).  None when there is no such
    capture -- a figure from another build must not be quoted."""
    import csv
    import glob
    tag = kernel_source_hash()
    for path in sorted(glob.glob(os.path.join(ROOT, "profiles", f"*_{tag}_{nlocal}_raw.csv"))):
        try:
            rows = list(csv.reader(open(path)))
            hdr = rows[0]
            kn, rd, wr = hdr.index("Kernel Name"), hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
            unit_r, unit_w = rows[1][rd], rows[1][wr]
            scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
            tot, seen = 0.0, {}
            for r in rows[2:]:
                if "f16_fused" in r[kn] or "f16_gemm" in r[kn]:
                    seen.setdefault(r[kn], []).append(float(r[rd]) * scale[unit_r] + float(r[wr]) * scale[unit_w])
            # the capture holds whole steps: the launches of one step = one entry per bucket launch; average over the steps captured
            per_step = sum(sum(v) for v in seen.values())
            steps = json.load(open(path.replace("_raw.csv", "_meta.json"))).get("steps_captured", 1) if os.path.exists(path.replace("_raw.csv", "_meta.json")) else 1
            return per_step / max(steps, 1), os.path.basename(path)
        except Exception:
            continue
    return None, None


def cpu_baseline_torch(model, M, cells=(16, 16, 16)):
    """the torch-CPU pyaev restatement (oracle/pyaev_torch.py: half pair list, pair/triple tensor formulation,
    autograd -- the closest thing to what the reference's `device cpu` path executes), all host threads, on a periodic water
    box of the same density and list radius (12 288 atoms + ghosts by default: bounded by the memory of the triple tensors)."""
    import torch
    from lammps_ani_b200 import aniw, domain
    from oracle.pyaev_torch import pyaev_forward
    params = aniw.load(model)
    ncores = os.cpu_count() or 1
    torch.set_num_threads(ncores)
    x, sp, box = domain.make_water_box(cells, seed=12345)
    sub = domain.decompose(x, sp, box)[0]
    idx = domain.half_neighbor_list(sub.x, sub.nlocal)
    t0 = time.perf_counter()
    pyaev_forward(params, sub.species, sub.x, idx, sub.nlocal, num_models=M)
    dt = time.perf_counter() - t0
    return sub.nlocal / dt, sub.nlocal, ncores, dt


def run_reference(args):
    """CPU arm: times the oracle port on all host threads, same workload/metric; rank 0 only."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    from lammps_ani_b200 import aniw
    model = model_path()
    cells, box, sub, halo, il, nn, jl = build_system(min(args.atoms, 40000), (1, 1, 1), 0)
    # size the per-step sample for ~3 s of CPU work
    rate, S0, ncores, dt = cpu_baseline(model, sub, il, nn, jl, args.models, 256)
    S = args.cpu_sample_atoms or int(max(256, min(sub.nlocal, rate * 3.0)))
    for _ in range(args.warmup):
        cpu_baseline(model, sub, il, nn, jl, args.models, S)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        cpu_baseline(model, sub, il, nn, jl, args.models, S)
    total = time.perf_counter() - t0
    value = S * args.steps / total
    line = {
        "impl": "reference", "metric": "ANI-2x water atom-timesteps/s", "value": value, "unit": "atom-timesteps/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"water box, ANI-2x geometry, {args.models} members, full list 7.1 A; CPU arm evaluates a "
                               f"{S}-atom sample of the box per step", "atoms_per_gpu": args.atoms},
        "cpu_baseline": {"value": value, "unit": "atom-timesteps/s", "cores": ncores, "kind": "port",
                         "sample": f"{S} centre atoms (with their full neighbour shells) of the water box per step, "
                                   f"{args.steps} steps; oracle/ani_oracle.c with OpenMP on {ncores} threads"},
        "e2e": {"value": value, "unit": "atom-timesteps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "ns_per_day_300k": value / 300000 * DT_FS * 86400e-6,
    }
    print(json.dumps(line))
    return 0


def main():
    args = parse_args()
    if args.impl == "reference":
        return run_reference(args)
    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        print("bench.py: no CUDA device (the product path has no CPU fallback)", file=sys.stderr)
        return 2
    from lammps_ani_b200 import aniw, build, domain, runner
    from lammps_ani_b200.ddstep import GpuDomainStep

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            print(f"bench.py: --gpus {args.gpus} needs torchrun with {args.gpus} ranks", file=sys.stderr)
            return 2
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    if rank == 0:
        build.build()
        model_path()
    if world > 1:
        dist.barrier()
    combustion = args.workload == "combustion"
    nbr_kind = args.nbr or ("half" if combustion else "full")
    if rank == 0 and combustion:
        combustion_model_path(not args.no_repulsion)
        if world > 1:
            pass
    model = combustion_model_path(not args.no_repulsion) if combustion else model_path()
    params = aniw.load(model)
    grid = domain.grid_for_ranks(world)
    t_setup = time.time()
    cells, box, sub, halo, il, nn, jl = build_system(args.atoms, grid, rank, args.workload)
    M = args.models
    ani = runner.ANI(model, local_rank, M, use_cuaev=True, use_fullnbr=(nbr_kind == "full"), use_single=True)
    mode_names = {"simt": runner.MLP_FP32_SIMT, "tf32x3": runner.MLP_TF32X3, "tf32": runner.MLP_TF32, "f16x3": runner.MLP_F16X3}
    if not args.mlp_mode:
        args.mlp_mode = "f16x3"
    if args.mlp_mode:
        ani.set_mlp_mode(mode_names[args.mlp_mode])
    idx12 = domain.half_neighbor_list(sub.x, sub.nlocal) if nbr_kind == "half" else None
    step = GpuDomainStep(ani, sub, halo, dev, il, nn, jl, atom_index12=idx12)
    # N > 1: the list an MD run has after its first rebuild, built on the device; it carries the interior / boundary split that
    # lets the library overlap the NCCL ghost exchange with the interior atoms' AEV kernels (ANI_BENCH_HALO_OVERLAP=0/1 forces
    # either path at any N)
    ho = os.environ.get("ANI_BENCH_HALO_OVERLAP", "auto")
    halo_overlap = nbr_kind == "full" and (ho == "1" or (ho == "auto" and world > 1))
    if halo_overlap:
        step.use_device_list(domain_list_cutoff())
    t_setup = time.time() - t_setup
    nlocal, ntotal = sub.nlocal, sub.ntotal

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---------------- device-resident timed region (value)
    for _ in range(max(args.warmup, 3)):
        step.step_gpu()
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(args.steps):
        step.step_gpu()
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    _, launches = ani.last_timing()
    launches += 2   # halo pack + unpack kernels
    # ---------------- end-to-end: the reference-facing host call (ANI::compute full-list overload, ani.cpp:183-265 ->
    # ani_compute): HOST x[ntotal] in, HOST f[ntotal] + energy out, H2D / D2H inside the timed region every step
    x_all = np.ascontiguousarray(sub.x, dtype=np.float64)
    f_host = np.empty((ntotal, 3), dtype=np.float64)     # the pair style's persistent out_force vector (pair_ani.cpp:89-95)
    ani.set_option("pin_user", 1)                         # both arrays live until release_host_buffers() below, as in the pair style
    for _ in range(3):
        ani.compute(sub.species, x_all, nlocal=nlocal, ago=1, out_force=f_host)
    barrier()
    g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    g0.record()
    for _ in range(args.steps):
        x_all[0, 0] += 1e-9                               # new host coordinates every step, in LAMMPS's own x array
        energy, f_all, _, _ = ani.compute(sub.species, x_all, nlocal=nlocal, ago=1, out_force=f_host)
    g1.record()
    barrier()
    ms_e2e = g0.elapsed_time(g1)
    ani.release_host_buffers()
    ani.set_option("pin_user", 0)
    clocks = sampler.stop()
    # ---------------- per-stage device times (library CUDA events on the same stream), a few extra steps
    ani.enable_timing(True)
    stage = np.zeros(5)
    nst = 5
    for _ in range(nst):
        step.step_gpu(True)
        t, _n = ani.last_timing()
        stage += np.array(t)
    stage /= nst
    ani.enable_timing(False)
    # ---------------- rebuild step (`neighbor->ago == 0`): list built on the device from the device coordinates (row f2)
    # against installing a host-built list (the reference's path: host flatten + H2D of the list, pair_ani.cpp:107-151)
    rebuild_dev_ms = rebuild_host_ms = None
    lo_bb, hi_bb = sub.x.min(0) - 1e-6, sub.x.max(0) + 1e-6
    if nbr_kind == "half":
        t0 = time.perf_counter()
        ani.set_neighbors_half(sub.species, idx12, nlocal)
        torch.cuda.synchronize()
        rebuild_host_ms = (time.perf_counter() - t0) * 1e3
    else:
        for _ in range(2):   # allocations: the first build counts and fills, the following ones write rows of a fixed stride in one pass
            ani.build_neighbors(sub.species, None, nlocal, domain_list_cutoff(), lo=lo_bb, hi=hi_bb, x_dev_ptr=step.x.data_ptr())
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        npairs_dev = ani.build_neighbors(sub.species, None, nlocal, domain_list_cutoff(), lo=lo_bb, hi=hi_bb, x_dev_ptr=step.x.data_ptr())
        torch.cuda.synchronize()
        rebuild_dev_ms = (time.perf_counter() - t0) * 1e3
        t0 = time.perf_counter()
        ani.set_neighbors_full(sub.species, il, nn, jl, nlocal)
        torch.cuda.synchronize()
        rebuild_host_ms = (time.perf_counter() - t0) * 1e3
        assert npairs_dev == int(jl.size), "device-built list differs from the host list"
    # ---------------- multi-GPU force check: identical replicas make the N-rank answer equal to the single periodic domain's
    force_check = None
    if world > 1:
        step.step_gpu()
        f_multi = step.f[:nlocal].clone()
        step.step_local_periodic(lambda s: s.pair_compute(False))
        f_single = step.f[:nlocal]
        dmax = float((f_multi - f_single).abs().max())
        t = torch.tensor([dmax], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dmax = float(t[0])
        assert dmax <= 1e-4 * 23.060548, f"forces through the {world}-rank halo exchange differ from the periodic single-domain result by {dmax} kcal/mol/A"
        force_check = {"status": "ok", "max_abs_diff_kcal_mol_A": dmax, "tolerance": 1e-4 * 23.060548,
                       "what": "local forces after the NCCL ghost exchange vs the same rank's single periodic domain (ghosts served locally)"}
    # ---------------- MD leg (value_md): the reference's benchmark protocol on the device -- fix nve + fix langevin (300 K, damp
    # 100 fs), dt 0.5 fs, moving coordinates, list rebuilt on the device every 10 steps inside the timed region
    ms_md, md_info = None, None
    if not args.no_md and nbr_kind == "full":
        from lammps_ani_b200 import nve
        masses = np.array([nve.MASS[params.symbols[int(sidx)]] for sidx in sub.species[:nlocal]])
        t_md, damp_md = (2500.0, 50.0) if combustion else (300.0, 100.0)      # examples/combustion/in.lammps:53-57 / benchmark/in.lammps:53-59
        md = nve.DeviceNVE(step, masses, DT_FS, temperature=t_md, seed=12345, rebuild_every=10, langevin_damp_fs=damp_md)
        for _ in range(10):
            md.advance()
        nmd = max(10, args.steps // 10 * 10)
        barrier()
        m0, m1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        r0 = md.rebuilds
        m0.record()
        for _ in range(nmd):
            md.advance()
        m1.record()
        barrier()
        ms_md = m0.elapsed_time(m1)
        ke = md.kinetic()
        md_info = {"steps": nmd, "rebuilds_in_timed_region": md.rebuilds - r0, "dt_fs": DT_FS,
                   "temperature_K_after": 2.0 * ke / (3.0 * nlocal * nve.BOLTZ),
                   "protocol": f"fix nve + fix langevin {t_md:g} {t_md:g} {damp_md:g} (uniform noise) as library kernels, neigh_modify every 10 "
                               "(device rebuild), examples/benchmark/in.lammps:24-26,53-59",
                   "note": "seeded random weights are not a physical potential: the thermostat is exercised, the trajectory means nothing"
                           if not combustion else "real ANI-1xnr weights"}
    # ---------------- reduce over ranks (max time)
    tm = torch.tensor([ms, ms_e2e, (ms_md or 0.0) / (md_info["steps"] if md_info else 1)], dtype=torch.float64, device=dev)
    cnt = torch.tensor([float(nlocal)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
        dist.all_reduce(cnt, op=dist.ReduceOp.SUM)
    ms, ms_e2e, ms_md_step = float(tm[0]), float(tm[1]), float(tm[2])
    natoms_total = float(cnt[0])
    value = natoms_total * args.steps / (ms * 1e-3)
    value_e2e = natoms_total * args.steps / (ms_e2e * 1e-3)

    if rank == 0:
        peaks = {}
        pk = os.path.join(ROOT, "MEASURED_PEAKS.json")
        if os.path.exists(pk):
            peaks = json.load(open(pk))
        hbm = peaks.get("hbm_gbs", 6650.0)
        bf16 = peaks.get("bf16_tflops_sustained", 1400.0)
        which = "MEASURED_PEAKS.json" if peaks else "fallback (B200_PROFILING.md)"
        Lc = active_aev_length(params, sub.species)
        fl_dense = flops_per_step(params, sub.species[:nlocal], M)
        fl = flops_per_step(params, sub.species[:nlocal], M, Lc)     # what the kernels execute (structural zeros skipped)
        mlp_ms = max(stage[2], 1e-9)
        mode = {0: "fp32 FFMA (SIMT)", 1: "tcgen05 tf32x3", 2: "tcgen05 tf32", 3: "tcgen05 f16x3"}.get(
            mode_names.get(args.mlp_mode, 0), "?")
        passes = {0: 1, 1: 3, 2: 1, 3: 3}[mode_names.get(args.mlp_mode, 0)]
        tf32_peak = bf16 if args.mlp_mode == "f16x3" else bf16 / 2.0
        ach = fl / (mlp_ms * 1e-3) / 1e12
        L = Lc
        aev_bytes = nlocal * (4.0 * float(jl.size) / nlocal + 8 + 16 + 4 * L)       # list + x + AEV row store
        bwd_bytes = nlocal * (4.0 * L + 4.0 * float(jl.size) / nlocal + 16 + 12)
        traffic, traffic_src = measured_traffic(nlocal) if (args.mlp_mode in ("", "f16x3") and M == 8) else (None, None)
        roofline = {"kernel": f"ensemble MLP fwd+bwd stage ({mode}; one fused kernel per species bucket: f16_fused_kernel)", "bound": "tensor",
                    "achieved": ach, "peak": tf32_peak, "unit": "TFLOP/s", "frac": ach / tf32_peak,
                    # dram__bytes_read + write over the stage's launches of one step, parsed from the ncu capture of exactly this
                    # kernel source (hash) and workload size; null when no such capture is committed
                    "traffic": traffic, "traffic_source": traffic_src, "kernel_source_hash": kernel_source_hash(),
                    "traffic_note": "DRAM bytes per step over the MLP stage's launches (ncu --set full); the algorithmic I/O of the stage is "
                                    "1 KB per atom with 128 active AEV columns (AEV planes in, dE/dAEV out)",
                    "peak_source": (f"{which}: bf16_tflops_sustained (fp16 runs at the bf16 rate)" if args.mlp_mode == "f16x3"
                                    else f"{which}: bf16_tflops_sustained/2 taken as the dense TF32 peak"),
                    "issued_flops_factor": passes, "frac_issued": passes * ach / tf32_peak,
                    "algorithmic_flops_per_step": fl, "ms": mlp_ms,
                    "active_aev_columns": f"{Lc} of {params.aev_length}",
                    "dense_definition_flops_per_step": fl_dense,
                    "achieved_on_dense_definition": fl_dense / (mlp_ms * 1e-3) / 1e12,
                    "note": "achieved/frac count only executed FLOPs: AEV columns of species absent from the box are "
                            "structurally zero and are not multiplied (exact); SURVEY 8d's 9.82 MFLOP/atom-step counts "
                            "them, see achieved_on_dense_definition"}
        stages = {"pack_ms": stage[0], "aev_fwd_ms": stage[1], "mlp_ms": stage[2], "aev_bwd_ms": stage[3], "reduce_d2h_ms": stage[4],
                  "aev_fwd_GBps": aev_bytes / (max(stage[1], 1e-9) * 1e-3) / 1e9, "aev_bwd_GBps": bwd_bytes / (max(stage[3], 1e-9) * 1e-3) / 1e9,
                  "hbm_peak_GBps": hbm,
                  "rebuild_device_list_ms": rebuild_dev_ms, "rebuild_host_list_install_ms": rebuild_host_ms,
                  "rebuild_note": "wall ms of one neigh rebuild: ani_build_neighbors (cell list + search on the device from device "
                                  "x) vs ani_set_neighbors_full of a ready host list (species bucketing + H2D of the list); the "
                                  "host-side construction of that list is not included in the second number"}
        if combustion:
            wl = (f"examples/combustion-style reactive box (config 5): {nlocal} atoms ({cells[0]} CH4 + {cells[1]} O2 at 0.25 g/cc, seeded "
                  f"jittered-lattice placement) + {ntotal - nlocal} ghosts per GPU, ANI-1xnr (real in-tree weights, 448-long AEV; {Lc} columns "
                  f"active){'' if args.no_repulsion else ' + RepulsionXTB'}, {M} members, fp32, {nbr_kind} neighbour list 7.1 A "
                  f"({jl.size / nlocal:.1f} full entries/atom), replicate grid {grid[0]}x{grid[1]}x{grid[2]}")
        line = {
            "metric": "ANI-1xnr combustion atom-timesteps/s" if combustion else "ANI-2x water atom-timesteps/s", "value": value,
            "unit": "atom-timesteps/s", "n_gpus": world,
            "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": value / A100_PUBLISHED_M8 if (M == 8 and not combustion) else None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": wl if combustion else f"examples/water-style periodic water box, {nlocal} atoms + {ntotal - nlocal} ghosts per GPU "
                                   f"(lattice {cells[0]}x{cells[1]}x{cells[2]} molecules), ANI-2x geometry (1008 AEV, 7 species; {Lc} columns active for the species present), "
                                   f"{M} ensemble members, fp32, full neighbour list 7.1 A ({jl.size / nlocal:.1f}/atom), "
                                   f"replicate grid {grid[0]}x{grid[1]}x{grid[2]}; `value` = pair path on static coordinates, list "
                                   f"installed once; `value_md` = nve + langevin + device rebuild every 10 steps inside the timed region",
                       "weights": "ANI-1xnr, in-tree NeuroChem ensemble" if combustion else "seeded random, ANI-2x shapes (real ANI-2x parameters unavailable offline)",
                       "halo_exchange": ("ani_compute_device_halo: NCCL send / recv groups on a second stream, overlapped with the interior atoms' "
                                         "AEV kernels (device-built list)" if halo_overlap else
                                         "ani_compute_device_halo, serialised on the step's stream (host-installed list)"),
                       "mlp_mode": mode, "l2": "inputs larger than L2 (AEV rows alone exceed 126 MB); no explicit flush",
                       "vs_baseline_note": "published number is 1xA100, Kokkos, 300k water, 8 models (BASELINE.md)",
                       "setup_s": t_setup},
            "ns_per_day": (args.steps / (ms * 1e-3)) * DT_FS * 86400e-6,
            "e2e": {"value": value_e2e, "unit": "atom-timesteps/s", "h2d_bytes_per_step": int(ntotal * 24 * world),
                    "d2h_bytes_per_step": int((ntotal * 24 + 16) * world), "ms_per_step": ms_e2e / args.steps,
                    "path": "ani_compute (C ABI, host pointers): host x[ntotal] f64 -> device (DMA from the caller's array, "
                            "page-locked in place once it is seen at a stable address), AEV+MLP+force scatter, f widened to "
                            "f64 on the device -> DMA into the caller's out_force + energy; ghosts supplied by the caller "
                            "as LAMMPS does",
                    "halo": False if world > 1 else None,
                    "halo_note": ("N independent host calls: the caller (LAMMPS's CommBrick) supplies ghost positions and folds ghost "
                                  "forces; the NCCL exchange is part of `value` / `value_md` only") if world > 1 else None},
            "gpu_launches": int(launches) * args.steps, "gpu_launches_per_step": int(launches),
            "clocks": clocks, "roofline": roofline, "stages": stages,
        }
        if md_info is not None:
            line["value_md"] = natoms_total / (ms_md_step * 1e-3)
            line["ms_per_step_md"] = ms_md_step
            line["md"] = md_info
        if force_check is not None:
            line["multi_gpu_force_check"] = force_check["status"]
            line["multi_gpu_force_check_detail"] = force_check
        if not args.no_cpu_baseline and world == 1:
            rate, S0, ncores, dt = cpu_baseline(model, sub, il, nn, jl, M, 256)
            S = args.cpu_sample_atoms or int(max(256, min(nlocal, rate * 12.0)))
            rate, S, ncores, dt = cpu_baseline(model, sub, il, nn, jl, M, S)
            line["cpu_baseline"] = {"value": rate, "unit": "atom-timesteps/s", "cores": ncores, "kind": "port",
                                    "sample": f"{S} centre atoms of the same box, one pass ({dt:.1f} s), oracle/ani_oracle.c OpenMP fp64"}
            try:
                r2, S2, nc2, dt2 = cpu_baseline_torch(model, M)
                line["cpu_baseline_torch"] = {"value": r2, "unit": "atom-timesteps/s", "cores": nc2, "kind": "port",
                                              "sample": f"{S2}-atom periodic water box (same density, 7.1 A half list), one pass ({dt2:.1f} s), oracle/pyaev_torch.py "
                                                        f"(pair/triple tensors + autograd, fp64, torch {nc2} threads): SURVEY.md 8d (i)"}
            except Exception as err:      # a reported baseline, never a reason to lose the bench line
                line["cpu_baseline_torch"] = {"unavailable": str(err)[:200]}
        print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
