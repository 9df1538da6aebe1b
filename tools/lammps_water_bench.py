"""The reference's own benchmark protocol (examples/benchmark/in.lammps: neighbor 2.0 bin, neigh_modify every 10 delay 0 check
yes, velocity 300 K, fix langevin + fix nve, dt 0.5 fs) run by a STOCK serial LAMMPS (tools/_lmp/lmp) with this repo's plugin:
`plugin load ani_plugin.so`, `pair_style ani 5.1 <model> cuda -1 cuaev full single`.  Everything outside the pair style
(neighbour build, integration, thermostat) is LAMMPS's own single-core CPU code here -- the reference runs those on the GPU through
Kokkos -- so the "Loop time" of this run is an upper bound on what the plain `ani` style costs end to end, and its Pair line
is the pair style's own cost (host flatten of the list on rebuild steps, H2D/D2H, kernels).

`--integrator device` replaces `fix langevin` + `fix nve` by `fix ani/md langevin ...`: integration and thermostat
run on the GPU next to `ani_neighbor device`, x / v / f stay there between rebuilds.

The model defaults to the ANI-2x-shape synthetic ensemble (what the water configs of BASELINE.json are quoted on; seeded random
weights, so the trajectory is not physical); `--model` takes any .aniw, e.g. the real ANI-1xnr weights.

usage: python tools/lammps_water_bench.py [--cells 19 19 19] [--steps 200] [--neighbor device] [--integrator device] [--model x.aniw]
"""
import argparse
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from lammps_ani_b200 import domain  # noqa: E402

MASSES = [1.008, 12.0107, 14.0067, 15.999, 32.06, 18.998403163, 35.45]

IN = """units          real
atom_style     atomic
neighbor       2.0 bin
neigh_modify   every 10 delay 0 check yes
newton         off off
boundary       p p p
read_data      {data}
plugin         load {plugin}
pair_style     ani 5.1 {model} cuda -1 cuaev {nbr} single
pair_coeff     * *
velocity       all create 300.0 12345 rot yes mom yes dist gaussian
{fixes}
timestep       0.5
thermo         {thermo}
run            {warm}
run            {steps}
"""


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cells", type=int, nargs=3, default=[19, 19, 19])      # 6859 molecules = 20 577 atoms (examples/water/20k)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warm", type=int, default=50)
    ap.add_argument("--neighbor", default="full", help="full | half | device (list built on the GPU by the library)")
    ap.add_argument("--integrator", default="lammps", help="lammps (fix langevin + fix nve on the host) | device (fix ani/md langevin)")
    ap.add_argument("--model", default="")
    ap.add_argument("--weight-scale", type=float, default=1.0,
                    help="scale of the synthetic model's hidden-layer weights; 0.25 gives forces so small that the rebuild frequency is set by "
                         "thermal motion, as with a trained potential (the default random weights heat the box to 10^4 K and every check rebuilds)")
    args = ap.parse_args()
    if not args.model and args.weight_scale != 1.0:
        from lammps_ani_b200 import aniw, extract
        args.model = os.path.join(ROOT, "lammps_ani_b200", "data", f"ani2x_synthetic_seed2024_w{args.weight_scale:g}.aniw")
        if not os.path.exists(args.model):
            os.makedirs(os.path.dirname(args.model), exist_ok=True)
            aniw.save(extract.synthetic_ani2x(seed=2024, weight_scale=args.weight_scale), args.model)
    if not args.model:
        import bench
        args.model = bench.model_path()
    label = "ANI-2x shapes, seeded random weights" if "synthetic" in os.path.basename(args.model) else os.path.basename(args.model)
    if args.weight_scale != 1.0:
        label += f" x{args.weight_scale:g} (tame forces)"
    if args.integrator == "device":
        fixes = "fix            1 all ani/md langevin 300.0 300.0 100.0 12345"
        if args.neighbor != "device":
            ap.error("--integrator device needs --neighbor device")
    else:
        fixes = "fix            1 all langevin 300.0 300.0 100.0 12345\nfix            2 all nve"
    lmp = os.path.join(ROOT, "tools", "_lmp", "lmp")
    plugin = os.path.join(ROOT, "lammps_ani_b200", "ani_plugin.so")
    x, sp, box = domain.make_water_box(tuple(args.cells), seed=12345, jitter=0.1)
    with tempfile.TemporaryDirectory() as tmp:
        data = os.path.join(tmp, "water.data")
        with open(data, "w") as f:
            f.write(f"# jittered-lattice water, {len(sp)} atoms (types: H=1 C=2 N=3 O=4 S=5 F=6 Cl=7)\n\n{len(sp)} atoms\n7 atom types\n")
            for d, n in zip(box, "xyz"):
                f.write(f"0.0 {d:.10f} {n}lo {n}hi\n")
            f.write("\nMasses\n\n" + "".join(f"{i + 1} {m}\n" for i, m in enumerate(MASSES)) + "\nAtoms\n\n")
            for i in range(len(sp)):
                f.write(f"{i + 1} {int(sp[i]) + 1} {x[i, 0]:.10f} {x[i, 1]:.10f} {x[i, 2]:.10f}\n")
        inp = os.path.join(tmp, "in.bench")
        open(inp, "w").write(IN.format(data=data, plugin=plugin, model=args.model, thermo=max(args.steps // 4, 1), warm=args.warm, nbr=args.neighbor,
                                       steps=args.steps, fixes=fixes))
        env = dict(os.environ, LD_LIBRARY_PATH=os.path.dirname(plugin) + ":" + os.environ.get("LD_LIBRARY_PATH", ""))
        r = subprocess.run([lmp, "-in", inp, "-log", "none"], cwd=tmp, env=env, capture_output=True, text=True)
        out = r.stdout
        keep = [ln for ln in out.splitlines() if any(k in ln for k in ("Loop time", "Performance", "Pair  ", "Neigh ", "Comm  ", "Modify", "Output",
                                                                      "Other ", "Neighbor list builds", "Step ", "atoms", "ERROR", "Ave neighs"))]
        print("\n".join(keep[-40:]))
        prof = [ln for ln in r.stderr.splitlines() if ln.startswith("[")]          # library profiling lines (ANI_B200_*_PROF)
        if prof:
            print("\n".join(prof[-30:]))
        if r.returncode != 0:
            print(out[-3000:], r.stderr[-2000:])
        loops = [ln for ln in out.splitlines() if ln.startswith("Loop time")]
        if loops:
            t = float(loops[-1].split()[3]); n = int(loops[-1].split()[-2]); st = int(loops[-1].split()[8])
            print(f"==> {n} atoms, {st} steps, {t:.3f} s: {n * st / t / 1e6:.3f} M atom-timesteps/s through stock serial LAMMPS (plain `ani` style, model: {label}, "
                  f"ani_neighbor {args.neighbor}, integrator {args.integrator})")
        return r.returncode


if __name__ == "__main__":
    sys.exit(main())
