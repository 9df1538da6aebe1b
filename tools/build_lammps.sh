#!/bin/bash
# Serial stock LAMMPS (28 Mar 2023) with the PLUGIN package, built OUT OF TREE from the sources where they lie
# (default /root/reference/external/lammps/src, read-only) into tools/_lmp/ (git-ignored; travels to the GPU box).
# It is the host application for tests/test_lammps_plugin.py: `plugin load ani_plugin.so` + `pair_style ani`.
# Not the reference's build system: a flat compile of src/*.cpp + src/PLUGIN/*.cpp + STUBS/mpi.cpp with the same defines the
# plugin is compiled with (-DLAMMPS_SMALLBIG, MPI STUBS), plus the generated style_*.h / lmpinstalledpkgs.h /
# lmpgitversion.h headers written to the build directory.
set -euo pipefail
SRC=${LAMMPS_SRC:-/root/reference/external/lammps/src}
HERE=$(cd "$(dirname "$0")" && pwd)
OUT=$HERE/_lmp
OBJ=$OUT/obj
GEN=$OUT/gen
JOBS=${JOBS:-$(nproc)}
[ -f "$SRC/lammps.cpp" ] || { echo "LAMMPS sources not found under $SRC"; exit 1; }
mkdir -p "$OBJ" "$GEN"
export LC_ALL=C

style() {  # $1 macro  $2 header prefix  $3 style name
  : > "$GEN/style_$3.h"
  for f in $(cd "$SRC" && grep -sl "$1" $2*.h) $(cd "$SRC/PLUGIN" && grep -sl "$1" $2*.h 2>/dev/null || true); do
    echo "#include \"$f\"" >> "$GEN/style_$3.h"
  done
}
style ANGLE_CLASS angle_ angle;       style ATOM_CLASS atom_vec_ atom;        style BODY_CLASS body_ body
style BOND_CLASS bond_ bond;          style COMMAND_CLASS "" command;         style COMPUTE_CLASS compute_ compute
style DIHEDRAL_CLASS dihedral_ dihedral; style DUMP_CLASS dump_ dump;         style FIX_CLASS fix_ fix
style GRAN_SUB_MOD_CLASS gran_sub_mod_ gran_sub_mod; style IMPROPER_CLASS improper_ improper
style INTEGRATE_CLASS "" integrate;   style KSPACE_CLASS "" kspace;           style MINIMIZE_CLASS min_ minimize
style NBIN_CLASS nbin_ nbin;          style NPAIR_CLASS npair_ npair;         style NSTENCIL_CLASS nstencil_ nstencil
style NTOPO_CLASS ntopo_ ntopo;       style PAIR_CLASS pair_ pair;            style READER_CLASS reader_ reader
style REGION_CLASS region_ region
for n in angle atom body bond command compute dihedral dump fix gran_sub_mod improper integrate kspace minimize nbin npair \
         nstencil ntopo pair reader region; do : > "$GEN/packages_$n.h"; done
cat > "$GEN/lmpinstalledpkgs.h" <<'EOF'
#ifndef LMP_INSTALLED_PKGS_H
#define LMP_INSTALLED_PKGS_H
const char * LAMMPS_NS::LAMMPS::installed_packages[] = { "PLUGIN", NULL };
#endif
EOF
cat > "$GEN/lmpgitversion.h" <<'EOF'
#ifndef LMP_GIT_VERSION_H
#define LMP_GIT_VERSION_H
bool LAMMPS_NS::LAMMPS::has_git_info() { return false; }
const char *LAMMPS_NS::LAMMPS::git_commit() { return "(unknown)"; }
const char *LAMMPS_NS::LAMMPS::git_branch() { return "(unknown)"; }
const char *LAMMPS_NS::LAMMPS::git_descriptor() { return "(unknown)"; }
#endif
EOF

CXXFLAGS="-O2 -std=c++17 -fPIC -DLAMMPS_SMALLBIG -DLAMMPS_MEMALIGN=64 -DLMP_PLUGIN -I$GEN -I$SRC -I$SRC/PLUGIN -I$SRC/STUBS"
compile() {
  local f=$1 o=$OBJ/$(basename "${1%.cpp}").o
  if [ ! -f "$o" ] || [ "$f" -nt "$o" ]; then g++ $CXXFLAGS -c "$f" -o "$o"; fi
}
export -f compile; export OBJ CXXFLAGS
ls "$SRC"/*.cpp "$SRC"/PLUGIN/*.cpp "$SRC"/STUBS/mpi.cpp | xargs -P "$JOBS" -I{} bash -c 'compile {}'
g++ -O2 -rdynamic "$OBJ"/*.o -o "$OUT/lmp" -ldl
echo "built $OUT/lmp"
"$OUT/lmp" -h | head -3
