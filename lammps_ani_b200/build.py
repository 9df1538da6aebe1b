"""Build libani_b200.so in-tree with nvcc for sm_100a (cross-compiles without a GPU)."""
from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libani_b200.so")
SOURCES = ["ani_b200.cu", "model_io.cpp"]
HEADERS = ["common.cuh", "aev.cuh", "mlp_simt.cuh", "mlp_tc.cuh", "mlp_f16.cuh", "mlp_f16_fused.cuh", "nbuild.cuh", "aev2.cuh", "misc.cuh", "md.cuh", "repulsion.cuh", "halo.cuh", "../../include/ani_b200.h"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "--use_fast_math",
              "-Xcompiler", "-fPIC", "-shared", "-ccbin", "/usr/bin/g++"]


def needs_build() -> bool:
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    return any(os.path.getmtime(os.path.join(CSRC, f)) > t for f in SOURCES + HEADERS)


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and not needs_build():
        return LIB
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + [os.path.join(CSRC, f) for f in SOURCES] + \
          ["-o", LIB, "-lcudart", "-lcuda", "-ldl"]
    subprocess.check_call(cmd)
    return LIB


LAMMPS_SRC = os.environ.get("LAMMPS_SRC", "/root/reference/external/lammps/src")
PLUGIN = os.path.join(HERE, "ani_plugin.so")


def build_plugin(force: bool = False) -> str:
    """C++ LAMMPS plugin (pair styles `ani`, `ani/kk`) against the LAMMPS headers; links only libani_b200."""
    src_dir = os.path.join(HERE, "lammps")
    srcs = [os.path.join(src_dir, f) for f in ("pair_ani.cpp", "fix_ani_md.cpp", "ani_plugin.cpp")]
    if not all(os.path.exists(s) for s in srcs):
        raise RuntimeError("plugin sources not present yet")
    if not os.path.exists(os.path.join(LAMMPS_SRC, "pair.h")):
        raise RuntimeError(f"LAMMPS headers not found under {LAMMPS_SRC}")
    if not force and os.path.exists(PLUGIN) and all(os.path.getmtime(s) < os.path.getmtime(PLUGIN) for s in srcs + [os.path.join(src_dir, h) for h in ("pair_ani.h", "fix_ani_md.h")]):
        return PLUGIN
    cmd = ["/usr/bin/g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-DLAMMPS_SMALLBIG", f"-I{LAMMPS_SRC}",
           f"-I{LAMMPS_SRC}/STUBS", f"-I{os.path.join(HERE, '..', 'include')}"] + srcs + \
          ["-o", PLUGIN, f"-L{HERE}", "-lani_b200", f"-Wl,-rpath,$ORIGIN"]
    subprocess.check_call(cmd)
    return PLUGIN


def check_kokkos_branch() -> None:
    """Type-check the `-DLMP_KOKKOS` branch of the pair style (`ani/kk` on Kokkos views, reference: src/pair_ani_kokkos.cpp:67-235)
    against the in-tree LAMMPS KOKKOS package and its vendored Kokkos 3.7.01, Serial backend, `g++ -fsyntax-only`.  The branch
    cannot be RUN here: a KOKKOS LAMMPS needs Kokkos's own build system, and Kokkos 3.7.01 has no sm_100 target."""
    import tempfile
    lammps = os.path.dirname(LAMMPS_SRC)
    kk = os.path.join(lammps, "lib", "kokkos")
    if not os.path.exists(os.path.join(kk, "core", "src", "Kokkos_Core.hpp")):
        raise RuntimeError(f"Kokkos headers not found under {kk}")
    with tempfile.TemporaryDirectory() as cfg:
        with open(os.path.join(cfg, "KokkosCore_config.h"), "w") as f:
            f.write("#pragma once\n#define KOKKOS_VERSION 30701\n#define KOKKOS_ENABLE_SERIAL\n#define KOKKOS_ENABLE_CXX17\n"
                    "#define KOKKOS_ENABLE_COMPLEX_ALIGN\n#define KOKKOS_ENABLE_LIBDL\n#define KOKKOS_ENABLE_IMPL_DESUL_ATOMICS\n")
        for name, body in (("DeclareBackend", "#include <decl/Kokkos_Declare_SERIAL.hpp>\n"), ("FwdBackend", "#include <fwd/Kokkos_Fwd_SERIAL.hpp>\n"),
                           ("SetupBackend", ""), ("PostInclude", "")):
            with open(os.path.join(cfg, f"KokkosCore_Config_{name}.hpp"), "w") as f:
                f.write(body)
        inc = [cfg, f"{kk}/core/src", f"{kk}/containers/src", f"{kk}/algorithms/src", f"{kk}/tpls/desul/include", f"{LAMMPS_SRC}/KOKKOS",
               LAMMPS_SRC, f"{LAMMPS_SRC}/STUBS", os.path.join(HERE, "..", "include")]
        cmd = ["/usr/bin/g++", "-std=c++17", "-fsyntax-only", "-DLMP_KOKKOS", "-DLAMMPS_SMALLBIG"] + [f"-I{d}" for d in inc] + \
              [os.path.join(HERE, "lammps", "pair_ani.cpp")]
        subprocess.check_call(cmd)


TEST_MODEL = os.path.join(HERE, "..", "tools", "test_model")


def build_test_model(force: bool = False) -> str:
    """The stand-alone C++ known-answer driver (tools/test_model.cpp; reference: src/ani_csrc/test_model.cpp)."""
    src = os.path.join(HERE, "..", "tools", "test_model.cpp")
    deps = [src, os.path.join(HERE, "..", "tools", "test_model_golden.h"), LIB]
    if not force and os.path.exists(TEST_MODEL) and all(os.path.getmtime(d) < os.path.getmtime(TEST_MODEL) for d in deps):
        return TEST_MODEL
    subprocess.check_call(["/usr/bin/g++", "-O2", "-std=c++17", src, "-o", TEST_MODEL, f"-L{HERE}", "-lani_b200",
                           f"-Wl,-rpath,{HERE}"])
    return TEST_MODEL


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
