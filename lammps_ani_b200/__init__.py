"""lammps_ani_b200: B200-native (sm_100a) ANI energy/force evaluation behind `pair_style ani`.

Only the hot path of roitberg-group/lammps-ani lives here:
  csrc/      CUDA kernels + the C ABI (include/ani_b200.h) -> libani_b200.so
  lammps/    the C++ LAMMPS pair style / plugin that binds the C ABI
  runner.py  Python mirror of the reference `class ANI` (src/ani_csrc/ani.h:11-84) over ctypes
  extract.py one-time parameter extraction to `.aniw`
"""
__all__ = ["aniw", "extract"]
