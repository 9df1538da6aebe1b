/* ani_plugin.cpp -- LAMMPS plugin entry point registering the pair styles `ani` and `ani/kk`
 * (same names, style and registration sequence as the reference's src/ani_plugin.cpp:21-40; loaded with
 * `plugin load ani_plugin.so` or through LAMMPS_PLUGIN_PATH). */
#include "lammpsplugin.h"
#include "version.h"

#include "fix_ani_md.h"
#include "pair_ani.h"

using namespace LAMMPS_NS;

static Pair* make_ani(LAMMPS* lmp) { return new PairANIB200(lmp); }
static Pair* make_ani_device(LAMMPS* lmp) { return new PairANIB200Device(lmp); }
static Fix* make_ani_md(LAMMPS* lmp, int narg, char** arg) { return new FixANIMD(lmp, narg, arg); }

extern "C" void lammpsplugin_init(void* lmp, void* handle, void* regfunc) {
  lammpsplugin_regfunc reg = (lammpsplugin_regfunc)regfunc;
  lammpsplugin_t p;
  p.version = LAMMPS_VERSION;
  p.style = "pair";
  p.author = "lammps_ani_b200";
  p.handle = handle;

  p.name = "ani";
  p.info = "ANI pair style, B200 CUDA library (libani_b200)";
  p.creator.v1 = (lammpsplugin_factory1*)&make_ani;
  (*reg)(&p, lmp);

  p.name = "ani/kk";
  p.info = "ANI pair style, device-resident variant (libani_b200)";
  p.creator.v1 = (lammpsplugin_factory1*)&make_ani_device;
  (*reg)(&p, lmp);

  // not in the reference: the device-resident integrator / thermostat next to `ani_neighbor device`
  p.style = "fix";
  p.name = "ani/md";
  p.info = "velocity Verlet + Langevin thermostat on the GPU for pair_style ani (libani_b200)";
  p.creator.v2 = (lammpsplugin_factory2*)&make_ani_md;
  (*reg)(&p, lmp);
}
