/* fix_ani_md.h -- `fix ID group ani/md [langevin Tstart Tstop damp seed]`: velocity-Verlet (fix nve) and, optionally, the
 * Langevin thermostat (fix langevin) ON THE GPU, next to `pair_style ani ... device`.
 *
 * What it replaces in the reference's runs: `fix nve` + `fix langevin` of examples/benchmark/in.lammps:53-59, which the
 * reference executes on the GPU through the KOKKOS package (external/lammps/src/verlet.cpp:253-256,347-349 call sites;
 * fix_nve.cpp, fix_langevin.cpp arithmetic).  Here x, v, f stay in device memory between neighbour rebuilds; the host copies
 * are refreshed only when LAMMPS needs them: before a rebuild check (neigh_modify every/check), on thermo / dump steps and at
 * the end of a run.  One rank, newton off, group all, constant temperature (Tstart == Tstop). */
#ifndef LMP_FIX_ANI_MD_H
#define LMP_FIX_ANI_MD_H

#include <vector>

#include "fix.h"

namespace LAMMPS_NS {

class PairANIB200;

class FixANIMD : public Fix {
 public:
  FixANIMD(class LAMMPS*, int, char**);
  ~FixANIMD() override;
  int setmask() override;
  void init() override;
  void setup(int) override;
  void initial_integrate(int) override;
  void post_force(int) override;
  void final_integrate() override;
  void end_of_step() override;
  void post_run() override;

 protected:
  PairANIB200* pair = nullptr;
  bool langevin = false;
  double t_target = 0.0, t_period = 0.0;
  unsigned long long seed = 0;
  double dtv = 0.0, dtf = 0.0, g1 = 0.0, g2 = 0.0;
  double *v_dev = nullptr, *mass_dev = nullptr;
  long long* tag_dev = nullptr;
  int cap = 0;
  bigint uploaded_step = -2;            // pair->res.rebuilt_step the device copy of v / mass / tag belongs to
  bool host_current = true;             // atom->x / atom->v hold the current state
  std::vector<double> h_mass;
  std::vector<long long> h_tag;

  void upload_state();
  void sync_host(bool with_forces);
};

}    // namespace LAMMPS_NS
#endif
