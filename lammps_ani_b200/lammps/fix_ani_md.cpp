/* fix_ani_md.cpp -- see fix_ani_md.h.  Per timestep (Verlet::run, external/lammps/src/verlet.cpp:229-360):
 *   initial_integrate : v += dtf f/m, x += dtv v on the device (ani_md_nve_initial).  If the coming neighbor->decide() will look
 *                       at positions (ago+1 reaches `every`) and, with `check yes`, the device-side displacement test says an
 *                       atom moved more than half the skin, x and v are copied to the host first: LAMMPS may then re-wrap,
 *                       sort and re-ghost them, and the pair style / this fix upload them again after the rebuild.
 *   [pair compute]    : PairANIB200::resident_compute -- ghosts from owners, forces into the device array.
 *   post_force        : Langevin forces on the device (ani_md_langevin), after re-uploading v / mass / tag if the atoms were
 *                       permuted by a rebuild.
 *   final_integrate   : v += dtf f/m on the device.
 *   end_of_step       : on thermo / dump steps (and the last step) x, v, f go back to the host arrays.
 */
#include "fix_ani_md.h"

#include <cmath>
#include <cstdlib>
#include <cstring>

#include "ani_b200.h"
#include "atom.h"
#include "comm.h"
#include "domain.h"
#include "error.h"
#include "force.h"
#include "neighbor.h"
#include "output.h"
#include "pair_ani.h"
#include "platform.h"
#include "update.h"

using namespace LAMMPS_NS;
using namespace FixConst;

FixANIMD::FixANIMD(LAMMPS* lmp, int narg, char** arg) : Fix(lmp, narg, arg) {
  if (narg != 3 && narg != 8) error->all(FLERR, "Illegal fix ani/md command: fix ID all ani/md [langevin Tstart Tstop damp seed]");
  if (strcmp(arg[1], "all") != 0) error->all(FLERR, "fix ani/md integrates group all");
  if (narg == 8) {
    if (strcmp(arg[3], "langevin") != 0) error->all(FLERR, "Illegal fix ani/md command");
    langevin = true;
    const double t0 = utils::numeric(FLERR, arg[4], false, lmp), t1 = utils::numeric(FLERR, arg[5], false, lmp);
    if (t0 != t1) error->all(FLERR, "fix ani/md langevin: Tstart must equal Tstop");
    t_target = t0;
    t_period = utils::numeric(FLERR, arg[6], false, lmp);
    seed = (unsigned long long)utils::inumeric(FLERR, arg[7], false, lmp);
    if (t_period <= 0.0 || t_target < 0.0) error->all(FLERR, "fix ani/md langevin: bad temperature / damping");
  }
  time_integrate = 1;
  dynamic_group_allow = 0;
  nevery = 1;
}

FixANIMD::~FixANIMD() {
  if (pair && pair->context()) {
    ani_dev_free(pair->context(), v_dev); ani_dev_free(pair->context(), mass_dev); ani_dev_free(pair->context(), tag_dev);
    pair->set_resident(false);
  }
}

int FixANIMD::setmask() { return INITIAL_INTEGRATE | POST_FORCE | FINAL_INTEGRATE | END_OF_STEP; }

void FixANIMD::init() {
  pair = dynamic_cast<PairANIB200*>(force->pair_match("^ani", 0));
  if (!pair) error->all(FLERR, "fix ani/md needs pair_style ani");
  if (atom->rmass_flag) error->all(FLERR, "fix ani/md uses per-type masses");
  pair->set_resident(true);
  dtv = update->dt;
  dtf = 0.5 * update->dt * force->ftm2v;                                     // FixNVE::init
  if (langevin) {                                                              // FixLangevin::init / post_force
    g1 = 1.0 / t_period / force->ftm2v;
    g2 = sqrt(24.0 * force->boltz / t_period / update->dt / force->mvv2e) / force->ftm2v * sqrt(t_target);
  }
}

void FixANIMD::upload_state() {
  static const bool prof = getenv("LAMMPS_ANI_PROFILING") != nullptr;
  const double t0 = prof ? platform::walltime() : 0.0;
  ani_ctx* c = pair->context();
  const int n = atom->nlocal;
  if (n > cap) {
    ani_dev_free(c, v_dev); ani_dev_free(c, mass_dev); ani_dev_free(c, tag_dev);
    cap = n + n / 8 + 64;
    if (ani_dev_alloc(c, (size_t)cap * 3 * sizeof(double), (void**)&v_dev) || ani_dev_alloc(c, (size_t)cap * sizeof(double), (void**)&mass_dev) ||
        ani_dev_alloc(c, (size_t)cap * sizeof(long long), (void**)&tag_dev))
      error->one(FLERR, "fix ani/md: device allocation failed: {}", ani_last_error());
  }
  h_mass.resize((size_t)n > 0 ? n : 1);
  h_tag.resize((size_t)n > 0 ? n : 1);
  for (int i = 0; i < n; i++) { h_mass[i] = atom->mass[atom->type[i]]; h_tag[i] = (long long)atom->tag[i]; }
  if (ani_dev_upload(c, v_dev, n > 0 ? &atom->v[0][0] : nullptr, (size_t)n * 3 * sizeof(double)) ||
      ani_dev_upload(c, mass_dev, h_mass.data(), (size_t)n * sizeof(double)) ||
      ani_dev_upload(c, tag_dev, h_tag.data(), (size_t)n * sizeof(long long)))
    error->one(FLERR, "fix ani/md: upload failed: {}", ani_last_error());
  uploaded_step = pair->res.rebuilt_step;
  if (prof) fprintf(stderr, "[fix ani/md] v / mass / tag upload after a rebuild %.3f ms\n", (platform::walltime() - t0) * 1e3);
}

void FixANIMD::sync_host(bool with_forces) {
  if (!pair->res.valid || uploaded_step != pair->res.rebuilt_step) return;     // the host arrays are the current state
  static const bool prof = getenv("LAMMPS_ANI_PROFILING") != nullptr;
  const double t0 = prof ? platform::walltime() : 0.0;
  pair->download_x();
  if (atom->nlocal > 0 && ani_dev_download(pair->context(), &atom->v[0][0], v_dev, (size_t)atom->nlocal * 3 * sizeof(double)))
    error->one(FLERR, "fix ani/md: download failed: {}", ani_last_error());
  if (with_forces) pair->download_f();
  host_current = true;
  if (prof) fprintf(stderr, "[fix ani/md] x / v%s to the host (waits for the queued steps) %.3f ms\n", with_forces ? " / f" : "", (platform::walltime() - t0) * 1e3);
}

/* Verlet::setup has just evaluated the forces from the host positions (the pair style uploaded x and built the list):
 * bring v, masses and tags up, and add the t = 0 Langevin force the way FixLangevin::setup does (post_force at setup). */
void FixANIMD::setup(int vflag) {
  upload_state();
  post_force(vflag);
  pair->download_f();
}

void FixANIMD::initial_integrate(int) {
  ani_ctx* c = pair->context();
  if (uploaded_step != pair->res.rebuilt_step) upload_state();
  if (ani_md_nve_initial(c, pair->res.x, v_dev, pair->res.f, mass_dev, atom->nlocal, dtv, dtf))
    error->one(FLERR, "fix ani/md: {}", ani_last_error());
  host_current = false;
  // Neighbor::decide(): ago++, then a rebuild (check no) or a displacement check on atom->x (check yes) when ago >= delay and
  // ago % every == 0 -- either way LAMMPS reads, and possibly permutes, the host arrays on this step
  const int ago1 = neighbor->ago + 1;
  if (ago1 >= neighbor->delay && ago1 % neighbor->every == 0) {
    // check yes, fixed box: the displacement test runs on the device first, and the host copy is only refreshed when it says
    // LAMMPS is going to rebuild (its own test on the stale host positions then says "no", consistently)
    const bool device_test = neighbor->dist_check && !domain->box_change && uploaded_step == pair->res.rebuilt_step;
    if (!device_test || pair->moved_half_skin()) sync_host(false);
  }
}

void FixANIMD::post_force(int) {
  if (uploaded_step != pair->res.rebuilt_step) upload_state();                  // a rebuild permuted the atoms: v, masses, tags again
  if (langevin && ani_md_langevin(pair->context(), v_dev, pair->res.f, mass_dev, (const int64_t*)tag_dev, atom->nlocal, g1, g2, seed,
                                  (int64_t)update->ntimestep))
    error->one(FLERR, "fix ani/md: {}", ani_last_error());
}

void FixANIMD::final_integrate() {
  if (ani_md_nve_final(pair->context(), v_dev, pair->res.f, mass_dev, atom->nlocal, dtf)) error->one(FLERR, "fix ani/md: {}", ani_last_error());
  host_current = false;
}

void FixANIMD::end_of_step() {
  if (update->ntimestep == output->next || update->ntimestep == update->laststep) sync_host(true);
}

void FixANIMD::post_run() { sync_host(true); }
