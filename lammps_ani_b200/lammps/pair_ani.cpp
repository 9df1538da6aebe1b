/* pair_ani.cpp -- per-timestep driver of `pair_style ani` over libani_b200.
 *
 * Behaviour follows the reference driver PairANI::compute (src/pair_ani.cpp:66-233): positions of local + ghost
 * atoms in, neighbour data re-flattened only when the list was rebuilt (neighbor->ago == 0, :107-151), forces on
 * local + ghost atoms out, ghost forces folded back with reverse_comm(this) on `out_force` when both newton
 * flags are off (:192-201), then f += out_force, eng_vdwl, eatom, virial in LAMMPS' xx,yy,zz,xy,xz,yz order
 * (:204-231).  What differs: no per-step std::vector allocations, no coordinate copy (atom->x is contiguous),
 * the model evaluation is a C call into the CUDA library.
 */
#include "pair_ani.h"

#include <algorithm>
#include <cstdlib>
#include <cstring>

#include "atom.h"
#include "comm.h"
#include "error.h"
#include "force.h"
#include "memory.h"
#include "neigh_list.h"
#include "neighbor.h"
#include "platform.h"
#include "domain.h"
#include "comm.h"
#include "update.h"

#ifdef LMP_KOKKOS
#include "atom_kokkos.h"
#include "atom_masks.h"
#include "kokkos.h"
#include "memory_kokkos.h"
#include "neigh_list_kokkos.h"
#include "neigh_request.h"
#endif

using namespace LAMMPS_NS;

PairANIB200::PairANIB200(LAMMPS* lmp)
    : Pair(lmp), cutoff(0.0), use_num_models(-1), use_cuaev(true), use_fullnbr(true), use_single(true), model(nullptr),
      ctx(nullptr), npairs(0) {
  writedata = 0;
  single_enable = 0;
  restartinfo = 1;
  manybody_flag = 1;
  if (strcmp(update->unit_style, "real") != 0) error->all(FLERR, "Pair ani requires real units");
  comm_reverse = 3;       // pair_ani.cpp:47-48
  comm_reverse_off = 3;
}

PairANIB200::~PairANIB200() {
  if (allocated) {
    memory->destroy(setflag);
    memory->destroy(cutsq);
  }
  close_model();
}

void PairANIB200::fail(const char* what) {
  error->one(FLERR, "pair ani: {}: {}", what, ani_last_error());
}

void PairANIB200::close_model() {
  resident_free();
  if (ctx) ani_ctx_destroy(ctx);
  if (model) ani_model_free(model);
  ctx = nullptr;
  model = nullptr;
}

/* rank -> GPU: node-local rank modulo the device count (pair_ani.cpp:255-283).  The device count comes from
 * CUDA_VISIBLE_DEVICES-independent probing inside ani_ctx_create; here only the local rank is needed. */
int PairANIB200::local_device(const std::string& device) {
  if (device == "cpu") error->all(FLERR, "pair ani: device 'cpu' is not available in the B200 build (no CPU path)");
  if (device != "cuda") error->all(FLERR, "pair ani: 3rd argument must be <cpu/cuda>");
  int rank = 0, local_rank = 0;
  MPI_Comm_rank(world, &rank);
#if defined(MPI_COMM_TYPE_SHARED) && !defined(MPI_STUBS)
  MPI_Comm local;
  MPI_Comm_split_type(world, MPI_COMM_TYPE_SHARED, rank, MPI_INFO_NULL, &local);
  MPI_Comm_rank(local, &local_rank);
  MPI_Comm_free(&local);
#else
  local_rank = rank;
#endif
  int ndev = 1;
  if (const char* e = getenv("ANI_B200_NUM_DEVICES")) ndev = atoi(e) > 0 ? atoi(e) : 1;
  else if (ani_device_count(&ndev) != 0 || ndev < 1) error->all(FLERR, "pair ani: no CUDA device available");
  return local_rank % ndev;
}

void PairANIB200::open_model() {
  close_model();
  const int dev = local_device(device_str);
  if (ani_model_load(model_file.c_str(), &model)) fail("cannot load the model");
  if (ani_ctx_create(model, dev, use_num_models, use_cuaev ? 1 : 0, use_fullnbr ? 1 : 0, use_single ? 1 : 0, &ctx))
    fail("cannot create the CUDA context");
  // atom->x and out_force keep their address between reallocations, and compute() releases the page-locks before either
  // moves (nmax / vector growth): let the library DMA straight from / to them
  ani_ctx_set_option(ctx, "pin_user", 1);
}

/* pair_style ani <cutoff> <model_file> <cpu|cuda> [num_models] [cuaev|pyaev] [full|half] [single|double] */
void PairANIB200::settings(int narg, char** arg) {
  if (narg < 3) error->all(FLERR, "Illegal pair_style command");
  cutoff = utils::numeric(FLERR, arg[0], false, lmp);
  model_file = arg[1];
  device_str = arg[2];
  use_num_models = narg > 3 ? utils::inumeric(FLERR, arg[3], false, lmp) : -1;
  use_cuaev = use_fullnbr = use_single = true;
  if (narg > 4) {
    if (strcmp(arg[4], "cuaev") == 0) use_cuaev = true;
    else if (strcmp(arg[4], "pyaev") == 0) use_cuaev = false;
    else error->all(FLERR, "ani_aev should be cuaev or pyaev");
  }
  if (narg > 5) {
    device_list = false;
    if (strcmp(arg[5], "full") == 0) use_fullnbr = true;
    else if (strcmp(arg[5], "half") == 0) use_fullnbr = false;
    else if (strcmp(arg[5], "device") == 0) { use_fullnbr = true; device_list = true; }   // extension: see init_style()
    else error->all(FLERR, "ani_neighbor should be full, half or device");
  }
  if (narg > 6) {
    if (strcmp(arg[6], "single") == 0) use_single = true;
    else if (strcmp(arg[6], "double") == 0) use_single = false;
    else error->all(FLERR, "precision should be single or double");
  }
  open_model();
}

void PairANIB200::allocate() {
  allocated = 1;
  const int n = atom->ntypes;
  memory->create(setflag, n + 1, n + 1, "pair:setflag");
  for (int i = 1; i <= n; i++)
    for (int j = i; j <= n; j++) setflag[i][j] = 0;
  memory->create(cutsq, n + 1, n + 1, "pair:cutsq");
}

void PairANIB200::coeff(int narg, char** arg) {
  if (!allocated) allocate();
  const char* msg = "Incorrect args for pair coefficients, it should be set as: pair_coeff * *";
  if (narg != 2) error->all(FLERR, msg);
  int ilo, ihi, jlo, jhi;
  const int n = atom->ntypes;
  utils::bounds(FLERR, arg[0], 1, n, ilo, ihi, error);
  utils::bounds(FLERR, arg[1], 1, n, jlo, jhi, error);
  if (ilo != 1 || jlo != 1 || ihi != n || jhi != n) error->all(FLERR, msg);
  for (int i = 1; i <= n; i++)
    for (int j = i; j <= n; j++) setflag[i][j] = 1;
}

void PairANIB200::init_style() {
  // the model needs every neighbour of a local atom, ghosts included: newton_pair off (pair_ani.cpp:347-356)
  if (force->newton_pair == 1)
    error->all(FLERR, use_fullnbr ? "Pair style ANI requires newton pair off when using full neighbor list"
                                  : "Pair style ANI requires newton pair off when using half neighbor list");
  // `ani_neighbor device` (not in the reference): no list is requested from LAMMPS at all; on rebuild steps the library builds
  // the same full list (all j within cutoff + skin of a local i, ghosts included) on the GPU from atom->x
  // (ani_build_neighbors).  LAMMPS still decides WHEN to rebuild (neigh_modify, displacement check) and still sizes the ghost
  // shell from cutoff + skin; its own CPU binning / list build and the host flatten + H2D of the list disappear.
  if (device_list) {
    // the device build bins in Cartesian space over [sublo - cutghost, subhi + cutghost): orthogonal boxes only
    if (domain->triclinic) error->all(FLERR, "Pair style ANI: ani_neighbor device needs an orthogonal box (use full or half)");
    return;
  }
  if (use_fullnbr) neighbor->add_request(this, NeighConst::REQ_FULL);
  else neighbor->add_request(this);
}

double PairANIB200::init_one(int, int) { return cutoff; }

void* PairANIB200::extract(const char*, int&) { return nullptr; }

double PairANIB200::memory_usage() {
  return (double)(species.capacity() * sizeof(int) + jlist.capacity() * sizeof(int) +
                  atom_index12.capacity() * sizeof(int64_t) + (out_force.capacity() + out_eatom.capacity()) * sizeof(double));
}

/* ------------------------------------------------------------------------------------------------ device-resident mode */
void PairANIB200::resident_free() {
  if (!ctx) return;
  ani_dev_free(ctx, res.x); ani_dev_free(ctx, res.f); ani_dev_free(ctx, res.shift); ani_dev_free(ctx, res.owner); ani_dev_free(ctx, res.x0);
  res = Resident();
}

void PairANIB200::set_resident(bool on) {
  if (on && !device_list) error->all(FLERR, "fix ani/md needs pair_style ani ... with ani_neighbor device");
  if (on && (force->newton || comm->nprocs > 1))
    error->all(FLERR, "fix ani/md: the device-resident mode covers one rank with newton off (ghosts are periodic images of local atoms)");
  res.on = on;
  res.valid = false;
}

void PairANIB200::download_x() {
  if (res.valid && atom->nlocal > 0 && ani_dev_download(ctx, &atom->x[0][0], res.x, (size_t)atom->nlocal * 3 * sizeof(double)))
    fail("download of the positions failed");
}

/* Neighbor::check_distance (external/lammps/src/neighbor.cpp) on the device copy: has any local atom moved more than half the
 * skin since the last build?  LAMMPS's own test reads atom->x, which is stale in the device-resident mode -- the fix calls
 * this first and copies x, v back only when the answer is yes, so that a check that would not rebuild costs 8 bytes. */
bool PairANIB200::moved_half_skin() {
  if (!res.valid || !res.x0) return true;
  double d2 = 0.0;
  if (ani_md_max_disp2(ctx, res.x, res.x0, atom->nlocal, &d2)) fail("displacement check failed");
  return d2 > 0.25 * neighbor->skin * neighbor->skin;
}

void PairANIB200::download_f() {
  if (res.valid && atom->nlocal > 0 && ani_dev_download(ctx, &atom->f[0][0], res.f, (size_t)atom->nlocal * 3 * sizeof(double)))
    fail("download of the forces failed");
}

/* One step with x / f on the device.  Rebuild steps (neighbor->ago == 0): LAMMPS has just re-wrapped, sorted and re-ghosted the
 * HOST arrays, so positions (local + ghost) go up once, each ghost's owner and displacement are recorded (every ghost of a
 * single periodic rank is an image of a local atom: atom->map(tag)), and the list is built on the device.  Other steps: the
 * local positions on the device were advanced by fix ani/md; ghosts = owner + shift (CommBrick::forward_comm's job).  Forces
 * are accumulated into the device array and the ghost part is folded onto the owners there (pair_ani.cpp:197-201). */
void PairANIB200::resident_compute(int eflag, int vflag) {
  const int nlocal = atom->nlocal, nghost = atom->nghost, ntotal = nlocal + nghost;
  double** x = atom->x;
  // LAMMPS_ANI_PROFILING (the reference's switch, pair_ani.cpp:49): wall time of the phases of a rebuild step, to stderr
  static const bool prof = getenv("LAMMPS_ANI_PROFILING") != nullptr;
  double tp[5] = {0, 0, 0, 0, 0};
  if (neighbor->ago == 0 || !res.valid) {
    if (prof) tp[0] = platform::walltime();
    // LAMMPS's per-atom arrays are pageable: page-lock them once per allocation so the rebuild-step transfers are direct DMA
    if (atom->nmax != last_nmax) {
      ani_ctx_release_host_buffers(ctx);
      last_nmax = atom->nmax;
      const size_t bytes = (size_t)atom->nmax * 3 * sizeof(double);
      if (atom->nmax > 0 && (ani_ctx_pin_host(ctx, &atom->x[0][0], bytes) || ani_ctx_pin_host(ctx, &atom->v[0][0], bytes) ||
                             ani_ctx_pin_host(ctx, &atom->f[0][0], bytes)))
        fail("page-locking the atom arrays failed");
    }
    if (nlocal > res.cap_x0) {
      ani_dev_free(ctx, res.x0);
      res.cap_x0 = nlocal + nlocal / 8 + 64;
      if (ani_dev_alloc(ctx, (size_t)res.cap_x0 * 3 * sizeof(double), (void**)&res.x0)) fail("device allocation failed");
    }
    if (ntotal > res.cap_total || nghost > res.cap_ghost) {
      ani_dev_free(ctx, res.x); ani_dev_free(ctx, res.f); ani_dev_free(ctx, res.shift); ani_dev_free(ctx, res.owner);
      res.cap_total = ntotal + ntotal / 8 + 64; res.cap_ghost = nghost + nghost / 8 + 64;
      if (ani_dev_alloc(ctx, (size_t)res.cap_total * 3 * sizeof(double), (void**)&res.x) ||
          ani_dev_alloc(ctx, (size_t)res.cap_total * 3 * sizeof(double), (void**)&res.f) ||
          ani_dev_alloc(ctx, (size_t)res.cap_ghost * 3 * sizeof(double), (void**)&res.shift) ||
          ani_dev_alloc(ctx, (size_t)res.cap_ghost * sizeof(int), (void**)&res.owner))
        fail("device allocation failed");
    }
    const int* type = atom->type;
    const tagint* tag = atom->tag;
    if ((size_t)ntotal > species.size()) species.resize((size_t)ntotal + ntotal / 2);
    for (int i = 0; i < ntotal; i++) species[i] = type[i] - 1;
    h_owner.resize((size_t)std::max(nghost, 1));
    h_shift.resize((size_t)std::max(nghost, 1) * 3);
    // every ghost of a single periodic rank is an image of a local atom: tag -> local index (no atom map needed)
    tagint maxtag = 0;
    for (int i = 0; i < nlocal; i++) maxtag = std::max(maxtag, tag[i]);
    h_tag2local.assign((size_t)maxtag + 1, -1);
    for (int i = 0; i < nlocal; i++) h_tag2local[tag[i]] = i;
    for (int g = 0; g < nghost; g++) {
      const tagint tg = tag[nlocal + g];
      const int o = (tg >= 0 && tg <= maxtag) ? h_tag2local[tg] : -1;
      if (o < 0) error->one(FLERR, "fix ani/md: a ghost atom has no local owner");
      h_owner[g] = o;
      for (int d = 0; d < 3; d++) h_shift[3 * (size_t)g + d] = x[nlocal + g][d] - x[o][d];
    }
    if (prof) tp[1] = platform::walltime();
    if (ani_dev_upload(ctx, res.x, ntotal > 0 ? &x[0][0] : nullptr, (size_t)ntotal * 3 * sizeof(double)) ||
        ani_dev_upload(ctx, res.x0, nlocal > 0 ? &x[0][0] : nullptr, (size_t)nlocal * 3 * sizeof(double)) ||
        ani_dev_upload(ctx, res.owner, h_owner.data(), (size_t)nghost * sizeof(int)) ||
        ani_dev_upload(ctx, res.shift, h_shift.data(), (size_t)nghost * 3 * sizeof(double)))
      fail("upload failed");
    double lo[3], hi[3];
    for (int d = 0; d < 3; d++) {
      lo[d] = domain->sublo[d] - comm->cutghost[d];
      hi[d] = domain->subhi[d] + comm->cutghost[d];
    }
    int64_t np = 0;
    if (prof) tp[2] = platform::walltime();
    if (ani_build_neighbors(ctx, nlocal, ntotal, species.data(), res.x, 1, neighbor->cutneighmax, lo, hi, &np))
      fail("device neighbour build failed");
    if (prof) {
      tp[3] = platform::walltime();
      fprintf(stderr, "[pair ani rebuild] host index work %.3f ms | uploads (x, x0, owners, shifts) %.3f | device list %.3f | %d local + %d ghost atoms\n",
              (tp[1] - tp[0]) * 1e3, (tp[2] - tp[1]) * 1e3, (tp[3] - tp[2]) * 1e3, nlocal, nghost);
    }
    npairs = np;
    res.nlocal = nlocal; res.nghost = nghost;
    res.valid = true;
    res.rebuilt_step = update->ntimestep;
  } else {
    if (nlocal != res.nlocal || nghost != res.nghost) error->one(FLERR, "fix ani/md: atom counts changed without a neighbour rebuild");
    if (ani_md_ghost_x(ctx, res.x, res.owner, res.shift, nlocal, nghost)) fail("ghost update failed");
  }
  if (ani_dev_zero(ctx, res.f, (size_t)ntotal * 3 * sizeof(double))) fail("force clear failed");
  double energy = 0.0, vir9[9];
  if (ani_compute_device(ctx, res.x, 0, res.f, 0, nullptr, eflag_global ? &energy : nullptr, vflag_either ? vir9 : nullptr))
    fail("evaluation failed");
  if (ani_md_fold_ghost_f(ctx, res.f, res.owner, nlocal, nghost)) fail("ghost force fold failed");
  if (eflag_atom) error->one(FLERR, "fix ani/md: per-atom energies are not available in the device-resident mode");
  if (eflag_global) eng_vdwl += energy;
  if (vflag_either) {
    virial[0] += vir9[0]; virial[1] += vir9[4]; virial[2] += vir9[8];
    virial[3] += vir9[1]; virial[4] += vir9[2]; virial[5] += vir9[5];
  }
}

void PairANIB200::compute(int eflag, int vflag) {
  ev_init(eflag, vflag);
  if (res.on) { resident_compute(eflag, vflag); return; }
  const int nlocal = atom->nlocal;
  const int ntotal = nlocal + atom->nghost;
  double** x = atom->x;
  double** f = atom->f;
  // the library page-locks atom->x and out_force in place once they are stable: let go before either is reallocated
  if (atom->nmax != last_nmax || (size_t)ntotal * 3 > out_force.size()) {
    ani_ctx_release_host_buffers(ctx);
    last_nmax = atom->nmax;
  }
  if ((size_t)ntotal * 3 > out_force.size()) out_force.resize((size_t)ntotal * 3 + (size_t)ntotal * 3 / 2);

  if (neighbor->ago == 0 && device_list) {
    const int* type = atom->type;
    if ((size_t)ntotal > species.size()) species.resize((size_t)ntotal + ntotal / 2);
    for (int i = 0; i < ntotal; i++) species[i] = type[i] - 1;
    double lo[3], hi[3];
    for (int d = 0; d < 3; d++) {
      lo[d] = domain->sublo[d] - comm->cutghost[d];
      hi[d] = domain->subhi[d] + comm->cutghost[d];
    }
    int64_t np = 0;
    const double* xf = ntotal > 0 ? &x[0][0] : nullptr;
    if (ani_build_neighbors(ctx, nlocal, ntotal, species.data(), xf, 0, neighbor->cutneighmax, lo, hi, &np))
      fail("device neighbour build failed");
    npairs = np;
  } else if (neighbor->ago == 0) {
    const int* type = atom->type;
    const int inum = list->inum;
    const int* ilist = list->ilist;
    const int* numneigh = list->numneigh;
    int** firstneigh = list->firstneigh;
    if ((size_t)ntotal > species.size()) species.resize((size_t)ntotal + ntotal / 2);
    for (int i = 0; i < ntotal; i++) species[i] = type[i] - 1;
    npairs = 0;
    for (int ii = 0; ii < inum; ii++) npairs += numneigh[ilist[ii]];
    if (use_fullnbr) {
      if ((size_t)npairs > jlist.size()) jlist.resize((size_t)npairs + npairs / 2);
      int64_t p = 0;
      for (int ii = 0; ii < inum; ii++) {
        const int i = ilist[ii];
        const int* js = firstneigh[i];
        const int jn = numneigh[i];
        for (int jj = 0; jj < jn; jj++) jlist[p++] = js[jj] & NEIGHMASK;
      }
      if (ani_set_neighbors_full(ctx, nlocal, ntotal, species.data(), inum, ilist, numneigh, jlist.data(), npairs, 0))
        fail("neighbour list upload failed");
    } else {
      if ((size_t)npairs * 2 > atom_index12.size()) atom_index12.resize((size_t)npairs * 2 + npairs);
      int64_t* I = atom_index12.data();
      int64_t* J = I + npairs;
      int64_t p = 0;
      for (int ii = 0; ii < inum; ii++) {
        const int i = ilist[ii];
        const int* js = firstneigh[i];
        const int jn = numneigh[i];
        for (int jj = 0; jj < jn; jj++, p++) { I[p] = i; J[p] = js[jj] & NEIGHMASK; }
      }
      if (ani_set_neighbors_half(ctx, nlocal, ntotal, species.data(), npairs, atom_index12.data(), 0))
        fail("neighbour list upload failed");
    }
  }

  double energy = 0.0, vir9[9];
  double* eat = nullptr;
  if (eflag_atom) {
    if ((size_t)nlocal > out_eatom.size()) out_eatom.resize((size_t)nlocal + nlocal / 2);
    eat = out_eatom.data();
  }
  const double* xflat = ntotal > 0 ? &x[0][0] : out_force.data();   // atom->x rows are contiguous (memory->create)
  if (ani_compute(ctx, xflat, eflag_global ? &energy : nullptr, out_force.data(), eat, vflag_either ? vir9 : nullptr))
    fail("evaluation failed");

  // newton off on both flags: Verlet does not reverse-communicate, so fold ghost forces here, on out_force and
  // not on f whose ghost entries are stale (pair_ani.cpp:192-201)
  if (!force->newton) comm->reverse_comm(this);

  // with newton off the ghost entries of f are never read again (their share was folded onto the owners just above)
  const int nf = force->newton ? ntotal : nlocal;
  for (int i = 0; i < nf; i++) {
    f[i][0] += out_force[3 * (size_t)i + 0];
    f[i][1] += out_force[3 * (size_t)i + 1];
    f[i][2] += out_force[3 * (size_t)i + 2];
  }
  if (eflag_global) eng_vdwl += energy;
  if (eflag_atom)
    for (int i = 0; i < nlocal; i++) eatom[i] += out_eatom[i];
  if (vflag_either) {
    virial[0] += vir9[0]; virial[1] += vir9[4]; virial[2] += vir9[8];
    virial[3] += vir9[1]; virial[4] += vir9[2]; virial[5] += vir9[5];
  }
}

/* restart record: same bytes as the reference (pair_ani.cpp:408-455): cutoff, use_num_models, the three flags as
 * `bool`, the two string sizes, the two strings */
void PairANIB200::write_restart(FILE* fp) {
  fwrite(&cutoff, sizeof(double), 1, fp);
  fwrite(&use_num_models, sizeof(int), 1, fp);
  fwrite(&use_cuaev, sizeof(bool), 1, fp);
  fwrite(&use_fullnbr, sizeof(bool), 1, fp);
  fwrite(&use_single, sizeof(bool), 1, fp);
  const int nm = (int)model_file.size(), nd = (int)device_str.size();
  fwrite(&nm, sizeof(int), 1, fp);
  fwrite(&nd, sizeof(int), 1, fp);
  fwrite(model_file.c_str(), sizeof(char), nm, fp);
  fwrite(device_str.c_str(), sizeof(char), nd, fp);
}

void PairANIB200::read_restart(FILE* fp) {
  utils::sfread(FLERR, &cutoff, sizeof(double), 1, fp, nullptr, error);
  utils::sfread(FLERR, &use_num_models, sizeof(int), 1, fp, nullptr, error);
  utils::sfread(FLERR, &use_cuaev, sizeof(bool), 1, fp, nullptr, error);
  utils::sfread(FLERR, &use_fullnbr, sizeof(bool), 1, fp, nullptr, error);
  utils::sfread(FLERR, &use_single, sizeof(bool), 1, fp, nullptr, error);
  int nm = 0, nd = 0;
  utils::sfread(FLERR, &nm, sizeof(int), 1, fp, nullptr, error);
  utils::sfread(FLERR, &nd, sizeof(int), 1, fp, nullptr, error);
  model_file.resize(nm);
  device_str.resize(nd);
  utils::sfread(FLERR, &model_file[0], sizeof(char), nm, fp, nullptr, error);
  utils::sfread(FLERR, &device_str[0], sizeof(char), nd, fp, nullptr, error);
  open_model();
}

int PairANIB200::pack_reverse_comm(int n, int first, double* buf) {
  memcpy(buf, out_force.data() + 3 * (size_t)first, (size_t)n * 3 * sizeof(double));
  return 3 * n;
}

void PairANIB200::unpack_reverse_comm(int n, int* idx, double* buf) {
  for (int i = 0; i < n; i++) {
    double* o = out_force.data() + 3 * (size_t)idx[i];
    o[0] += buf[3 * i + 0]; o[1] += buf[3 * i + 1]; o[2] += buf[3 * i + 2];
  }
}

/* ------------------------------------------------------------------------------------------------ ani/kk */

#ifdef LMP_KOKKOS
/* a Kokkos style as pair_ani_kokkos.cpp:38-46 declares itself: runs in the device execution space, reads x / type, adds to f
 * (and eatom / vatom there), so that VerletKokkos keeps the DualViews on the device around compute() */
PairANIB200Device::PairANIB200Device(LAMMPS* lmp) : PairANIB200(lmp) {
  respa_enable = 0;
  kokkosable = 1;
  atomKK = (AtomKokkos*)atom;
  execution_space = Device;
  datamask_read = X_MASK | F_MASK | TYPE_MASK | ENERGY_MASK | VIRIAL_MASK;
  datamask_modify = F_MASK | ENERGY_MASK | VIRIAL_MASK;
}
#else
PairANIB200Device::PairANIB200Device(LAMMPS* lmp) : PairANIB200(lmp) {}
#endif

void PairANIB200Device::init_style() {
#ifdef LMP_KOKKOS
  if (!use_fullnbr) error->all(FLERR, "Pair style ANI requires full neighbor list when using kokkos");
  neighbor->add_request(this, NeighConst::REQ_FULL);
  auto request = neighbor->find_request(this);
  request->set_kokkos_host(0);
  request->set_kokkos_device(1);
  if (lmp->kokkos->neighflag != FULL) error->all(FLERR, "Pair style ANI requires full neighbor list when using kokkos");
  if (force->newton_pair == 0) error->all(FLERR, "Pair style ANI requires newton pair on when using kokkos");
#else
  error->all(FLERR, "pair style ani/kk needs a LAMMPS built with the KOKKOS package (CUDA backend); use pair style ani");
#endif
}

#ifdef LMP_KOKKOS
/* Device-resident step (pair_ani_kokkos.cpp:67-214): x, f, type and the padded neighbour table are read where
 * Kokkos keeps them; the table is de-padded on the device only when the list was rebuilt. */
void PairANIB200Device::compute(int eflag, int vflag) {
  ev_init(eflag, vflag, 0);
  atomKK->sync(execution_space, X_MASK | F_MASK | TYPE_MASK);
  atomKK->modified(execution_space, F_MASK);
  auto x = atomKK->k_x.view<LMPDeviceType>();
  auto f = atomKK->k_f.view<LMPDeviceType>();
  auto type = atomKK->k_type.view<LMPDeviceType>();
  const int nlocal = atom->nlocal, ntotal = nlocal + atom->nghost;
  const int x_is_float = sizeof(typename decltype(x)::non_const_value_type) == 4;
  if (neighbor->ago == 0) {
    auto* kl = static_cast<NeighListKokkos<LMPDeviceType>*>(list);
    auto nb = kl->d_neighbors;
    const bool left = std::is_same<typename decltype(nb)::array_layout, Kokkos::LayoutLeft>::value;
    const int64_t rs = left ? 1 : (int64_t)nb.extent(1), cs = left ? (int64_t)nb.extent(0) : 1;
    if (ani_set_neighbors_table(ctx, nlocal, ntotal, type.data(), 1, list->inum, kl->d_ilist.data(), kl->d_numneigh.data(),
                                nb.data(), rs, cs, NEIGHMASK, 1))
      fail("neighbour table upload failed");
  }
  double energy = 0.0, vir9[9];
  double* eat = nullptr;
  if (eflag_atom) {
    memoryKK->destroy_kokkos(k_eatom, eatom);
    memoryKK->create_kokkos(k_eatom, eatom, maxeatom, "pair:eatom");
    eat = k_eatom.view<LMPDeviceType>().data();
  }
  if (ani_compute_device(ctx, x.data(), x_is_float, f.data(), x_is_float, eat, eflag_global ? &energy : nullptr,
                         vflag_either ? vir9 : nullptr))
    fail("evaluation failed");
  if (eflag_global) eng_vdwl += energy;
  if (vflag_either) {
    virial[0] += vir9[0]; virial[1] += vir9[4]; virial[2] += vir9[8];
    virial[3] += vir9[1]; virial[4] += vir9[2]; virial[5] += vir9[5];
  }
  if (eflag_atom) { k_eatom.template modify<LMPDeviceType>(); k_eatom.template sync<LMPHostType>(); }
}
#endif
