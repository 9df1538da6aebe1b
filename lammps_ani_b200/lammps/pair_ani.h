/* pair_ani.h -- `pair_style ani` / `ani/kk` for LAMMPS on top of libani_b200 (C ABI, include/ani_b200.h).
 *
 * Same user surface as the reference pair style (src/pair_ani.h:22-57, src/pair_ani.cpp:285-341):
 *   pair_style ani <cutoff> <model_file> <cpu|cuda> [num_models] [cuaev|pyaev] [full|half] [single|double]
 *   pair_coeff * *
 * The class links only libani_b200 and cudart-free LAMMPS headers: no libtorch, no Kokkos.  `device cpu` is
 * rejected (there is no CPU path in this build).
 */
#ifndef LMP_PAIR_ANI_B200_H
#define LMP_PAIR_ANI_B200_H

#include <cstdint>
#include <string>
#include <vector>

#include "ani_b200.h"
#include "pair.h"
#ifdef LMP_KOKKOS
#include "kokkos_type.h"
#endif

namespace LAMMPS_NS {

class PairANIB200 : public Pair {
 public:
  PairANIB200(class LAMMPS*);
  ~PairANIB200() override;
  void compute(int, int) override;
  void settings(int, char**) override;
  void coeff(int, char**) override;
  void init_style() override;
  double init_one(int, int) override;
  void* extract(const char*, int&) override;
  void write_restart(FILE*) override;
  void read_restart(FILE*) override;
  int pack_reverse_comm(int, int, double*) override;
  void unpack_reverse_comm(int, int*, double*) override;
  double memory_usage() override;

 protected:
  double cutoff;
  std::string model_file, device_str;
  int use_num_models;
  bool use_cuaev, use_fullnbr, use_single;
  int last_nmax = -1;
  bool device_list = false;   // `ani_neighbor device`: the library builds the list on the GPU, none is requested from LAMMPS
  ani_model* model;
  ani_ctx* ctx;
  // flattened neighbour data of the last rebuild (grow-only, x1.5 like pair_ani.cpp:120-127)
  std::vector<int> species, jlist;
  std::vector<int64_t> atom_index12;
  int64_t npairs;
  std::vector<double> out_force, out_eatom;

  // ---- device-resident mode, switched on by `fix ani/md`: x, f (and the fix's v) live on the GPU between
  // neighbour rebuilds; ghost positions are regenerated from their owners and ghost forces folded on the device; the host
  // arrays are refreshed only when LAMMPS needs them (rebuild checks, thermo / dump steps)
 public:
  struct Resident {
    bool on = false, valid = false;     // requested by a fix / device arrays hold the current state
    double *x = nullptr, *f = nullptr, *shift = nullptr;
    double* x0 = nullptr;               // local positions at the last neighbour build (device-side displacement check)
    int cap_x0 = 0;
    int* owner = nullptr;
    int cap_total = 0, cap_ghost = 0, nlocal = 0, nghost = 0;
    bigint rebuilt_step = -1;           // timestep of the last upload after a rebuild (the fix re-uploads v when it changes)
  } res;
  ani_ctx* context() { return ctx; }
  void set_resident(bool on);
  void download_x();                    // device -> atom->x (local atoms)
  void download_f();                    // device -> atom->f (local atoms, overwrite)
  bool moved_half_skin();               // Neighbor::check_distance on the device arrays: true = LAMMPS will want to rebuild

 protected:
  void resident_compute(int eflag, int vflag);
  void resident_free();
  std::vector<int> h_owner, h_tag2local;
  std::vector<double> h_shift;

  virtual void allocate();
  int local_device(const std::string& device);
  void open_model();
  void close_model();
  void fail(const char* what);
};

/* Device-resident variant registered as `ani/kk`: positions go to the GPU once per step as one pinned transfer
 * and forces come back the same way when LAMMPS is built without KOKKOS; with -DLMP_KOKKOS the Kokkos views'
 * device pointers are handed to ani_compute_device directly (pair_ani_kokkos.cpp:67-214). */
class PairANIB200Device : public PairANIB200 {
 public:
  PairANIB200Device(class LAMMPS* lmp);
  void init_style() override;
#ifdef LMP_KOKKOS
  void compute(int, int) override;

 protected:
  DAT::tdual_efloat_1d k_eatom;
#endif
};

}    // namespace LAMMPS_NS
#endif
