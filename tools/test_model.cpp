// test_model -- stand-alone known-answer driver over the C ABI of libani_b200 (include/ani_b200.h), no LAMMPS, no Python.
//
//     test_model <model.aniw> <cpu|cuda> <single|double>
//
// Same command line, inputs, call and thresholds as the reference's src/ani_csrc/test_model.cpp:64-174 (the cheapest
// end-to-end oracle of SURVEY.md 3.5): the 30-atom water cluster, its hard-coded half pair list, `ANI` constructed with
// use_cuaev = false, use_fullnbr = false (test_model.cpp:83), one half-list compute (test_model.cpp:151), energy error in
// Hartree and force error in kcal/mol/A against 1e-8 (double) / 3e-4 (single) (test_model.cpp:164).  The expectations are
// chosen by the model that was loaded: ANI-2x (1008-long AEV) -> the reference's own numbers (test_model.cpp:121-137);
// ANI-1xnr (448) -> the CPU oracle's fp64 result on the same inputs (tests/golden/water30_ani1xnr.json).  `cpu` fails:
// libani_b200 has no CPU path.  The model is a .aniw blob (python -m lammps_ani_b200.extract <ani2x.pt|*.info> out.aniw).
#include <cmath>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "../include/ani_b200.h"
#include "test_model_golden.h"

static const double hartree2kcalmol = 627.5094738898777;   // src/ani_csrc/ani.h:9

static int fail(const char* what) {
  fprintf(stderr, "Error: %s: %s\n", what, ani_last_error());
  return 1;
}

int main(int argc, const char* argv[]) {
  if (argc != 4) {
    fprintf(stderr, "Error: usage is test_model <model.aniw> <cpu/cuda> <single/double>\n");
    return 255;
  }
  const std::string model_file = argv[1], device_str = argv[2], precision = argv[3];
  if (device_str != "cpu" && device_str != "cuda") { fprintf(stderr, "Error: 2nd argument must be <cpu/cuda>\n"); return 255; }
  if (precision != "single" && precision != "double") { fprintf(stderr, "3rd argument must be <single/double>\n"); return 255; }
  const int local_rank = device_str == "cpu" ? -1 : 0;
  const bool single = precision == "single";

  ani_model* model = nullptr;
  ani_ctx* ctx = nullptr;
  if (ani_model_load(model_file.c_str(), &model)) return fail("cannot load the model");
  int S = 0, L = 0, M = 0;
  double Rcr = 0, Rca = 0;
  ani_model_info(model, &S, &L, &M, &Rcr, &Rca);
  printf("model %s: %d species, AEV length %d, %d members, Rcr %.2f Rca %.2f\n", model_file.c_str(), S, L, M, Rcr, Rca);
  if (ani_ctx_create(model, local_rank, -1, /*use_cuaev=*/0, /*use_fullnbr=*/0, single ? 1 : 0, &ctx)) return fail("cannot create the context");

  const int nlocal = 30, ntotal = 30;
  const int64_t npairs = (int64_t)(sizeof(kAtomIndex12) / sizeof(kAtomIndex12[0]) / 2);
  const double* expected_force;
  double expected_energy;
  if (L == 1008) { expected_force = kAni2xForce; expected_energy = kAni2xEnergyHartree * hartree2kcalmol; }
  else if (L == 448) { expected_force = kAni1xnrForce; expected_energy = kAni1xnrEnergyKcalmol; }
  else { fprintf(stderr, "Error: no known answers for a model with AEV length %d\n", L); return 1; }

  double worst_e = 0, worst_f = 0;
  for (int call = 0; call < 2; call++) {      // second call: cached neighbour data, as in an MD step with ago != 0
    if (call == 0 && ani_set_neighbors_half(ctx, nlocal, ntotal, kSpecies, npairs, kAtomIndex12, 0)) return fail("neighbour list upload failed");
    double out_energy = 0;
    std::vector<double> out_force((size_t)ntotal * 3, 0.0);
    if (ani_compute(ctx, kCoords, &out_energy, out_force.data(), nullptr, nullptr)) return fail("evaluation failed");
    const double energy_err = std::fabs(out_energy - expected_energy) / hartree2kcalmol;   // Hartree
    double force_err = 0;
    for (size_t i = 0; i < out_force.size(); i++) force_err = std::fmax(force_err, std::fabs(expected_force[i] - out_force[i]));
    printf("%s call : energy %.10f (kcal/mol), error: %.3e (hartree)\n", call ? "Second" : "First ", out_energy, energy_err);
    printf("%s call : force %.10e, %.10e, %.10e, error: %.3e (kcal/mol/A)\n", call ? "Second" : "First ", out_force[0], out_force[1], out_force[2],
           force_err);
    worst_e = std::fmax(worst_e, energy_err);
    worst_f = std::fmax(worst_f, force_err);
  }
  const double threshold = single ? 3e-4 : 1e-8;
  ani_ctx_destroy(ctx);
  ani_model_free(model);
  if (!(worst_e < threshold)) { fprintf(stderr, "Wrong Energy\n"); return 2; }
  if (!(worst_f < threshold)) { fprintf(stderr, "Wrong Forces\n"); return 3; }
  printf("OK\n");
  return 0;
}
