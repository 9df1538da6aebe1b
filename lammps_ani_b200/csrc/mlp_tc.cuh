// mlp_tc.cuh -- tcgen05 / TMEM / TMA ensemble MLP (placeholder interface; kernels land in the next milestone).
#pragma once
#include "common.cuh"

namespace anib {

struct TcSpeciesWeights { bool ready = false; };
struct TcWork {};
template <typename T> struct SpeciesNet;

inline int64_t tc_padded_rows(int nlocal, int S) { return (int64_t)nlocal + 128 * (S + 1); }

inline int tc_prepare_weights(const HostModel&, int, std::vector<SpeciesNet<float>>&) { return 0; }

inline int tc_mlp_species(cudaStream_t, bool, TcSpeciesWeights&, TcWork&, int, const std::vector<int>&, const float*,
                          const float*, int64_t, int, int, float*, double*, double, int*) {
  set_error("tensor-core MLP path is not built into this library yet");
  return 1;
}

}  // namespace anib
