// common.cuh -- shared declarations of libani_b200 (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <string>
#include <vector>

namespace anib {

constexpr int kMaxSpecies = 8;
constexpr int kMaxShfR = 32;
constexpr int kMaxShfA = 16;
constexpr int kMaxShfZ = 16;
constexpr int kMaxLayers = 6;
constexpr double kHartree2KcalMol = 627.5094738898777;  // reference src/ani_csrc/ani.h:9
constexpr double kCeluAlpha = 0.1;

// Host copy of one `.aniw` file (see lammps_ani_b200/aniw.py for the layout).
struct HostModel {
  int S = 0, nR = 0, nA = 0, nZ = 0, M = 0, use_repulsion = 0;
  double Rcr = 0, Rca = 0, EtaR = 0, EtaA = 0, Zeta = 0;
  std::vector<double> ShfR, ShfA, ShfZ, sae;
  std::vector<std::string> symbols;
  std::vector<std::vector<int>> dims;  // per species: in, h1, ..., 1
  // W[m][s][l] row-major [out][in]; b[m][s][l]
  std::vector<std::vector<std::vector<std::vector<float>>>> W, b;
  // RepulsionXTB pair term (use_repulsion): E = sum_pairs y_ab / r * exp(-sqrt_alpha_ab r^k_ab) * fc(r), r in Bohr, E in Hartree
  double rep_cutoff = 0.0;
  int rep_cutoff_fn = 1;               // 0 cosine, 1 smooth (order 2)
  std::vector<double> rep_y, rep_sa, rep_k;   // [S*S]
  int rad_len() const { return S * nR; }
  int ang_sub() const { return nA * nZ; }
  int aev_len() const { return S * nR + (S * (S + 1) / 2) * nA * nZ; }
};

bool load_aniw(const std::string& path, HostModel& out, std::string& err);

// AEV constants in the arithmetic type of the kernels.
template <typename T>
struct AevConst {
  T Rcr, Rca, Rcr2, Rca2, EtaR, EtaA, Zeta;
  T pi_over_Rcr, pi_over_Rca;
  T ShfR[kMaxShfR];
  T ShfA[kMaxShfA];
  T cosZ[kMaxShfZ], sinZ[kMaxShfZ];
  int S, nR, nA, nZ, rad_len, ang_sub, aev_len, npairs;
  int triu[kMaxSpecies][kMaxSpecies];
};

void set_error(const std::string& msg);

#define ANIB_CUDA_OK(expr)                                                                       \
  do {                                                                                           \
    cudaError_t _e = (expr);                                                                     \
    if (_e != cudaSuccess) {                                                                     \
      ::anib::set_error(std::string(#expr) + " failed: " + cudaGetErrorString(_e) + " at " +     \
                        __FILE__ + ":" + std::to_string(__LINE__));                              \
      return 1;                                                                                  \
    }                                                                                            \
  } while (0)

#ifdef __CUDACC__
// Packed fp32 pairs (sm_100: FFMA2 / FMUL2 on a 64-bit register pair, one issue slot for two lanes of arithmetic; the AEV
// kernels and the MLP epilogues are bound by instruction issue).  An operand built with pk2(w, w) is issued in the scalar-broadcast form.
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pk2(float a, float b) {
  f32x2 r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b));
  return r;
}
__device__ __forceinline__ void upk2(f32x2 r, float& a, float& b) { asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(r)); }
__device__ __forceinline__ f32x2 ffma2(f32x2 a, f32x2 b, f32x2 c) {
  f32x2 d;
  asm("fma.rn.ftz.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
__device__ __forceinline__ f32x2 pk2u(uint32_t a, uint32_t b) {
  f32x2 r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "r"(a), "r"(b));
  return r;
}
__device__ __forceinline__ f32x2 fmul2(f32x2 a, f32x2 b) {
  f32x2 d;
  asm("mul.rn.ftz.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
#endif

}  // namespace anib
