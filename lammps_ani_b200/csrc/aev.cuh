// aev.cuh -- AEV forward (neighbour list -> AEV rows) and AEV backward / force + virial scatter kernels.
//
// Replaces the cuaev kernels the reference reaches through
// aev_computer._compute_cuaev_with_full_nbrlist (models/lammps_ani.py:277) and autograd's backward through
// them (lammps_ani.py:203-214): K1/K2/K3 and K8/K9/K10 of SURVEY.md 2.1.  Not a port: there is no global
// pair compaction, no CUB pass and no atomics in the forward.  One warp owns one centre atom:
//   * the atom's slice of the LAMMPS list is read coalesced, neighbour coordinates come as one 16/32-byte
//     vector load each (x,y,z,species packed), survivors of the Rcr / Rca filters are compacted into shared
//     memory by ballot;
//   * forward: lanes = the nA*nZ angular (ShfA,ShfZ) terms, the warp walks the triples serially and
//     accumulates the AEV row in shared memory with lane-private addresses, then stores it with one
//     coalesced pass (optionally also as TF32 hi/lo planes for the tensor-core MLP);
//   * backward: lanes = neighbours; each lane owns the gradient of its own neighbour, so no cross-lane
//     reduction is needed inside the triple loop; forces leave through fp32/fp64 red.global.add.
// cos(theta - ShfZ) is expanded as cos(theta)cos(ShfZ) + sin(theta)sin(ShfZ) with cos(theta) = 0.95 cos_jik,
// which is the same function as the reference's acos/cos pair without the two transcendental calls.
#pragma once
#include <cuda_fp16.h>

#include "common.cuh"

namespace anib {

template <typename T> struct Vec4T;
template <> struct Vec4T<float> { using type = float4; };
template <> struct Vec4T<double> { using type = double4; };

__device__ __forceinline__ float pack_species(int s, float) { return __int_as_float(s); }
__device__ __forceinline__ double pack_species(int s, double) { return (double)s; }
__device__ __forceinline__ int unpack_species(float w) { return __float_as_int(w); }
__device__ __forceinline__ int unpack_species(double w) { return (int)w; }

// ---- arithmetic helpers: fp32 uses the SFU approximations (ex2/lg2/sin/cos), fp64 the accurate library.
__device__ __forceinline__ float t_sqrt(float x) { return sqrtf(x); }
__device__ __forceinline__ double t_sqrt(double x) { return sqrt(x); }
__device__ __forceinline__ float t_exp(float x) { return exp2f(x * 1.4426950408889634f); }
__device__ __forceinline__ double t_exp(double x) { return exp(x); }
__device__ __forceinline__ void t_sincos(float x, float* s, float* c) { __sincosf(x, s, c); }
__device__ __forceinline__ void t_sincos(double x, double* s, double* c) { sincos(x, s, c); }
__device__ __forceinline__ float t_cos(float x) { return __cosf(x); }
__device__ __forceinline__ double t_cos(double x) { return cos(x); }
// u^p for u >= 0
__device__ __forceinline__ float t_pow(float u, float p) { return exp2f(p * __log2f(u)); }
__device__ __forceinline__ double t_pow(double u, double p) { return u > 0.0 ? pow(u, p) : 0.0; }

// exp(-eta d^2) with the constant k = -eta log2(e) folded on the caller's side (fp32: two multiplies and one ex2)
__device__ __forceinline__ float t_gauss(float d, float k, float) { return exp2f(d * d * k); }
__device__ __forceinline__ double t_gauss(double d, double, double eta) { return exp(-eta * d * d); }

// round-to-nearest TF32 (10-bit mantissa) kept in an fp32 container: exactly representable for tcgen05 kind::tf32
__device__ __forceinline__ float round_tf32(float f) {
  uint32_t u;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(u) : "f"(f));
  return __uint_as_float(u);
}

template <typename T>
__global__ void pack_x_kernel(const void* __restrict__ x_in, int in_is_float, const int* __restrict__ species, int ntotal,
                              typename Vec4T<T>::type* __restrict__ x4) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= ntotal) return;
  typename Vec4T<T>::type v;
  if (in_is_float) {
    const float* p = (const float*)x_in + 3 * (size_t)i;
    v.x = (T)p[0]; v.y = (T)p[1]; v.z = (T)p[2];
  } else {
    const double* p = (const double*)x_in + 3 * (size_t)i;
    v.x = (T)p[0]; v.y = (T)p[1]; v.z = (T)p[2];   // single mode: same rounding as `.to(dtype)` (ani.cpp:206-209)
  }
  v.w = pack_species(species[i], T(0));
  x4[i] = v;
}

struct AevSmemLayout {
  // per-warp byte offsets
  int off_d, off_s, off_j, off_g, off_a, off_afc, off_adfc, off_as, off_ai, off_acc, off_triu, bytes;
};

// Shared-memory layout shared by host and device.  cap = neighbour capacity (>= max numneigh), capA = angular cap.
template <typename T>
__host__ __device__ inline AevSmemLayout aev_smem_layout(int cap, int capA, int acc_elems, bool backward) {
  AevSmemLayout L;
  int o = 0;
  auto take = [&](int bytes, int align) { o = (o + align - 1) / align * align; int r = o; o += bytes; return r; };
  L.off_d = take(cap * 4 * (int)sizeof(T), 4 * (int)sizeof(T));
  L.off_a = take(capA * 4 * (int)sizeof(T), 4 * (int)sizeof(T));
  L.off_acc = take(acc_elems * (int)sizeof(T), 16);
  L.off_afc = take(capA * (int)sizeof(T), 8);
  L.off_adfc = backward ? take(capA * (int)sizeof(T), 8) : 0;
  L.off_g = backward ? take(cap * 3 * (int)sizeof(T), 8) : 0;
  L.off_s = take(cap * 4, 4);
  L.off_j = backward ? take(cap * 4, 4) : 0;
  L.off_as = take(capA * 4, 4);
  L.off_ai = backward ? take(capA * 4, 4) : 0;
  L.off_triu = take(kMaxSpecies * kMaxSpecies * 4, 4);
  L.bytes = (o + 31) / 32 * 32;
  return L;
}

constexpr int kAevWarps = 4;  // warps (= centre atoms) per CTA

// Gather + filter the neighbour slice of centre `i` into shared memory.  Returns (n, na) to all lanes.
template <typename T, bool BWD>
__device__ __forceinline__ void aev_gather(const AevConst<T>& P, bool radial_filter,
                                           const typename Vec4T<T>::type* __restrict__ x4,
                                           const int* __restrict__ jl, int cnt, typename Vec4T<T>::type xi, int lane,
                                           typename Vec4T<T>::type* sh_d, int* sh_s, int* sh_j,
                                           typename Vec4T<T>::type* sh_a, T* sh_afc, T* sh_adfc, int* sh_as, int* sh_ai,
                                           int capA, int& n_out, int& na_out) {
  using V4 = typename Vec4T<T>::type;
  const unsigned lt = (1u << lane) - 1u;
  int n = 0, na = 0;
  for (int base = 0; base < cnt; base += 32) {
    int e = base + lane;
    bool valid = e < cnt;
    int j = valid ? __ldg(jl + e) : 0;
    V4 p = xi;
    if (valid) p = x4[j];
    T dx = p.x - xi.x, dy = p.y - xi.y, dz = p.z - xi.z;
    T r2 = dx * dx + dy * dy + dz * dz;
    bool keep = valid && (!radial_filter || r2 <= P.Rcr2) && r2 > T(0);
    bool keepa = keep && r2 <= P.Rca2;
    if (!radial_filter) keep = keep || keepa;
    unsigned m = __ballot_sync(0xffffffffu, keep);
    unsigned ma = __ballot_sync(0xffffffffu, keepa);
    int pos = n + __popc(m & lt);
    int posa = na + __popc(ma & lt);
    if (keep) {
      T r = t_sqrt(r2);
      V4 d; d.x = dx; d.y = dy; d.z = dz; d.w = r;
      sh_d[pos] = d;
      int s = unpack_species(p.w);
      sh_s[pos] = s;
      if (BWD) sh_j[pos] = j;
      if (keepa && posa < capA) {
        T inv = T(1) / r;
        V4 u; u.x = dx * inv; u.y = dy * inv; u.z = dz * inv; u.w = r;
        sh_a[posa] = u;
        T sn, cs;
        t_sincos(P.pi_over_Rca * r, &sn, &cs);
        sh_afc[posa] = T(0.5) * cs + T(0.5);
        if (BWD) { sh_adfc[posa] = T(-0.5) * P.pi_over_Rca * sn; sh_ai[posa] = pos; }
        sh_as[posa] = s;
      }
    }
    n += __popc(m);
    na += __popc(ma);
  }
  n_out = n;
  na_out = na;
}

// ------------------------------------------------------------------------------------------------ forward
// aev rows are written in BUCKET order: row `pos` belongs to atom order[pos]; ld = row pitch in elements.
// If aev_hi != nullptr (fp32 only) the row is also written as TF32 hi / lo planes for the tcgen05 MLP.
template <typename T>
__global__ void __launch_bounds__(kAevWarps * 32)
aev_forward_kernel(const __grid_constant__ AevConst<T> P, int radial_filter,
                   const typename Vec4T<T>::type* __restrict__ x4, const int64_t* __restrict__ nbr_beg,
                   const int* __restrict__ nbr_cnt, const int* __restrict__ jlist, const int* __restrict__ order,
                   int nrows, T* __restrict__ aev,
                   int64_t ld, void* __restrict__ aev_hi, void* __restrict__ aev_lo, int64_t ld_planes, int plane_mode,
                   float plane_scale, int cap, int capA, int smem_per_warp, int* __restrict__ errflag) {
  using V4 = typename Vec4T<T>::type;
  extern __shared__ __align__(32) unsigned char smem_raw[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row = blockIdx.x * kAevWarps + warp;
  if (row >= nrows) return;
  const int nRpad = P.nR <= 16 ? 16 : 32;   // lanes per neighbour in the radial pass
  const int npl = 32 / nRpad;               // neighbours per warp iteration
  const int acc_elems = P.aev_len + (npl - 1) * P.rad_len;
  const AevSmemLayout L = aev_smem_layout<T>(cap, capA, acc_elems, false);
  unsigned char* base = smem_raw + (size_t)warp * smem_per_warp;
  V4* sh_d = (V4*)(base + L.off_d);
  V4* sh_a = (V4*)(base + L.off_a);
  T* acc = (T*)(base + L.off_acc);
  T* sh_afc = (T*)(base + L.off_afc);
  int* sh_s = (int*)(base + L.off_s);
  int* sh_as = (int*)(base + L.off_as);

  const int i = order[row];
  const V4 xi = x4[i];
  const int64_t beg = nbr_beg[i];
  const int cnt = nbr_cnt[i];
  for (int t = lane; t < acc_elems; t += 32) acc[t] = T(0);
  int n, na;
  aev_gather<T, false>(P, radial_filter != 0, x4, jlist + beg, cnt, xi, lane, sh_d, sh_s, nullptr, sh_a, sh_afc, nullptr,
                       sh_as, nullptr, capA, n, na);
  if (na > capA) { if (lane == 0) atomicExch(errflag, 1); na = capA; }
  __syncwarp();

  // ---- radial: lane = (sub-neighbour, shift k)
  {
    const int k = lane % nRpad, sub = lane / nRpad;
    const T shf = P.ShfR[k < P.nR ? k : 0];
    T* racc = (sub == 0) ? acc : acc + P.aev_len + (sub - 1) * P.rad_len;
    for (int e0 = 0; e0 < n; e0 += npl) {
      int e = e0 + sub;
      if (e < n && k < P.nR) {
        T r = sh_d[e].w;
        T fc = T(0.5) * t_cos(P.pi_over_Rcr * r) + T(0.5);
        T dr = r - shf;
        racc[sh_s[e] * P.nR + k] += T(0.25) * t_exp(-P.EtaR * dr * dr) * fc;
      }
    }
  }
  // ---- angular: lane = (ShfA index ia, ShfZ index iz); triples walked serially, lane-private accumulators
  if (lane < P.ang_sub) {
    const int ia = lane / P.nZ, iz = lane % P.nZ;
    const T shfA = P.ShfA[ia], cz = P.cosZ[iz], sz = P.sinZ[iz];
    T* aacc = acc + P.rad_len + lane;
    for (int a = 0; a + 1 < na; a++) {
      const V4 A = sh_a[a];
      const T fca2 = T(2) * sh_afc[a];
      const int* trow = P.triu[sh_as[a]];
      for (int b = a + 1; b < na; b++) {
        const V4 B = sh_a[b];
        T c = T(0.95) * (A.x * B.x + A.y * B.y + A.z * B.z);
        T sn = t_sqrt(fmax(T(0), T(1) - c * c));
        T u = fmax(T(0), T(0.5) + T(0.5) * (c * cz + sn * sz));
        T dr = T(0.5) * (A.w + B.w) - shfA;
        T term = fca2 * sh_afc[b] * t_pow(u, P.Zeta) * t_exp(-P.EtaA * dr * dr);
        aacc[trow[sh_as[b]] * P.ang_sub] += term;
      }
    }
  }
  __syncwarp();
  // ---- one coalesced store of the row: fp32/fp64 row (aev != nullptr) and / or operand planes for the tensor-core MLP
  // (plane_mode 1: TF32 hi/lo in fp32 containers; 2: fp16 hi/lo of plane_scale * value)
  T* out = aev ? aev + (size_t)row * ld : nullptr;
  for (int t = lane; t < P.aev_len; t += 32) {
    T v = acc[t];
    if (t < P.rad_len)
      for (int s2 = 1; s2 < npl; s2++) v += acc[P.aev_len + (s2 - 1) * P.rad_len + t];
    if (out) out[t] = v;
    if (plane_mode == 1) {
      float f = (float)v;
      float hi = round_tf32(f);
      ((float*)aev_hi)[(size_t)row * ld_planes + t] = hi;
      ((float*)aev_lo)[(size_t)row * ld_planes + t] = round_tf32(f - hi);
    } else if (plane_mode == 2) {
      const float f = (float)v * plane_scale;
      const __half hi = __float2half_rn(f);
      ((__half*)aev_hi)[(size_t)row * ld_planes + t] = hi;
      ((__half*)aev_lo)[(size_t)row * ld_planes + t] = __float2half_rn(f - __half2float(hi));
    }
  }
}

// ------------------------------------------------------------------------------------------------ backward
// daev: dE/dAEV rows in bucket order (Hartree).  force: T[ntotal*3], accumulated with atomics (kcal/mol/A).
// virial_slots: double[kVirialSlots*9], accumulated (kcal/mol), or nullptr.
constexpr int kVirialSlots = 128;

template <typename T> __device__ __forceinline__ void atomic_add_t(T* p, T v) { atomicAdd(p, v); }

template <typename T, int NZ>
__global__ void __launch_bounds__(kAevWarps * 32)
aev_backward_kernel(const __grid_constant__ AevConst<T> P, int radial_filter,
                    const typename Vec4T<T>::type* __restrict__ x4, const int64_t* __restrict__ nbr_beg,
                    const int* __restrict__ nbr_cnt, const int* __restrict__ jlist, const int* __restrict__ order,
                    int nrows, const T* __restrict__ daev, int64_t ld, T* __restrict__ force, double* __restrict__ virial_slots,
                    T out_scale, int cap, int capA, int smem_per_warp) {
  using V4 = typename Vec4T<T>::type;
  extern __shared__ __align__(32) unsigned char smem_raw[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row = blockIdx.x * kAevWarps + warp;
  if (row >= nrows) return;
  const int rstride = P.nR + 1;            // padded radial rows (bank spread across species)
  const int astride = P.ang_sub + 1;       // padded angular rows
  const int acc_elems = P.S * rstride + P.npairs * astride;
  const AevSmemLayout L = aev_smem_layout<T>(cap, capA, acc_elems, true);
  unsigned char* base = smem_raw + (size_t)warp * smem_per_warp;
  V4* sh_d = (V4*)(base + L.off_d);
  V4* sh_a = (V4*)(base + L.off_a);
  T* Arad = (T*)(base + L.off_acc);
  T* Aang = Arad + P.S * rstride;
  T* sh_afc = (T*)(base + L.off_afc);
  T* sh_adfc = (T*)(base + L.off_adfc);
  T* sh_g = (T*)(base + L.off_g);
  int* sh_s = (int*)(base + L.off_s);
  int* sh_j = (int*)(base + L.off_j);
  int* sh_as = (int*)(base + L.off_as);
  int* sh_ai = (int*)(base + L.off_ai);
  int* sh_triu = (int*)(base + L.off_triu);

  const int i = order[row];
  const V4 xi = x4[i];
  const int64_t beg = nbr_beg[i];
  const int cnt = nbr_cnt[i];
  // upstream gradient row -> shared (coalesced read, padded rows)
  const T* g_in = daev + (size_t)row * ld;
  for (int t = lane; t < P.aev_len; t += 32) {
    T v = g_in[t];
    if (t < P.rad_len) Arad[(t / P.nR) * rstride + (t % P.nR)] = v;
    else { int q = t - P.rad_len; Aang[(q / P.ang_sub) * astride + (q % P.ang_sub)] = v; }
  }
  for (int t = lane; t < kMaxSpecies * kMaxSpecies; t += 32) sh_triu[t] = P.triu[t / kMaxSpecies][t % kMaxSpecies];
  int n, na;
  aev_gather<T, true>(P, radial_filter != 0, x4, jlist + beg, cnt, xi, lane, sh_d, sh_s, sh_j, sh_a, sh_afc, sh_adfc,
                      sh_as, sh_ai, capA, n, na);
  if (na > capA) na = capA;   // the forward pass raised the error flag
  __syncwarp();

  // ---- radial: lane = neighbour
  for (int e = lane; e < n; e += 32) {
    const V4 d = sh_d[e];
    const T r = d.w;
    T sn, cs;
    t_sincos(P.pi_over_Rcr * r, &sn, &cs);
    const T fc = T(0.5) * cs + T(0.5), dfc = T(-0.5) * P.pi_over_Rcr * sn;
    const T* Ar = Arad + sh_s[e] * rstride;
    T acc = T(0);
    for (int k = 0; k < P.nR; k++) {
      T dr = r - P.ShfR[k];
      acc += Ar[k] * t_exp(-P.EtaR * dr * dr) * (dfc - T(2) * P.EtaR * dr * fc);
    }
    const T coef = T(0.25) * acc / r;
    sh_g[3 * e + 0] = coef * d.x; sh_g[3 * e + 1] = coef * d.y; sh_g[3 * e + 2] = coef * d.z;
  }
  __syncwarp();
  // ---- angular: lane = neighbour b (chunks of 32); loop over all a != b; lane accumulates its own gradient
  for (int b0 = 0; b0 < na; b0 += 32) {
    const int b = b0 + lane;
    const bool has = b < na;
    V4 B; B.x = B.y = B.z = T(0); B.w = T(1);
    T fcb = T(0), dfcb = T(0);
    int sb = 0;
    if (has) { B = sh_a[b]; fcb = sh_afc[b]; dfcb = sh_adfc[b]; sb = sh_as[b]; }
    T gx = T(0), gy = T(0), gz = T(0);
    for (int a = 0; a < na; a++) {
      if (!has || a == b) continue;
      const V4 A = sh_a[a];
      const T fca = sh_afc[a];
      const T cosv = A.x * B.x + A.y * B.y + A.z * B.z;
      const T c = T(0.95) * cosv;
      const T sn = t_sqrt(fmax(T(0), T(1) - c * c));
      const T c_over_sn = c / sn;
      const T* w = Aang + sh_triu[sh_as[a] * kMaxSpecies + sb] * astride;
      T f1[NZ], df1[NZ];
#pragma unroll
      for (int z = 0; z < NZ; z++) {
        T u = fmax(T(0), T(0.5) + T(0.5) * (c * P.cosZ[z] + sn * P.sinZ[z]));
        T um1 = t_pow(u, P.Zeta - T(1));                       // u^(Zeta-1)
        f1[z] = um1 * u;
        df1[z] = P.Zeta * um1 * T(0.5) * (P.cosZ[z] - c_over_sn * P.sinZ[z]);   // d f1 / d c
      }
      const T ravg = T(0.5) * (A.w + B.w);
      T s_rb = T(0), s_cos = T(0);
      for (int ia = 0; ia < P.nA; ia++) {
        T dr = ravg - P.ShfA[ia];
        T f2 = t_exp(-P.EtaA * dr * dr);
        T df2 = T(-2) * P.EtaA * dr * f2;
        T Pz = T(0), Qz = T(0);
#pragma unroll
        for (int z = 0; z < NZ; z++) {
          T wv = w[ia * NZ + z];
          Pz += wv * f1[z];
          Qz += wv * df1[z];
        }
        s_rb += Pz * (T(0.5) * df2 * fcb + f2 * dfcb);
        s_cos += Qz * f2;
      }
      const T dE_drb = T(2) * fca * s_rb;
      const T dE_dcos = T(2) * fca * fcb * T(0.95) * s_cos / B.w;
      gx += dE_drb * B.x + dE_dcos * (A.x - cosv * B.x);
      gy += dE_drb * B.y + dE_dcos * (A.y - cosv * B.y);
      gz += dE_drb * B.z + dE_dcos * (A.z - cosv * B.z);
    }
    if (has) {
      int e = sh_ai[b];
      sh_g[3 * e + 0] += gx; sh_g[3 * e + 1] += gy; sh_g[3 * e + 2] += gz;
    }
  }
  __syncwarp();
  // ---- scatter: F_j -= g, F_i += sum g, virial -= g (x) d
  T fix = T(0), fiy = T(0), fiz = T(0);
  double vir[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
  for (int e = lane; e < n; e += 32) {
    const T gx = sh_g[3 * e + 0] * out_scale, gy = sh_g[3 * e + 1] * out_scale, gz = sh_g[3 * e + 2] * out_scale;
    const int j = sh_j[e];
    atomic_add_t(force + 3 * (size_t)j + 0, -gx);
    atomic_add_t(force + 3 * (size_t)j + 1, -gy);
    atomic_add_t(force + 3 * (size_t)j + 2, -gz);
    fix += gx; fiy += gy; fiz += gz;
    if (virial_slots != nullptr) {
      const V4 d = sh_d[e];
      vir[0] -= (double)(gx * d.x); vir[1] -= (double)(gx * d.y); vir[2] -= (double)(gx * d.z);
      vir[3] -= (double)(gy * d.x); vir[4] -= (double)(gy * d.y); vir[5] -= (double)(gy * d.z);
      vir[6] -= (double)(gz * d.x); vir[7] -= (double)(gz * d.y); vir[8] -= (double)(gz * d.z);
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    fix += __shfl_xor_sync(0xffffffffu, fix, o);
    fiy += __shfl_xor_sync(0xffffffffu, fiy, o);
    fiz += __shfl_xor_sync(0xffffffffu, fiz, o);
  }
  if (lane == 0) {
    atomic_add_t(force + 3 * (size_t)i + 0, fix);
    atomic_add_t(force + 3 * (size_t)i + 1, fiy);
    atomic_add_t(force + 3 * (size_t)i + 2, fiz);
  }
  if (virial_slots != nullptr) {
#pragma unroll
    for (int q = 0; q < 9; q++) {
      double v = vir[q];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      if (lane == 0) atomicAdd(virial_slots + (size_t)(blockIdx.x % kVirialSlots) * 9 + q, v);
    }
  }
}

}  // namespace anib
