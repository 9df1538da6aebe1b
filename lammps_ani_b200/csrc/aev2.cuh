// aev2.cuh -- second-generation AEV forward / backward kernels (one warp per centre atom, lanes = angular TRIPLES).
//
// Same contract as the kernels in aev.cuh (which stay as the reference implementation inside the library and are
// selectable with ANI_B200_AEV=1); what changes is where the transcendental work is done.  An angular term factorises
//     2 fc(r_a) fc(r_b) * f1_z(theta_ab) * f2_a((r_a + r_b)/2),      f1_z = ((1 + cos(theta - ShfZ_z))/2)^Zeta,
//                                                                    f2_a = exp(-EtaA (ravg - ShfA_a)^2)
// so a triple needs nZ powers and nA exponentials, not nZ*nA of each.  Forward: a batch of 32 triples is evaluated with
// one triple per lane (stage 1: f1*fc and f2 into shared memory), then the nA*nZ products are accumulated with one
// (ShfA,ShfZ) term per lane in registers (stage 2), flushing to the shared AEV row only when the species pair changes.
// Backward: one triple per lane computes both neighbours' gradients at once from the factorised sums
//     S0 = sum w f1 f2, S1 = sum w f1 f2', S2 = sum w f1' f2        (w = dE/dAEV of the triple's species pair)
// and adds them to the per-neighbour gradient in shared memory; triples are enumerated along wrapped diagonals
// (a, (a+k) mod n) so the lanes of a batch touch distinct neighbours and the shared-memory atomics do not serialise.
// Replaces cuaev's angular kernels K2/K9 reached via models/lammps_ani.py:277 and :203-214.
#pragma once
#include "aev.cuh"

namespace anib {

// row pitch of the transposed triple-batch buffers f1T[z][triple], f2T[a][triple]: 32 triples + 4 so that the vector reads of
// the (ShfZ, ShfA) lanes fall into different banks
constexpr int kAev2TS = 36;

struct Aev2Layout {
  int off_r, off_rfc, off_s, off_a, off_afc, off_as, off_acc, off_f1, off_f2, off_ud, off_us, off_seg;   // forward
  int off_d, off_j, off_adfc, off_ai, off_g, off_gi;                                                      // backward extras
  int bytes;
};

// Forward: unsorted gather buffers (ud, us) -> species-sorted radial (r, rfc) and angular (a, afc) lists; the triple
// batch buffers f1 / f2 alias the unsorted buffer, which is dead once the lists are sorted.
// shared-memory words of the backward kernel's copy of the upstream gradient row: S radial rows of nR + 1 words (rounded up to
// a multiple of four), then one row of ang_sub + 4 words per species pair
__host__ __device__ constexpr int aev2_bwd_astride(int ang_sub) { return ang_sub + 4; }
__host__ __device__ inline int aev2_bwd_acc_elems(int S, int nR, int npairs, int ang_sub) {
  return (S * (nR + 1) + 3) / 4 * 4 + npairs * aev2_bwd_astride(ang_sub);
}

template <typename T>
__host__ __device__ inline Aev2Layout aev2_layout(int cap, int capA, int acc_elems, int nZ, int nA, bool backward) {
  Aev2Layout L{};
  int o = 0;
  auto take = [&](int bytes, int align) { o = (o + align - 1) / align * align; int r = o; o += bytes; return r; };
  const int ts = (int)sizeof(T);
  L.off_a = take(capA * 4 * ts, 4 * ts);
  if (backward) L.off_d = take(cap * 4 * ts, 4 * ts);
  else {
    const int ud = cap * 4 * ts, fb = kAev2TS * (nZ + nA) * ts;
    L.off_ud = take(ud > fb ? ud : fb, 4 * ts);
    L.off_f1 = L.off_ud;
    L.off_f2 = L.off_ud + kAev2TS * nZ * ts;
  }
  L.off_acc = take(acc_elems * ts, 16);
  L.off_afc = take(capA * ts, 8);
  if (!backward) {
    L.off_r = take(cap * ts, 8);
    L.off_rfc = take(cap * ts, 8);
    L.off_us = take(cap * 4, 4);
    L.off_seg = take(4 * (kMaxSpecies + 1) * 4, 4);   // roff, aoff, run_r, run_a
  } else {
    L.off_as = take(capA * 4, 4);
    L.off_adfc = take(capA * ts, 8);
    L.off_ai = take(capA * 4, 4);
    L.off_g = take(cap * 3 * ts, 8);
    L.off_gi = take(ts == 4 ? capA * 6 * 4 : 0, 8);   // fp32: fixed-point accumulators of the angular gradients (hi[3], lo[3])
    L.off_j = take(cap * 4, 4);
    L.off_s = take(cap * 4, 4);
  }
  L.bytes = (o + 31) / 32 * 32;
  return L;
}

constexpr int kAev2Warps = 4;

// Row selection of the two-pass scheme.  Pass 1 runs every row with shared-memory buffers sized from the survivor counts seen in
// recent steps (more resident warps than with list-sized buffers) and appends a row that does not fit to ovf_rows; pass 2 runs
// just those rows (sel_rows, *sel_cnt) with list-sized buffers, which always hold (n <= list count).  All pointers null = one
// pass with the buffers given.
struct Aev2Rows {
  int* stats = nullptr;           // [2] running max of (radial, angular) survivors (forward pass only)
  int* ovf_cnt = nullptr;
  int* ovf_rows = nullptr;
  const int* sel_cnt = nullptr;
  const int* sel_rows = nullptr;
  int* range_flag = nullptr;     // backward pass, fp32: bit 1 is set when an angular gradient term leaves the fixed-point range
};

// ------------------------------------------------------------------------------------------------ forward
template <typename T, int NZ, int NA>
__device__ __forceinline__ void
aev2_forward_row(const AevConst<T>& P, int radial_filter, const typename Vec4T<T>::type* __restrict__ x4,
                 const int64_t* __restrict__ nbr_beg, const int* __restrict__ nbr_cnt, const int* __restrict__ jlist,
                 const int* __restrict__ order, int nrows, T* __restrict__ aev, int64_t ld, void* __restrict__ aev_hi,
                 void* __restrict__ aev_lo, int64_t ld_planes, int plane_mode, float plane_scale, int cap, int capA,
                 int smem_per_warp, int* __restrict__ errflag, const Aev2Rows& rs, int row) {
  using V4 = typename Vec4T<T>::type;
  extern __shared__ __align__(32) unsigned char smem_raw[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const Aev2Layout L = aev2_layout<T>(cap, capA, P.aev_len, NZ, NA, false);
  unsigned char* base = smem_raw + (size_t)warp * smem_per_warp;
  V4* sh_a = (V4*)(base + L.off_a);
  V4* sh_ud = (V4*)(base + L.off_ud);
  T* acc = (T*)(base + L.off_acc);
  T* sh_afc = (T*)(base + L.off_afc);
  T* sh_r = (T*)(base + L.off_r);
  T* sh_rfc = (T*)(base + L.off_rfc);
  T* sh_f1 = (T*)(base + L.off_f1);
  T* sh_f2 = (T*)(base + L.off_f2);
  int* sh_us = (int*)(base + L.off_us);
  int* roff = (int*)(base + L.off_seg);            // [S+1] start of each species' segment in the radial list
  int* aoff = roff + (kMaxSpecies + 1);            // [S+1] same for the angular list
  int* run_r = aoff + (kMaxSpecies + 1);
  int* run_a = run_r + (kMaxSpecies + 1);

  const int i = order[row];
  const V4 xi = x4[i];
  const int* jl = jlist + nbr_beg[i];
  const int cnt = nbr_cnt[i];
  for (int t = lane; t < P.aev_len; t += 32) acc[t] = T(0);
  if (lane <= kMaxSpecies) { run_r[lane] = 0; run_a[lane] = 0; }
  __syncwarp();
  // ---- gather: coalesced list read, one vector load per neighbour, ballot compaction of the Rcr survivors (unsorted),
  //      per-species counts of the radial / angular survivors (one leader lane per species present in the chunk)
  const unsigned lt = (1u << lane) - 1u;
  int n = 0, na = 0;
  int jn = lane < cnt ? __ldg(jl + lane) : 0;
  V4 pn = xi;
  if (lane < cnt) pn = x4[jn];
  for (int b0 = 0; b0 < cnt; b0 += 32) {
    const int e = b0 + lane;
    const bool valid = e < cnt;
    const V4 p = pn;
    if (e + 32 < cnt) { jn = __ldg(jl + e + 32); pn = x4[jn]; }   // next chunk's index and coordinates are in flight meanwhile
    const T dx = p.x - xi.x, dy = p.y - xi.y, dz = p.z - xi.z;
    const T r2 = dx * dx + dy * dy + dz * dz;
    const bool keep = valid && (!radial_filter || r2 <= P.Rcr2) && r2 > T(0);
    const bool keepa = keep && r2 <= P.Rca2;
    const unsigned m = __ballot_sync(0xffffffffu, keep);
    const unsigned ma = __ballot_sync(0xffffffffu, keepa);
    if (keep) {
      const int pos = n + __popc(m & lt);
      const int s = unpack_species(p.w);
      V4 d; d.x = dx; d.y = dy; d.z = dz; d.w = t_sqrt(r2);
      if (pos < cap) {
        sh_ud[pos] = d;
        sh_us[pos] = s | (keepa ? 256 : 0);
      }
      const unsigned grp = __match_any_sync(m, s);
      if ((grp & lt) == 0) run_r[s] += __popc(grp);
      if (keepa) {
        const unsigned ga = grp & ma;
        if ((ga & lt) == 0) run_a[s] += __popc(ga);
      }
    }
    __syncwarp();
    n += __popc(m);
    na += __popc(ma);
  }
  if (rs.stats != nullptr && lane == 0) {   // running maxima of the survivor counts: the host sizes the next steps' buffers from them
    if (n > rs.stats[0]) atomicMax(rs.stats + 0, n);
    if (na > rs.stats[1]) atomicMax(rs.stats + 1, na);
  }
  if (n > cap || na > capA) {
    if (rs.ovf_rows != nullptr) {           // buffers sized from recent steps do not hold this atom: defer it to the second pass
      if (lane == 0) rs.ovf_rows[atomicAdd(rs.ovf_cnt, 1)] = row;
      return;
    }
    if (lane == 0) atomicExch(errflag, 1);
    if (n > cap) n = cap;                   // unreachable with list-sized buffers (n <= cnt <= cap)
  }
  // ---- segment offsets (exclusive prefix over species), run counters reset
  if (lane == 0) {
    int o_r = 0, o_a = 0;
    for (int s = 0; s < P.S; s++) {
      const int cr = run_r[s], ca = run_a[s];
      roff[s] = o_r; aoff[s] = o_a; run_r[s] = 0; run_a[s] = 0;
      o_r += cr; o_a += ca;
    }
    roff[P.S] = o_r; aoff[P.S] = o_a;
  }
  __syncwarp();
  // ---- scatter into species-sorted lists; cutoff factors are computed once per neighbour here
  for (int b0 = 0; b0 < n; b0 += 32) {
    const int e = b0 + lane;
    const bool valid = e < n;
    V4 d = xi; int s = 0;
    bool isa = false;
    if (valid) { d = sh_ud[e]; const int us = sh_us[e]; s = us & 255; isa = (us & 256) != 0; }
    const unsigned m = __ballot_sync(0xffffffffu, valid);
    const unsigned ma = __ballot_sync(0xffffffffu, isa);
    if (valid) {
      const unsigned grp = __match_any_sync(m, s);
      const int pos = roff[s] + run_r[s] + __popc(grp & lt);
      const unsigned ga = grp & ma;
      const int posa = aoff[s] + run_a[s] + __popc(ga & lt);
      __syncwarp(m);
      if ((grp & lt) == 0) run_r[s] += __popc(grp);
      if (isa && (ga & lt) == 0) run_a[s] += __popc(ga);
      const T r = d.w;
      sh_r[pos] = r;
      sh_rfc[pos] = T(0.125) * t_cos(P.pi_over_Rcr * r) + T(0.125);   // 0.25 * fc(r; Rcr)
      if (isa && posa < capA) {
        const T inv = T(1) / r;
        V4 u; u.x = d.x * inv; u.y = d.y * inv; u.z = d.z * inv; u.w = r;
        sh_a[posa] = u;
        sh_afc[posa] = T(0.5) * t_cos(P.pi_over_Rca * r) + T(0.5);
      }
    }
    __syncwarp();
  }

  // ---- radial: lane = (sub-neighbour, shift k); one species segment at a time, accumulated in a register
  {
    const T kR = -P.EtaR * T(1.4426950408889634);
    const int nRpad = P.nR <= 16 ? 16 : 32;
    const int npl = 32 / nRpad;
    const int k = lane % nRpad, sub = lane / nRpad;
    const T shf = P.ShfR[k < P.nR ? k : 0];
    for (int s = 0; s < P.S; s++) {
      const int e1 = roff[s + 1];
      if (roff[s] == e1) continue;
      T accr = T(0);
#pragma unroll 2
      for (int e = roff[s] + sub; e < e1; e += npl) {
        const T dr = sh_r[e] - shf;
        accr += sh_rfc[e] * t_gauss(dr, kR, P.EtaR);
      }
      if (npl == 2) accr += __shfl_xor_sync(0xffffffffu, accr, 16);
      if (lane < P.nR) acc[s * P.nR + lane] = accr;
    }
  }
  // ---- angular: one species pair (s <= t) at a time; batches of 32 triples: stage 1 lane = triple, stage 2 lane = term
  {
    const int iz = lane % NZ, ia = lane / NZ;
    const T kA = -P.EtaA * T(1.4426950408889634);
    for (int s = 0; s < P.S; s++) {
      const int as0 = min(aoff[s], capA), ns = min(aoff[s + 1], capA) - as0;
      if (ns == 0) continue;
      for (int t = s; t < P.S; t++) {
        const int at0 = min(aoff[t], capA), nt = min(aoff[t + 1], capA) - at0;
        const int ntri = (s == t) ? ns * (ns - 1) / 2 : ns * nt;
        if (ntri == 0) continue;
        T accr = T(0);
        // index decode without integer division: floor((idx + 0.5) / d) in fp32 is exact here (idx < 2^13, d <= capA)
        const float inv_ns = 1.0f / (float)ns, inv_nt = 1.0f / (float)nt;
        const int nfull = ns * ((ns - 1) / 2);        // s == t: triples on the complete wrapped diagonals (a, (a+k+1) mod ns)
        for (int t0 = 0; t0 < ntri; t0 += 32) {
          const int idx = t0 + lane;
          T f1[NZ], f2[NA];
#pragma unroll
          for (int z = 0; z < NZ; z++) f1[z] = T(0);
#pragma unroll
          for (int q = 0; q < NA; q++) f2[q] = T(0);
          if (idx < ntri) {
            int a, b;
            if (s == t) {
              int ta, tb;
              if (idx < nfull) {
                const int k = __float2int_rd(((float)idx + 0.5f) * inv_ns);
                ta = idx - k * ns; tb = ta + k + 1; if (tb >= ns) tb -= ns;
              } else { ta = idx - nfull; tb = ta + ns / 2; }
              a = as0 + ta; b = as0 + tb;
            } else {
              const int qa = __float2int_rd(((float)idx + 0.5f) * inv_nt);
              a = as0 + qa; b = at0 + (idx - qa * nt);
            }
            const V4 A = sh_a[a], B = sh_a[b];
            const T c = T(0.95) * (A.x * B.x + A.y * B.y + A.z * B.z);
            const T sn = t_sqrt(fmax(T(0), T(1) - c * c));
            const T fcp = T(2) * sh_afc[a] * sh_afc[b];
            const T hc = T(0.5) * c, hs = T(0.5) * sn;
#pragma unroll
            for (int z = 0; z < NZ; z++) {
              const T u = fmax(T(0), hc * P.cosZ[z] + (hs * P.sinZ[z] + T(0.5)));
              f1[z] = fcp * t_pow(u, P.Zeta);
            }
            const T ravg = T(0.5) * (A.w + B.w);
#pragma unroll
            for (int q = 0; q < NA; q++) {
              const T dr = ravg - P.ShfA[q];
              f2[q] = t_gauss(dr, kA, P.EtaA);
            }
          }
          // transposed batch buffers (idle lanes write zeros): f1T[z][lane], f2T[a][lane]; consecutive lanes -> consecutive banks
#pragma unroll
          for (int z = 0; z < NZ; z++) sh_f1[z * kAev2TS + lane] = f1[z];
#pragma unroll
          for (int q = 0; q < NA; q++) sh_f2[q * kAev2TS + lane] = f2[q];
          __syncwarp();
          const int nb4 = (min(32, ntri - t0) + 3) >> 2;
          if (lane < NA * NZ) {
            const V4* pa = (const V4*)(sh_f1 + iz * kAev2TS);
            const V4* pb = (const V4*)(sh_f2 + ia * kAev2TS);
            if constexpr (sizeof(T) == 4) {
              f32x2 acc2 = pk2((float)accr, 0.f);
#pragma unroll 2
              for (int tt = 0; tt < nb4; tt++) {     // four triples per pair of vector loads, two packed multiply-adds
                const V4 u = pa[tt], v = pb[tt];
                acc2 = ffma2(pk2((float)u.x, (float)u.y), pk2((float)v.x, (float)v.y), acc2);
                acc2 = ffma2(pk2((float)u.z, (float)u.w), pk2((float)v.z, (float)v.w), acc2);
              }
              float lo, hi;
              upk2(acc2, lo, hi);
              accr = (T)(lo + hi);
            } else {
              for (int tt = 0; tt < nb4; tt++) {     // four triples per pair of vector loads
                const V4 u = pa[tt], v = pb[tt];
                accr += u.x * v.x + u.y * v.y + u.z * v.z + u.w * v.w;
              }
            }
          }
          __syncwarp();
        }
        if (lane < NA * NZ) acc[P.rad_len + P.triu[s][t] * (NA * NZ) + lane] = accr;
      }
    }
  }
  __syncwarp();
  // ---- one coalesced store of the row: fp32/fp64 row and / or operand planes for the tensor-core MLP
  T* out = aev ? aev + (size_t)row * ld : nullptr;
  if (plane_mode == 2 && out == nullptr && (P.aev_len & 1) == 0) {
    __half2* ph = (__half2*)((__half*)aev_hi + (size_t)row * ld_planes);
    __half2* pl = (__half2*)((__half*)aev_lo + (size_t)row * ld_planes);
    for (int t = 2 * lane; t < P.aev_len; t += 64) {
      const float v0 = (float)acc[t] * plane_scale, v1 = (float)acc[t + 1] * plane_scale;
      const __half2 h = __floats2half2_rn(v0, v1);
      const float2 hf = __half22float2(h);
      ph[t >> 1] = h;
      pl[t >> 1] = __floats2half2_rn(v0 - hf.x, v1 - hf.y);
    }
    return;
  }
  for (int t = lane; t < P.aev_len; t += 32) {
    const T v = acc[t];
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

// One warp per row; pass 1 covers all rows with one row per warp, pass 2 strides a small grid over the deferred rows.
template <typename T, int NZ, int NA, bool STRIDED = false>
__global__ void __launch_bounds__(kAev2Warps * 32, sizeof(T) == 4 ? 9 : 1)   // fp32: 56 registers -> 36 warps per SM
aev2_forward_kernel(const __grid_constant__ AevConst<T> P, int radial_filter, const typename Vec4T<T>::type* __restrict__ x4,
                    const int64_t* __restrict__ nbr_beg, const int* __restrict__ nbr_cnt, const int* __restrict__ jlist,
                    const int* __restrict__ order, int nrows, T* __restrict__ aev, int64_t ld, void* __restrict__ aev_hi,
                    void* __restrict__ aev_lo, int64_t ld_planes, int plane_mode, float plane_scale, int cap, int capA,
                    int smem_per_warp, int* __restrict__ errflag, Aev2Rows rs) {
  if (!STRIDED) {   // pass 1 (or the only pass): one row per warp
    const int row = blockIdx.x * kAev2Warps + (threadIdx.x >> 5);
    if (row < nrows)
      aev2_forward_row<T, NZ, NA>(P, radial_filter, x4, nbr_beg, nbr_cnt, jlist, order, nrows, aev, ld, aev_hi, aev_lo, ld_planes,
                                  plane_mode, plane_scale, cap, capA, smem_per_warp, errflag, rs, row);
    return;
  }
  const int nsel = min(*rs.sel_cnt, nrows);
  for (int slot = blockIdx.x * kAev2Warps + (threadIdx.x >> 5); slot < nsel; slot += gridDim.x * kAev2Warps) {
    aev2_forward_row<T, NZ, NA>(P, radial_filter, x4, nbr_beg, nbr_cnt, jlist, order, nrows, aev, ld, aev_hi, aev_lo, ld_planes,
                                plane_mode, plane_scale, cap, capA, smem_per_warp, errflag, rs, rs.sel_rows[slot]);
    __syncwarp();
  }
}

// ------------------------------------------------------------------------------------------------ backward
// daev: dE/dAEV rows in bucket order (Hartree).  force: T[ntotal*3], accumulated with atomics (kcal/mol/A).
// virial_slots: double[kVirialSlots*9], accumulated (kcal/mol), or nullptr.
template <typename T, int NZ, int NA>
__device__ __forceinline__ void
aev2_backward_row(const AevConst<T>& P, int radial_filter, const typename Vec4T<T>::type* __restrict__ x4,
                  const int64_t* __restrict__ nbr_beg, const int* __restrict__ nbr_cnt, const int* __restrict__ jlist,
                  const int* __restrict__ order, int nrows, const T* __restrict__ daev, int64_t ld, T* __restrict__ force,
                  double* __restrict__ virial_slots, T out_scale, int cap, int capA, int smem_per_warp, const Aev2Rows& rs,
                  int row) {
  using V4 = typename Vec4T<T>::type;
  extern __shared__ __align__(32) unsigned char smem_raw[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int rstride = P.nR + 1;            // padded rows: species / species pairs land in different banks
  constexpr int astride = aev2_bwd_astride(NA * NZ);     // ... and a species-pair row starts on a 16-byte boundary (vector loads)
  const int acc_elems = aev2_bwd_acc_elems(P.S, P.nR, P.npairs, NA * NZ);
  const Aev2Layout L = aev2_layout<T>(cap, capA, acc_elems, NZ, NA, true);
  unsigned char* base = smem_raw + (size_t)warp * smem_per_warp;
  V4* sh_a = (V4*)(base + L.off_a);
  V4* sh_d = (V4*)(base + L.off_d);
  T* Arad = (T*)(base + L.off_acc);
  T* Aang = Arad + (P.S * rstride + 3) / 4 * 4;
  T* sh_afc = (T*)(base + L.off_afc);
  T* sh_adfc = (T*)(base + L.off_adfc);
  T* sh_g = (T*)(base + L.off_g);
  int* sh_s = (int*)(base + L.off_s);
  int* sh_j = (int*)(base + L.off_j);
  int* sh_as = (int*)(base + L.off_as);
  int* sh_ai = (int*)(base + L.off_ai);
  int* sh_gi = (int*)(base + L.off_gi);

  const int i = order[row];
  const V4 xi = x4[i];
  const int* jl = jlist + nbr_beg[i];
  const int cnt = nbr_cnt[i];
  // upstream gradient row -> shared, asynchronously (cp.async, 4/8-byte elements into the padded rows): the copy is in flight
  // while the neighbour list is gathered below
  {
    const T* g_in = daev + (size_t)row * ld;
    const bool pow2 = (P.nR & (P.nR - 1)) == 0;
    const int sh = 31 - __clz(P.nR);
    for (int t = lane; t < P.aev_len; t += 32) {
      T* dst;
      if (t < P.rad_len) {
        const int s = pow2 ? (t >> sh) : (t / P.nR);
        dst = Arad + s * rstride + (t - s * P.nR);
      } else {
        const int q = t - P.rad_len;
        dst = Aang + (q / (NA * NZ)) * astride + (q % (NA * NZ));
      }
      const uint32_t d32 = (uint32_t)__cvta_generic_to_shared(dst);
      if (sizeof(T) == 4) asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d32), "l"(g_in + t) : "memory");
      else asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(d32), "l"(g_in + t) : "memory");
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  }
  // ---- gather (as in the forward pass, plus what the gradients need)
  const unsigned lt = (1u << lane) - 1u;
  int n = 0, na = 0;
  int jn = lane < cnt ? __ldg(jl + lane) : 0;
  V4 pn = xi;
  if (lane < cnt) pn = x4[jn];
  for (int b0 = 0; b0 < cnt; b0 += 32) {
    const int e = b0 + lane;
    const bool valid = e < cnt;
    const int j = jn;
    const V4 p = pn;
    if (e + 32 < cnt) { jn = __ldg(jl + e + 32); pn = x4[jn]; }   // next chunk's index and coordinates are in flight meanwhile
    const T dx = p.x - xi.x, dy = p.y - xi.y, dz = p.z - xi.z;
    const T r2 = dx * dx + dy * dy + dz * dz;
    const bool keep = valid && (!radial_filter || r2 <= P.Rcr2) && r2 > T(0);
    const bool keepa = keep && r2 <= P.Rca2;
    const unsigned m = __ballot_sync(0xffffffffu, keep);
    const unsigned ma = __ballot_sync(0xffffffffu, keepa);
    if (keep) {
      const int pos = n + __popc(m & lt);
      const T r = t_sqrt(r2);
      V4 d; d.x = dx; d.y = dy; d.z = dz; d.w = r;
      const int s = unpack_species(p.w);
      if (pos < cap) {
        sh_d[pos] = d;
        sh_s[pos] = s;
        sh_j[pos] = j;
      }
      if (keepa) {
        const int posa = na + __popc(ma & lt);
        if (posa < capA) {
          const T inv = T(1) / r;
          V4 u; u.x = dx * inv; u.y = dy * inv; u.z = dz * inv; u.w = r;
          sh_a[posa] = u;
          T sn, cs;
          t_sincos(P.pi_over_Rca * r, &sn, &cs);
          sh_afc[posa] = T(0.5) * cs + T(0.5);
          sh_adfc[posa] = T(-0.5) * P.pi_over_Rca * sn;
          sh_as[posa] = s;
          sh_ai[posa] = pos;
        }
      }
    }
    n += __popc(m);
    na += __popc(ma);
  }
  asm volatile("cp.async.wait_all;" ::: "memory");
  __syncwarp();
  if (n > cap || na > capA) {
    if (rs.ovf_rows != nullptr) {           // deferred to the second pass (list-sized buffers)
      if (lane == 0) rs.ovf_rows[atomicAdd(rs.ovf_cnt, 1)] = row;
      return;
    }
    if (na > capA) na = capA;               // the forward pass raised the error flag
    if (n > cap) n = cap;
  }

  // fp32: the per-neighbour angular gradients are summed in 2^-36 fixed point, split into two int32 words (q = hi * 2^24 + lo,
  // lo in [0, 2^24)): shared-memory integer adds are native, float adds are compare-and-swap loops; a neighbour receives at most
  // na - 1 <= 95 contributions, so neither word overflows for |term| < 5000, and the sum does not depend on the order.
  // The bound is checked: a lane that saw a triple whose six terms sum (in absolute value) to 4096 or more, or to NaN, raises
  // bit 1 of rs.range_flag, which the host reports as status 3 (two atoms nearly on top of each other: the step is invalid).
  constexpr bool kFixed = sizeof(T) == 4;
  bool bad = false;
  const T kR = -P.EtaR * T(1.4426950408889634), kA = -P.EtaA * T(1.4426950408889634), m2EtaA = T(-2) * P.EtaA;
  if (kFixed)
    for (int e = lane; e < na * 6; e += 32) sh_gi[e] = 0;
  // ---- radial: lane = neighbour
  for (int e = lane; e < n; e += 32) {
    const V4 d = sh_d[e];
    const T r = d.w;
    T sn, cs;
    t_sincos(P.pi_over_Rcr * r, &sn, &cs);
    const T fc = T(0.5) * cs + T(0.5), dfc = T(-0.5) * P.pi_over_Rcr * sn;
    const T* Ar = Arad + sh_s[e] * rstride;
    const T c2 = T(-2) * P.EtaR * fc;
    T acc = T(0);
    for (int k = 0; k < P.nR; k++) {
      const T dr = r - P.ShfR[k];
      acc += Ar[k] * t_gauss(dr, kR, P.EtaR) * (c2 * dr + dfc);
    }
    const T coef = T(0.25) * acc / r;
    sh_g[3 * e + 0] = coef * d.x; sh_g[3 * e + 1] = coef * d.y; sh_g[3 * e + 2] = coef * d.z;
  }
  __syncwarp();
  // ---- angular: lane = triple, enumerated along wrapped diagonals (a, (a+k) mod na), k = 1 .. na/2
  {
    const int ntri = na * (na - 1) / 2;
    const int nfull = na * ((na - 1) / 2);        // triples on the complete diagonals (all of them when na is odd)
    const float inv_na = 1.0f / (float)max(na, 1);   // floor((t + 0.5) / na) in fp32 is exact for t < 2^13: no integer division
    for (int t = lane; t < ntri; t += 32) {
      int a, b;
      if (t < nfull) { const int k = __float2int_rd(((float)t + 0.5f) * inv_na); a = t - k * na; b = a + k + 1; if (b >= na) b -= na; }
      else { a = t - nfull; b = a + na / 2; }
      const V4 A = sh_a[a], B = sh_a[b];
      const T fca = sh_afc[a], fcb = sh_afc[b];
      const T cosv = A.x * B.x + A.y * B.y + A.z * B.z;
      const T c = T(0.95) * cosv;
      const T sn = t_sqrt(fmax(T(0), T(1) - c * c));
      const T c_over_sn = c / sn;
      const T hc = T(0.5) * c, hs = T(0.5) * sn;
      T f1[NZ], df1[NZ];
#pragma unroll
      for (int z = 0; z < NZ; z++) {
        const T u = fmax(T(0), hc * P.cosZ[z] + (hs * P.sinZ[z] + T(0.5)));
        const T um1 = t_pow(u, P.Zeta - T(1));
        f1[z] = um1 * u;
        df1[z] = P.Zeta * um1 * T(0.5) * (P.cosZ[z] - c_over_sn * P.sinZ[z]);   // d f1 / d c
      }
      const T ravg = T(0.5) * (A.w + B.w);
      const T* w = Aang + P.triu[sh_as[a]][sh_as[b]] * astride;
      T S0 = T(0), S1 = T(0), S2 = T(0);
      if constexpr (sizeof(T) == 4 && NZ % 4 == 0) {
        // fp32: (P, Q) and (S0, S2) as packed pairs, the weights of a shift as 128-bit loads: 1.6 instructions per (q, z) term
        // instead of 3
        f32x2 fd[NZ];
#pragma unroll
        for (int z = 0; z < NZ; z++) fd[z] = pk2((float)f1[z], (float)df1[z]);
        f32x2 S02 = pk2(0.f, 0.f);
        const float4* w4 = reinterpret_cast<const float4*>(w);
#pragma unroll
        for (int q = 0; q < NA; q++) {
          const float dr = (float)ravg - (float)P.ShfA[q];
          const float f2 = t_gauss(dr, (float)kA, 0.f);
          const float df2 = (float)m2EtaA * dr * f2;
          f32x2 PQ = pk2(0.f, 0.f);
#pragma unroll
          for (int z4 = 0; z4 < NZ / 4; z4++) {
            const float4 wq = w4[q * (NZ / 4) + z4];
            PQ = ffma2(fd[4 * z4 + 0], pk2(wq.x, wq.x), PQ);
            PQ = ffma2(fd[4 * z4 + 1], pk2(wq.y, wq.y), PQ);
            PQ = ffma2(fd[4 * z4 + 2], pk2(wq.z, wq.z), PQ);
            PQ = ffma2(fd[4 * z4 + 3], pk2(wq.w, wq.w), PQ);
          }
          S02 = ffma2(PQ, pk2(f2, f2), S02);
          float Pz, Qz;
          upk2(PQ, Pz, Qz);
          S1 += (T)(Pz * df2);
        }
        float s0, s2;
        upk2(S02, s0, s2);
        S0 = (T)s0; S2 = (T)s2;
      } else {
#pragma unroll
        for (int q = 0; q < NA; q++) {
          const T dr = ravg - P.ShfA[q];
          const T f2 = t_gauss(dr, kA, P.EtaA);
          const T df2 = m2EtaA * dr * f2;
          T Pz = T(0), Qz = T(0);
#pragma unroll
          for (int z = 0; z < NZ; z++) {
            const T wv = w[q * NZ + z];
            Pz += wv * f1[z];
            Qz += wv * df1[z];
          }
          S0 += Pz * f2; S1 += Pz * df2; S2 += Qz * f2;
        }
      }
      const T half_s1 = fca * fcb * S1;                                  // 2 * fca fcb * 0.5 * S1
      const T dE_dra = T(2) * sh_adfc[a] * fcb * S0 + half_s1;
      const T dE_drb = T(2) * fca * sh_adfc[b] * S0 + half_s1;
      const T dE_dcos = T(2) * fca * fcb * T(0.95) * S2;
      const T ka = dE_dcos / A.w, kb = dE_dcos / B.w;
      const T gax = dE_dra * A.x + ka * (B.x - cosv * A.x), gay = dE_dra * A.y + ka * (B.y - cosv * A.y),
              gaz = dE_dra * A.z + ka * (B.z - cosv * A.z);
      const T gbx = dE_drb * B.x + kb * (A.x - cosv * B.x), gby = dE_drb * B.y + kb * (A.y - cosv * B.y),
              gbz = dE_drb * B.z + kb * (A.z - cosv * B.z);
      if (kFixed) {
        const float l1 = fabsf((float)gax) + fabsf((float)gay) + fabsf((float)gaz) + fabsf((float)gbx) + fabsf((float)gby) + fabsf((float)gbz);
        bad |= !(l1 < 4096.0f);                                                  // true for NaN as well
        auto fx_add = [&](int word, T v) {
          const long long q = __float2ll_rn((float)v * 68719476736.0f);          // 2^36
          atomicAdd(sh_gi + word, (int)(q >> 24));
          atomicAdd(sh_gi + word + 3, (int)(q & 0xFFFFFF));
        };
        fx_add(6 * a + 0, gax); fx_add(6 * a + 1, gay); fx_add(6 * a + 2, gaz);
        fx_add(6 * b + 0, gbx); fx_add(6 * b + 1, gby); fx_add(6 * b + 2, gbz);
      } else {
        const int ea = sh_ai[a], eb = sh_ai[b];
        atomicAdd(sh_g + 3 * ea + 0, gax); atomicAdd(sh_g + 3 * ea + 1, gay); atomicAdd(sh_g + 3 * ea + 2, gaz);
        atomicAdd(sh_g + 3 * eb + 0, gbx); atomicAdd(sh_g + 3 * eb + 1, gby); atomicAdd(sh_g + 3 * eb + 2, gbz);
      }
    }
  }
  __syncwarp();
  if (kFixed) {   // fold the fixed-point sums into the per-neighbour gradients (one angular slot per neighbour: no conflicts)
    if (bad && rs.range_flag != nullptr) atomicOr(rs.range_flag, 2);
    for (int a = lane; a < na; a += 32) {
      const int ea = sh_ai[a];
#pragma unroll
      for (int q = 0; q < 3; q++) {
        const long long v = ((long long)sh_gi[6 * a + q] << 24) + (long long)sh_gi[6 * a + 3 + q];
        sh_g[3 * ea + q] += (T)(__ll2float_rn(v) * 1.4551915228366852e-11f);   // 2^-36
      }
    }
    __syncwarp();
  }
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


template <typename T, int NZ, int NA, bool STRIDED = false>
__global__ void __launch_bounds__(kAev2Warps * 32, sizeof(T) == 4 ? 8 : 1)   // fp32: 64 registers -> 32 warps per SM
aev2_backward_kernel(const __grid_constant__ AevConst<T> P, int radial_filter, const typename Vec4T<T>::type* __restrict__ x4,
                     const int64_t* __restrict__ nbr_beg, const int* __restrict__ nbr_cnt, const int* __restrict__ jlist,
                     const int* __restrict__ order, int nrows, const T* __restrict__ daev, int64_t ld, T* __restrict__ force,
                     double* __restrict__ virial_slots, T out_scale, int cap, int capA, int smem_per_warp, Aev2Rows rs) {
  if (!STRIDED) {
    const int row = blockIdx.x * kAev2Warps + (threadIdx.x >> 5);
    if (row < nrows)
      aev2_backward_row<T, NZ, NA>(P, radial_filter, x4, nbr_beg, nbr_cnt, jlist, order, nrows, daev, ld, force, virial_slots,
                                   out_scale, cap, capA, smem_per_warp, rs, row);
    return;
  }
  const int nsel = min(*rs.sel_cnt, nrows);
  for (int slot = blockIdx.x * kAev2Warps + (threadIdx.x >> 5); slot < nsel; slot += gridDim.x * kAev2Warps) {
    aev2_backward_row<T, NZ, NA>(P, radial_filter, x4, nbr_beg, nbr_cnt, jlist, order, nrows, daev, ld, force, virial_slots,
                                 out_scale, cap, capA, smem_per_warp, rs, rs.sel_rows[slot]);
    __syncwarp();
  }
}

}  // namespace anib
