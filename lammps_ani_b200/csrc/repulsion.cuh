// repulsion.cuh -- the RepulsionXTB pair term of the `*_repulsion.pt` models.
//
// Replaces LammpsANI.compute_repulsion (models/lammps_ani.py:301-330, wired in at :186-192): torchani's
// potentials.RepulsionXTB.compute_from_neighbors on the half list + its autograd backward.  GFN2-xTB repulsion,
//     e(r) = y_ab / r_b * exp(-sqrt_alpha_ab * r_b^k_ab) * fc(r),   r_b = r in Bohr, e in Hartree,
// with torchani's smooth (order 2) or cosine cutoff function.  One warp per local atom walks its neighbour row; every ENTRY
// carries half of the pair energy (local-local pairs appear in two rows; pairs with a ghost once, which is torchani's
// `ghost_flags` halving), so energy, forces on local + ghost atoms and the virial come out of one pass with analytic
// derivatives.  The energy goes to the total only (the reference adds it to `energies`, not to atomic_energies).
// PARITY UNPINNED: restated from the published functional form; the torchani source and a model file with `rep_calc` are not
// available offline.
#pragma once
#include "aev.cuh"

namespace anib {

template <typename T>
struct RepConst {
  T y[kMaxSpecies][kMaxSpecies], sa[kMaxSpecies][kMaxSpecies], k[kMaxSpecies][kMaxSpecies];   // indexed by COMPACT species
  T rc, rc2;
  int fn;
};

template <typename T>
__global__ void repulsion_kernel(RepConst<T> R, const typename Vec4T<T>::type* __restrict__ x4, const int64_t* __restrict__ beg,
                                 const int* __restrict__ cnt, const int* __restrict__ jlist, int nlocal, T* __restrict__ force,
                                 double* __restrict__ virial_slots, double* __restrict__ e_accum, T out_scale) {
  using V4 = typename Vec4T<T>::type;
  const int i = (int)(((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5), lane = threadIdx.x & 31;
  if (i >= nlocal) return;
  const V4 xi = x4[i];
  const int si = unpack_species(xi.w);
  const int64_t b = beg[i];
  const int n = cnt[i];
  T fix = T(0), fiy = T(0), fiz = T(0);
  double e_half = 0.0, vir[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
  for (int e = lane; e < n; e += 32) {
    const int j = jlist[b + e];
    const V4 xj = x4[j];
    const T dx = xj.x - xi.x, dy = xj.y - xi.y, dz = xj.z - xi.z;
    const T r2 = dx * dx + dy * dy + dz * dz;
    if (!(r2 <= R.rc2) || !(r2 > T(0))) continue;
    const int sj = unpack_species(xj.w);
    const T r = sqrt(r2), rinv = T(1) / r;
    const T rb = r * T(1.8897261258369282);
    const T y = R.y[si][sj], sa = R.sa[si][sj], kk = R.k[si][sj];
    const T p = pow(rb, kk);
    const T base = y / rb * exp(-sa * p);
    const T dbase = base * (-rinv - sa * kk * p * rinv);
    T fc, dfc;
    if (R.fn == 1) {
      const T t = r2 / R.rc2;
      if (!(t < T(1))) continue;
      const T u = T(1) / (T(1) - t);
      fc = exp(T(1) - u);
      dfc = -fc * u * u * (T(2) * r / R.rc2);
    } else {
      const T a = T(3.14159265358979323846) * r / R.rc;
      fc = T(0.5) * cos(a) + T(0.5);
      dfc = T(-0.5) * T(3.14159265358979323846) / R.rc * sin(a);
    }
    const T en = base * fc, de = dbase * fc + base * dfc;
    e_half += 0.5 * (double)en;
    const T gs = T(0.5) * de * rinv * out_scale;            // dE/dd = gs * d   (d = x_j - x_i), kcal/mol/A
    const T gx = gs * dx, gy = gs * dy, gz = gs * dz;
    atomic_add_t(force + 3 * (size_t)j + 0, -gx);
    atomic_add_t(force + 3 * (size_t)j + 1, -gy);
    atomic_add_t(force + 3 * (size_t)j + 2, -gz);
    fix += gx; fiy += gy; fiz += gz;
    if (virial_slots != nullptr) {
      vir[0] -= (double)(gx * dx); vir[1] -= (double)(gx * dy); vir[2] -= (double)(gx * dz);
      vir[3] -= (double)(gy * dx); vir[4] -= (double)(gy * dy); vir[5] -= (double)(gy * dz);
      vir[6] -= (double)(gz * dx); vir[7] -= (double)(gz * dy); vir[8] -= (double)(gz * dz);
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    fix += __shfl_xor_sync(0xffffffffu, fix, o);
    fiy += __shfl_xor_sync(0xffffffffu, fiy, o);
    fiz += __shfl_xor_sync(0xffffffffu, fiz, o);
    e_half += __shfl_xor_sync(0xffffffffu, e_half, o);
  }
  if (lane == 0) {
    atomic_add_t(force + 3 * (size_t)i + 0, fix);
    atomic_add_t(force + 3 * (size_t)i + 1, fiy);
    atomic_add_t(force + 3 * (size_t)i + 2, fiz);
    if (e_accum != nullptr && e_half != 0.0) atomicAdd(e_accum, e_half);
  }
  if (virial_slots != nullptr) {
#pragma unroll
    for (int q = 0; q < 9; q++) {
      double v = vir[q];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      if (lane == 0 && v != 0.0) atomicAdd(virial_slots + (size_t)(blockIdx.x % kVirialSlots) * 9 + q, v);
    }
  }
}

// scalars[0] += scale * scalars[1]   (total energy += repulsion energy; one thread)
__global__ void add_scaled_scalar_kernel(double* scalars, double scale) { scalars[0] += scale * scalars[1]; }

}  // namespace anib
