// misc.cuh -- small data-movement kernels around the hot path (outputs, halo pack/unpack).
#pragma once
#include "common.cuh"

namespace anib {

// eatom (bucket order, Hartree incl. self energy) -> per-atom array in atom order (kcal/mol); accumulate != 0: +=
__global__ void scatter_eatom_kernel(const double* __restrict__ e_pos, const int* __restrict__ order, int n, double scale,
                                     double* __restrict__ out, int accumulate) {
  int pos = blockIdx.x * blockDim.x + threadIdx.x;
  if (pos >= n) return;
  int i = order[pos];
  double v = e_pos[pos] * scale;
  if (accumulate) out[i] += v; else out[i] = v;
}

// f += force  (pair_ani_kokkos.cpp:190-191)
template <typename T, typename U>
__global__ void axpy_kernel(const T* __restrict__ src, U* __restrict__ dst, int64_t n) {
  int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (k < n) dst[k] += (U)src[k];
}

// dst = (U)src   (device-side widening of the force array before a direct D2H into the caller's buffer)
template <typename T, typename U>
__global__ void convert_kernel(const T* __restrict__ src, U* __restrict__ dst, int64_t n) {
  int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (k < n) dst[k] = (U)src[k];
}

// forward halo: out[k] = x[send_index[k]] + shift[k]   (CommBrick::forward_comm pack with PBC shift, comm_brick.cpp:549-593)
template <typename T>
__global__ void halo_pack_kernel(const T* __restrict__ x, const int64_t* __restrict__ idx, const T* __restrict__ shift, int64_t n,
                                 T* __restrict__ out) {
  int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n * 3) return;
  int64_t k = t / 3; int cpt = (int)(t % 3);
  out[t] = x[idx[k] * 3 + cpt] + (shift ? shift[t] : T(0));
}

// reverse halo: f[index[k]] += buf[k]  (PairANI::unpack_reverse_comm, pair_ani.cpp:473-484); an owner can appear
// several times in one buffer (periodic self images), hence the atomic
template <typename T>
__global__ void halo_unpack_add_kernel(T* __restrict__ f, const int64_t* __restrict__ idx, const T* __restrict__ buf, int64_t n) {
  int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n * 3) return;
  int64_t k = t / 3; int cpt = (int)(t % 3);
  atomicAdd(f + idx[k] * 3 + cpt, buf[t]);
}

// padded neighbour table -> flat list (one warp per centre atom): jlist[beg[i] + k] = table[i*rs + k*cs] & mask
__global__ void depad_neighbors_kernel(const int* __restrict__ table, int64_t rs, int64_t cs, const int64_t* __restrict__ beg,
                                       const int* __restrict__ cnt, int nlocal, int mask, int* __restrict__ jlist) {
  const int i = (int)(((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5);
  const int lane = threadIdx.x & 31;
  if (i >= nlocal) return;
  const int64_t b = beg[i];
  const int n = cnt[i];
  for (int k = lane; k < n; k += 32) jlist[b + k] = table[(int64_t)i * rs + (int64_t)k * cs] & mask;
}

}  // namespace anib
