// nbuild.cuh -- full neighbour list built on the device.
//
// What it replaces: `neighbor->build` (external/lammps/src/npair_full_bin.cpp / KOKKOS/npair_kokkos.cpp) followed by the host
// flatten of src/pair_ani.cpp:107-151 and the H2D copy of the list (ani.cpp:216-218): 590 MB per rebuild for a 1 M-atom
// subdomain.  Here the list never exists on the host: atoms are binned into cells of one cutoff (stable radix sort by cell id,
// so atoms of a cell stay in index order), every local atom scans its 27 cells twice (count, then fill at its CSR offset), and
// the CSR goes straight to the engine.  Integer work: the per-atom neighbour SETS are identical to LAMMPS's (all j != i, local or
// ghost, with |x_i - x_j|^2 <= cutoff^2 in fp64); the order inside a row is (cell, index), which the AEV kernels do not depend
// on beyond fp32 summation order.
#pragma once
#include <cub/cub.cuh>
#include <thrust/iterator/transform_iterator.h>

#include "common.cuh"

namespace anib {

struct NbToI64 {
  __host__ __device__ __forceinline__ int64_t operator()(const int& v) const { return (int64_t)v; }
};

// total number of stored pairs from the exclusive scan (a one-thread kernel instead of a host pass over nlocal counts)
__global__ void nb_total_kernel(const int64_t* __restrict__ beg, const int* __restrict__ cnt, int n, int64_t* __restrict__ out) {
  *out = n > 0 ? beg[n - 1] + (int64_t)cnt[n - 1] : 0;
}

struct NbGrid {
  double lo[3], inv_cell;
  int n[3];
};

__device__ __forceinline__ int nb_cell_coord(double x, double lo, double inv, int n) {
  int c = (int)floor((x - lo) * inv);
  return c < 0 ? 0 : (c >= n ? n - 1 : c);
}

__global__ void nb_cell_id_kernel(const double* __restrict__ x, int ntotal, NbGrid G, int* __restrict__ cell_id, int* __restrict__ atom_id) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= ntotal) return;
  const int cx = nb_cell_coord(x[3 * (size_t)i + 0], G.lo[0], G.inv_cell, G.n[0]);
  const int cy = nb_cell_coord(x[3 * (size_t)i + 1], G.lo[1], G.inv_cell, G.n[1]);
  const int cz = nb_cell_coord(x[3 * (size_t)i + 2], G.lo[2], G.inv_cell, G.n[2]);
  cell_id[i] = (cz * G.n[1] + cy) * G.n[0] + cx;
  atom_id[i] = i;
}

// sorted_cell ascending: cell_beg[c] .. cell_end[c] is the slice of cell c (both arrays zeroed beforehand: empty cells are 0..0)
__global__ void nb_cell_range_kernel(const int* __restrict__ sorted_cell, int ntotal, int* __restrict__ cell_beg, int* __restrict__ cell_end) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= ntotal) return;
  const int c = sorted_cell[k];
  if (k == 0 || sorted_cell[k - 1] != c) cell_beg[c] = k;
  if (k == ntotal - 1 || sorted_cell[k + 1] != c) cell_end[c] = k + 1;
}

// one thread per local atom; FILL = false: count, FILL = true: write at beg[i]
template <bool FILL>
__global__ void nb_search_kernel(const double* __restrict__ x, int nlocal, NbGrid G, const int* __restrict__ cell_beg,
                                 const int* __restrict__ cell_end, const int* __restrict__ sorted_atom, double cut2,
                                 int* __restrict__ cnt, const int64_t* __restrict__ beg, int* __restrict__ jlist) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nlocal) return;
  const double xi = x[3 * (size_t)i], yi = x[3 * (size_t)i + 1], zi = x[3 * (size_t)i + 2];
  const int cx = nb_cell_coord(xi, G.lo[0], G.inv_cell, G.n[0]);
  const int cy = nb_cell_coord(yi, G.lo[1], G.inv_cell, G.n[1]);
  const int cz = nb_cell_coord(zi, G.lo[2], G.inv_cell, G.n[2]);
  int n = 0;
  int* out = FILL ? jlist + beg[i] : nullptr;
  for (int dz = -1; dz <= 1; dz++) {
    const int z = cz + dz;
    if (z < 0 || z >= G.n[2]) continue;
    for (int dy = -1; dy <= 1; dy++) {
      const int y = cy + dy;
      if (y < 0 || y >= G.n[1]) continue;
      const int x0 = max(cx - 1, 0), x1 = min(cx + 1, G.n[0] - 1);
      const int row = (z * G.n[1] + y) * G.n[0];
      // cells x0..x1 of one row are consecutive in the sorted order: one contiguous slice
      int kb = -1, ke = -1;
      for (int c = x0; c <= x1; c++) {
        const int b = cell_beg[row + c], e = cell_end[row + c];
        if (e > b) { if (kb < 0) kb = b; ke = e; }
      }
      for (int k = kb; k < ke; k++) {
        const int j = sorted_atom[k];
        const double ddx = x[3 * (size_t)j] - xi, ddy = x[3 * (size_t)j + 1] - yi, ddz = x[3 * (size_t)j + 2] - zi;
        if (j != i && ddx * ddx + ddy * ddy + ddz * ddz <= cut2) {
          if (FILL) out[n] = j;
          n++;
        }
      }
    }
  }
  if (!FILL) cnt[i] = n;
}

}  // namespace anib
