// nbuild.cuh -- full neighbour list built on the device.
//
// What it replaces: `neighbor->build` (external/lammps/src/npair_full_bin.cpp / KOKKOS/npair_kokkos.cpp) followed by the host
// flatten of src/pair_ani.cpp:107-151 and the H2D copy of the list (ani.cpp:216-218): 590 MB per rebuild for a 1 M-atom
// subdomain.  Here the list never exists on the host: atoms are binned into cells of one cutoff (stable radix sort by cell id,
// so atoms of a cell stay in index order), a warp per local atom scans the 27 cells around it -- once, writing into a row of
// fixed stride, when the longest row of the previous build is known; count, scan, fill otherwise -- and the CSR goes straight to
// the engine, together with the species bucketing of the rows (also done here, on the device).  Integer work: the per-atom neighbour SETS are identical to LAMMPS's (all j != i, local or
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

// Species bookkeeping of a rebuild on the device (the host used to loop five times over the atoms): the sort key of every local
// atom, 2 * species + (1 if it has a ghost neighbour: a BOUNDARY atom), the histogram of the keys, the set of species present
// among local + ghost atoms, and a flag for indices outside the model.  Rows sorted by this key are bucketed by species with
// the interior atoms first in each bucket: their AEVs do not depend on ghost positions (and no ghost force comes from them),
// which is what lets the halo exchange overlap their share of the AEV kernels.
// stats: [0..15] histogram of the keys, [16] present mask, [17] invalid flag
constexpr int kNbStats = 18;
struct NbSpeciesMap { int cmap[8]; };
__global__ void nb_species_stats_kernel(const int* __restrict__ species, const int* __restrict__ bnd, int nlocal, int ntotal, int S,
                                        int* __restrict__ key, int* __restrict__ stats) {
  __shared__ int sh[kNbStats];
  if (threadIdx.x < kNbStats) sh[threadIdx.x] = 0;
  __syncthreads();
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < ntotal; i += gridDim.x * blockDim.x) {
    const int s = species[i];
    if (s < 0 || s >= S) { sh[17] = 1; if (i < nlocal) key[i] = 0; continue; }
    if (i < nlocal) {
      const int k = 2 * s + (bnd[i] != 0 ? 1 : 0);
      key[i] = k;
      atomicAdd(&sh[k], 1);
    }
    atomicOr(&sh[16], 1 << s);
  }
  __syncthreads();
  if (threadIdx.x < 16 && sh[threadIdx.x]) atomicAdd(&stats[threadIdx.x], sh[threadIdx.x]);
  if (threadIdx.x == 16 && sh[16]) atomicOr(&stats[16], sh[16]);
  if (threadIdx.x == 17 && sh[17]) atomicOr(&stats[17], 1);
}
// rows (positions in the sorted order) of the interior atoms, then of the boundary atoms: list[0 .. n_int), list[n_int .. nlocal)
struct NbRowSplit { int boff[9], nint[8], int_before[8], bnd_before[8], n_int, S; };
__global__ void nb_row_lists_kernel(NbRowSplit sp, int nlocal, int* __restrict__ list) {
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= nlocal) return;
  int s = 0;
  while (s + 1 < sp.S && r >= sp.boff[s + 1]) s++;
  const int k = r - sp.boff[s];
  if (k < sp.nint[s]) list[sp.int_before[s] + k] = r;
  else list[sp.n_int + sp.bnd_before[s] + (k - sp.nint[s])] = r;
}
// compact species index of every atom (the kernels' species), and the sort key of the local atoms
__global__ void nb_species_compact_kernel(const int* __restrict__ species, int ntotal, NbSpeciesMap m, int* __restrict__ cspecies) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < ntotal) cspecies[i] = m.cmap[species[i] & 7];
}

// positions in cell-sorted order for the search below: pk[k] = (x, y, z relative to the grid origin in fp32, atom index as
// bits) -- one 16-byte load per candidate -- and the caller's doubles, one array per component, for the borderline cases
__global__ void nb_gather_sorted_kernel(const double* __restrict__ x, const int* __restrict__ sorted_atom, int ntotal, NbGrid G,
                                        double* __restrict__ xs, double* __restrict__ ys, double* __restrict__ zs,
                                        float4* __restrict__ pk) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= ntotal) return;
  const int j = sorted_atom[k];
  const double a = x[3 * (size_t)j], b = x[3 * (size_t)j + 1], c = x[3 * (size_t)j + 2];
  xs[k] = a; ys[k] = b; zs[k] = c;
  pk[k] = make_float4((float)(a - G.lo[0]), (float)(b - G.lo[1]), (float)(c - G.lo[2]), __int_as_float(j));
}

__global__ void nb_strided_beg_kernel(int64_t* __restrict__ beg, int n, int stride) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) beg[i] = (int64_t)i * stride;
}

// One WARP per atom, atoms taken in cell-sorted order (ghosts leave at once): the 32 lanes test 32 candidates of a slice at a
// time (one coalesced 16-byte load each; neighbouring warps read the same slices, which stay in L1), and a ballot compacts the
// hits, so a row is written as contiguous runs and in the same order (ascending sorted position) as a serial scan would.
// MODE 0: count.  MODE 1: write at beg[i] (after a count and a scan).  MODE 2: count AND write in one pass into rows of a fixed
// stride (the longest row of the previous build plus a margin); a row that does not fit is only counted, and the caller, who
// sees the longest row, then falls back to modes 0 + 1.
// The criterion is the fp64 one, |x_j - x_i|^2 <= cut2 on the caller's doubles; the test is first made in fp32 on positions
// relative to the grid origin, whose error is bounded by eps (set by the caller from the extent of the grid), and only a
// candidate within eps of the cutoff -- one in 10^4 -- is decided in fp64.  (The thread-per-atom all-fp64 version this
// replaces took 0.73 + 1.16 ms on 300 k atoms: 2 100 instructions per atom and pass, scattered 4-byte writes.)
template <int MODE>
__global__ void nb_search_warp_kernel(const double* __restrict__ xs, const double* __restrict__ ys, const double* __restrict__ zs,
                                      const float4* __restrict__ pk, int ntotal, int nlocal, NbGrid G, const int* __restrict__ cell_beg,
                                      const int* __restrict__ cell_end, double cut2, float eps, int* __restrict__ cnt,
                                      int* __restrict__ bnd, const int64_t* __restrict__ beg, int stride, int* __restrict__ jlist) {
  const int w = (int)(((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5), lane = threadIdx.x & 31;
  if (w >= ntotal) return;
  const float4 pi = pk[w];
  const int i = __float_as_int(pi.w);
  if (i >= nlocal) return;
  const float c2lo = (float)cut2 - eps, c2hi = (float)cut2 + eps;
  // the cell of atom i from its double position, exactly as nb_cell_id_kernel computed it
  const int cx = nb_cell_coord(xs[w], G.lo[0], G.inv_cell, G.n[0]);
  const int cy = nb_cell_coord(ys[w], G.lo[1], G.inv_cell, G.n[1]);
  const int cz = nb_cell_coord(zs[w], G.lo[2], G.inv_cell, G.n[2]);
  const unsigned lt = (1u << lane) - 1u;
  int n = 0;
  bool ghost_nb = false;                          // a neighbour of this atom is a ghost (index >= nlocal)
  int* out = MODE == 1 ? jlist + beg[i] : MODE == 2 ? jlist + (size_t)i * stride : nullptr;
  const int room = MODE == 2 ? stride : 0x7fffffff;
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
      for (int k0 = kb; k0 < ke; k0 += 32) {
        const int k = k0 + lane;
        bool hit = false;
        int j = -1;
        if (k < ke) {
          const float4 pj = pk[k];
          j = __float_as_int(pj.w);
          const float ax = pj.x - pi.x, ay = pj.y - pi.y, az = pj.z - pi.z;
          const float d2 = ax * ax + ay * ay + az * az;
          hit = d2 <= c2lo;
          if (!hit && d2 <= c2hi) {                      // too close to call in fp32
            const double ddx = xs[k] - xs[w], ddy = ys[k] - ys[w], ddz = zs[k] - zs[w];
            hit = ddx * ddx + ddy * ddy + ddz * ddz <= cut2;
          }
          hit = hit && j != i;
        }
        const unsigned m = __ballot_sync(0xffffffffu, hit);
        if (MODE != 1) ghost_nb |= __any_sync(0xffffffffu, hit && j >= nlocal);
        if (MODE != 0 && hit) {
          const int pos = n + __popc(m & lt);
          if (pos < room) out[pos] = j;
        }
        n += __popc(m);
      }
    }
  }
  if (MODE != 1 && lane == 0) { cnt[i] = n; bnd[i] = ghost_nb ? 1 : 0; }
}

}  // namespace anib
