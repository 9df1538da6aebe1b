// md.cuh -- the time integration that surrounds the pair compute, on the device.
//
// Replaces what LAMMPS's run loop does on the host around PairANI::compute (external/lammps/src/verlet.cpp:253-256,347-349):
// `fix nve` (fix_nve.cpp initial_integrate / final_integrate) and `fix langevin` (fix_langevin.cpp post_force, the default
// uniform-noise variant), in LAMMPS's own unit conventions (the caller passes dtv, dtf = 0.5 dt ftm2v, and the langevin
// prefactors), so that x, v and f never leave the GPU between neighbour rebuilds.  HBM-bound elementwise kernels, one thread
// per atom, 24-byte rows.
#pragma once
#include "common.cuh"

namespace anib {

// v += dtf f / m ;  x += dtv v          (FixNVE::initial_integrate)
__global__ void md_nve_initial_kernel(double* __restrict__ x, double* __restrict__ v, const double* __restrict__ f,
                                      const double* __restrict__ mass, int n, double dtv, double dtf) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double dtfm = dtf / mass[i];
#pragma unroll
  for (int c = 0; c < 3; c++) {
    const double vv = v[3 * (size_t)i + c] + dtfm * f[3 * (size_t)i + c];
    v[3 * (size_t)i + c] = vv;
    x[3 * (size_t)i + c] += dtv * vv;
  }
}

// v += dtf f / m                          (FixNVE::final_integrate)
__global__ void md_nve_final_kernel(double* __restrict__ v, const double* __restrict__ f, const double* __restrict__ mass, int n,
                                    double dtf) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double dtfm = dtf / mass[i];
#pragma unroll
  for (int c = 0; c < 3; c++) v[3 * (size_t)i + c] += dtfm * f[3 * (size_t)i + c];
}

// counter-based uniform random numbers: one 64-bit mix of (seed, step, atom id, component) per number, so the stream does not
// depend on how atoms are distributed over GPUs or ordered in memory (LAMMPS's RanMars is a per-rank sequential stream; the
// statistics -- uniform on [0,1), the variant fix_langevin.cpp uses by default -- are what is matched, not the bits)
__device__ __forceinline__ double md_uniform(unsigned long long seed, unsigned long long step, unsigned long long id, unsigned c) {
  unsigned long long z = seed * 0x9E3779B97F4A7C15ull + step * 0xBF58476D1CE4E5B9ull + (id * 3ull + c) * 0x94D049BB133111EBull;
  z ^= z >> 30; z *= 0xBF58476D1CE4E5B9ull;
  z ^= z >> 27; z *= 0x94D049BB133111EBull;
  z ^= z >> 31; z *= 0x9E3779B97F4A7C15ull;
  z ^= z >> 29;
  return (double)(z >> 11) * (1.0 / 9007199254740992.0);
}

// f += gamma1 v + gamma2 (u - 0.5),  gamma1 = -m g1, gamma2 = sqrt(m) g2     (FixLangevin::post_force_templated, no tally / zero /
// gjf; g1 = 1 / t_period / ftm2v, g2 = sqrt(24 boltz / t_period / dt / mvv2e) / ftm2v * sqrt(t_target), fix_langevin.cpp init())
__global__ void md_langevin_kernel(const double* __restrict__ v, double* __restrict__ f, const double* __restrict__ mass,
                                   const long long* __restrict__ tag, int n, double g1, double g2, unsigned long long seed,
                                   unsigned long long step) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double m = mass[i];
  const double gamma1 = -m * g1, gamma2 = sqrt(m) * g2;
  const unsigned long long id = tag ? (unsigned long long)tag[i] : (unsigned long long)i;
#pragma unroll
  for (int c = 0; c < 3; c++)
    f[3 * (size_t)i + c] += gamma1 * v[3 * (size_t)i + c] + gamma2 * (md_uniform(seed, step, id, (unsigned)c) - 0.5);
}

// sum_i m_i |v_i|^2 (fp64), one partial per block; the caller's second stage adds them up
__global__ void md_mv2_kernel(const double* __restrict__ v, const double* __restrict__ mass, int n, double* __restrict__ partial) {
  __shared__ double sh[256];
  double s = 0.0;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const double a = v[3 * (size_t)i], b = v[3 * (size_t)i + 1], c = v[3 * (size_t)i + 2];
    s += mass[i] * (a * a + b * b + c * c);
  }
  sh[threadIdx.x] = s;
  __syncthreads();
  for (int w = 128; w > 0; w >>= 1) {
    if ((int)threadIdx.x < w) sh[threadIdx.x] += sh[threadIdx.x + w];
    __syncthreads();
  }
  if (threadIdx.x == 0) partial[blockIdx.x] = sh[0];
}

// max_i |x_i - x0_i|^2 over the local atoms (Neighbor::check_distance, external/lammps/src/neighbor.cpp, on the device): the bits
// of a non-negative double order like the number, so one atomicMax on the 64-bit pattern per block does it
__global__ void md_max_disp2_kernel(const double* __restrict__ x, const double* __restrict__ x0, int n, unsigned long long* __restrict__ out) {
  __shared__ double sh[256];
  double m = 0.0;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const double a = x[3 * (size_t)i] - x0[3 * (size_t)i], b = x[3 * (size_t)i + 1] - x0[3 * (size_t)i + 1],
                 c = x[3 * (size_t)i + 2] - x0[3 * (size_t)i + 2];
    m = fmax(m, a * a + b * b + c * c);
  }
  sh[threadIdx.x] = m;
  __syncthreads();
  for (int w = 128; w > 0; w >>= 1) {
    if ((int)threadIdx.x < w) sh[threadIdx.x] = fmax(sh[threadIdx.x], sh[threadIdx.x + w]);
    __syncthreads();
  }
  if (threadIdx.x == 0) atomicMax(out, (unsigned long long)__double_as_longlong(sh[0]));
}

// ---- allocation-free variant of the sum (ani_md_sum_mv2_static): the per-block partials live in a file-scope device array
// instead of a stream-ordered allocation.  Every GPU that loads the library holds these 8 KB until the process exits.
constexpr int kMv2StaticBlocks = 1024;                 // partials held by the static scratch
constexpr int kMv2StaticMaxAtoms = 1 << 24;            // stated limit of the entry point (16 777 216 atoms: 64 grid-stride passes)
static __device__ __align__(256) double g_mv2_static_scratch[kMv2StaticBlocks];

// second stage: one block adds the nb <= 1024 partials in a fixed order (the result does not depend on the launch) -> out[0]
__global__ void md_mv2_final_kernel(const double* __restrict__ partial, int nb, double* __restrict__ out) {
  __shared__ double sh[256];
  double s = 0.0;
  for (int i = threadIdx.x; i < nb; i += 256) s += partial[i];
  sh[threadIdx.x] = s;
  __syncthreads();
  for (int w = 128; w > 0; w >>= 1) {
    if ((int)threadIdx.x < w) sh[threadIdx.x] += sh[threadIdx.x + w];
    __syncthreads();
  }
  if (threadIdx.x == 0) out[0] = sh[0];
}

// ghost-force fold for a single periodic domain: f[owner[g]] += f[nlocal + g]   (pair_ani.cpp:461-484 with every ghost a self image)
__global__ void md_fold_ghost_kernel(double* __restrict__ f, const int* __restrict__ owner, int nlocal, int nghost) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (long long)nghost * 3) return;
  const int g = (int)(t / 3), c = (int)(t % 3);
  atomicAdd(f + (size_t)owner[g] * 3 + c, f[(size_t)(nlocal + g) * 3 + c]);
}

// ghost positions of a single periodic domain: x[nlocal + g] = x[owner[g]] + shift[g]   (CommBrick::forward_comm, comm_brick.cpp:549-593)
__global__ void md_ghost_x_kernel(double* __restrict__ x, const int* __restrict__ owner, const double* __restrict__ shift, int nlocal,
                                  int nghost) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (long long)nghost * 3) return;
  const int g = (int)(t / 3), c = (int)(t % 3);
  x[(size_t)(nlocal + g) * 3 + c] = x[(size_t)owner[g] * 3 + c] + shift[t];
}

}  // namespace anib
