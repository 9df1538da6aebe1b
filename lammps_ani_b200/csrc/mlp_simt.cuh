// mlp_simt.cuh -- per-species ensemble MLP, forward + dE/dAEV backward, as FFMA / DFMA tiled GEMMs with the
// bias / CELU / CELU' work fused into their epilogues.  Used for the `double` precision mode and as the strict
// fp32 path (ANI_MLP_FP32_SIMT); the tcgen05 path (mlp_tc.cuh) has the same data flow.
//
// Replaces BmmEnsemble / ANINetworks forward (torch.bmm -> cuBLAS SGEMM + separate bias and CELU kernels) and the
// autograd backward through them (models/lammps_ani.py:219-257,203-214; K4-K7 of SURVEY.md 2.1).
//
// Layout: activations of a chunk of R atoms of one species are stored as [R][M][width] (member-major columns),
// so member m's GEMM operand is a strided view (lda = M*width) and the final dE/dAEV GEMM is ONE GEMM with the
// ensemble mean folded in by concatenating the members along K:  dAEV[R][L] = (1/M) G1[R][M*h1] . W1cat[M*h1][L].
#pragma once
#include "common.cuh"

namespace anib {

enum { EPI_BIAS_CELU = 0, EPI_MUL_AUX = 1, EPI_SCALE = 2 };

template <typename T> __device__ __forceinline__ T celu_fwd(T z);
template <> __device__ __forceinline__ float celu_fwd<float>(float z) {
  return z > 0.f ? z : 0.1f * (exp2f(z * 14.426950408889634f) - 1.f);   // exp(z/0.1)
}
template <> __device__ __forceinline__ double celu_fwd<double>(double z) { return z > 0.0 ? z : 0.1 * expm1(z / 0.1); }

// CELU'(z).  fp64 follows torch autograd's formula elu_backward(grad, alpha, 1, 1.0/alpha.toFloat()):
// 0.1*k*exp(k z) with k = 1/(double)(float)0.1, because the reference's fp64 numbers contain it (oracle/ani_oracle.c,
// dcelu).  In fp32 that k rounds to exactly 10.
template <typename T> __device__ __forceinline__ T celu_bwd(T z);
template <> __device__ __forceinline__ float celu_bwd<float>(float z) {
  return z > 0.f ? 1.f : exp2f(z * 14.426950408889634f);
}
template <> __device__ __forceinline__ double celu_bwd<double>(double z) {
  const double k = 1.0 / (double)0.1f;
  return z > 0.0 ? 1.0 : 0.1 * k * exp(k * z);
}

// C[r][n] = epi( sum_k A[r][k] * B[n][k] ),  A: [R][K] pitch lda, B: [N][K] pitch ldb  (C = A . B^T), blockIdx.z = member.
//   EPI_BIAS_CELU: z = acc + bias[n];  C = celu(z);  C2 = celu'(z)
//   EPI_MUL_AUX  : C = acc * aux[r][n]
//   EPI_SCALE    : C = acc * scale
// Requires K % 4 == 0 and 16-byte aligned rows (all ANI layer widths are multiples of 16).
template <typename T, int EPI>
__global__ void __launch_bounds__(256)
gemm_nt_kernel(int R, int N, int K, const T* __restrict__ A, int64_t lda, int64_t sAz, const T* __restrict__ B, int64_t ldb,
               int64_t sBz, T* __restrict__ C, int64_t ldc, int64_t sCz, T* __restrict__ C2, const T* __restrict__ bias,
               int64_t sbz, const T* __restrict__ aux, int64_t ldx, int64_t sXz, T scale) {
  constexpr int BM = 64, BN = 64, BK = 16;
  __shared__ __align__(16) T As[BK][BM + 4];
  __shared__ __align__(16) T Bs[BK][BN + 4];
  const int z = blockIdx.z;
  A += (size_t)z * sAz; B += (size_t)z * sBz; C += (size_t)z * sCz;
  if (EPI == EPI_BIAS_CELU) { C2 += (size_t)z * sCz; bias += (size_t)z * sbz; }
  if (EPI == EPI_MUL_AUX) aux += (size_t)z * sXz;
  const int r0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int lrow = tid >> 2, lk = (tid & 3) * 4;   // loader: 64 rows x 4 k-quads
  T acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; i++)
#pragma unroll
    for (int j = 0; j < 4; j++) acc[i][j] = T(0);
  for (int k0 = 0; k0 < K; k0 += BK) {
    T a[4] = {T(0), T(0), T(0), T(0)}, b[4] = {T(0), T(0), T(0), T(0)};
    if (r0 + lrow < R && k0 + lk < K) {
      const T* p = A + (size_t)(r0 + lrow) * lda + k0 + lk;
#pragma unroll
      for (int q = 0; q < 4; q++) a[q] = p[q];
    }
    if (n0 + lrow < N && k0 + lk < K) {
      const T* p = B + (size_t)(n0 + lrow) * ldb + k0 + lk;
#pragma unroll
      for (int q = 0; q < 4; q++) b[q] = p[q];
    }
    __syncthreads();
#pragma unroll
    for (int q = 0; q < 4; q++) { As[lk + q][lrow] = a[q]; Bs[lk + q][lrow] = b[q]; }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < BK; k++) {
      T av[4], bv[4];
#pragma unroll
      for (int i = 0; i < 4; i++) av[i] = As[k][ty * 4 + i];
#pragma unroll
      for (int j = 0; j < 4; j++) bv[j] = Bs[k][tx * 4 + j];
#pragma unroll
      for (int i = 0; i < 4; i++)
#pragma unroll
        for (int j = 0; j < 4; j++) acc[i][j] = fma(av[i], bv[j], acc[i][j]);
    }
  }
#pragma unroll
  for (int i = 0; i < 4; i++) {
    const int r = r0 + ty * 4 + i;
    if (r >= R) continue;
#pragma unroll
    for (int j = 0; j < 4; j++) {
      const int n = n0 + tx * 4 + j;
      if (n >= N) continue;
      T v = acc[i][j];
      if (EPI == EPI_BIAS_CELU) {
        T zz = v + bias[n];
        C[(size_t)r * ldc + n] = celu_fwd<T>(zz);
        C2[(size_t)r * ldc + n] = celu_bwd<T>(zz);
      } else if (EPI == EPI_MUL_AUX) {
        C[(size_t)r * ldc + n] = v * aux[(size_t)r * ldx + n];
      } else {
        C[(size_t)r * ldc + n] = v * scale;
      }
    }
  }
}

// Output layer (width h -> 1) + start of the backward pass, one warp per atom row:
//   e_m = b4[m] + sum_n H[r][m][n] w4[m][n];   eatom[row0 + r] = mean_m e_m + sae;   G[r][m][n] = w4[m][n] * D[r][m][n]
template <typename T>
__global__ void __launch_bounds__(128)
mlp_out_kernel(int R, int M, int h, const T* __restrict__ H, const T* __restrict__ D, const T* __restrict__ w4,
               const T* __restrict__ b4, T* __restrict__ G, double* __restrict__ eatom, double sae) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int r = blockIdx.x * 4 + warp;
  if (r >= R) return;
  double esum = 0.0;
  for (int m = 0; m < M; m++) {
    const T* hrow = H + ((size_t)r * M + m) * h;
    const T* drow = D + ((size_t)r * M + m) * h;
    T* grow = G + ((size_t)r * M + m) * h;
    const T* w = w4 + (size_t)m * h;
    T acc = T(0);
    for (int n = lane; n < h; n += 32) {
      T wv = w[n];
      acc = fma(hrow[n], wv, acc);
      grow[n] = wv * drow[n];
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    esum += (double)(acc + b4[m]);
  }
  if (lane == 0) eatom[r] = esum / (double)M + sae;
}

// Deterministic two-stage sum of doubles.
__global__ void reduce_sum_stage1(const double* __restrict__ in, int64_t n, double* __restrict__ partial) {
  __shared__ double sh[256];
  double s = 0.0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) s += in[i];
  sh[threadIdx.x] = s;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if ((int)threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) partial[blockIdx.x] = sh[0];
}
__global__ void reduce_sum_stage2(const double* __restrict__ partial, int n, double* __restrict__ out, double scale) {
  __shared__ double sh[256];
  double s = 0.0;
  for (int i = threadIdx.x; i < n; i += blockDim.x) s += partial[i];
  sh[threadIdx.x] = s;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if ((int)threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) out[0] = sh[0] * scale;
}

}  // namespace anib
