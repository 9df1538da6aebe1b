// mlp_tc.cuh -- the ensemble MLP on Blackwell tensor cores: TMA-fed tcgen05.mma (kind::tf32) with the
// accumulator in TMEM and bias / CELU / CELU' / ensemble-mean fused into the epilogues.
//
// Replaces torch.bmm -> cuBLAS SGEMM + separate bias / CELU kernels and their autograd backward
// (BmmEnsemble via models/lammps_ani.py:219-257,203-214; K5/K7 of SURVEY.md 2.1).
//
// fp32 parity on a tensor core that has no fp32 mode: every operand is kept as two TF32-exact planes,
//   hi = rna_tf32(x), lo = rna_tf32(x - hi), and  A.B ~= A_hi.B_hi + A_lo.B_hi + A_hi.B_lo   (3 MMAs, fp32 accumulate),
// which leaves ~2^-21 relative error per product (ANI_MLP_TF32X3).  ANI_MLP_TF32 issues only the hi.hi product
// (the reference's LAMMPS_ANI_ALLOW_TF32=1 behaviour).
//
// One persistent warp-specialised kernel does every GEMM of the network:
//   warp 0   : TMA producer  (cp.async.bulk.tensor.2d, SWIZZLE_128B, 32 fp32 = 128 B per row, mbarrier tx-count)
//   warp 1   : MMA issuer    (one lane; UMMA M=128, N = layer width <= 256, K=8 per instruction; tcgen05.commit)
//   warp 2   : TMEM allocator (512 columns = two 256-column accumulator stages, epilogue overlaps the next tile)
//   warps 4-7: epilogue      (tcgen05.ld 32x32b -> registers -> fused math -> hi/lo planes or fp32 rows in HBM)
// Tile = 128 atom rows x full layer width, one ensemble member; tiles of the same atom rows and different members
// are adjacent in the schedule so their A tiles (the AEV rows) are shared through L2.
#pragma once
#include <cuda.h>

#include "common.cuh"

namespace anib {

struct DevBufTc {
  void* p = nullptr;
  size_t cap = 0;
  ~DevBufTc() { if (p) cudaFree(p); }
  bool reserve(size_t bytes) {
    if (bytes <= cap) return true;
    if (p) { cudaFree(p); p = nullptr; cap = 0; }
    if (cudaMalloc(&p, bytes + 1024) != cudaSuccess) { p = nullptr; return false; }
    cap = bytes;
    return true;
  }
  float* f() const { return (float*)p; }
};

// Per-species TF32 planes of the weights (built once per context).
struct TcSpeciesWeights {
  bool ready = false;
  int nh = 0;                                  // hidden layers
  std::vector<DevBufTc> Wf_hi, Wf_lo;          // [j] j=0..nh-1 : [M*dims[j+1]][dims[j]]
  std::vector<DevBufTc> WT_hi, WT_lo;          // [j] j=1..nh-1 : [M*dims[j]][dims[j+1]]
  DevBufTc Wcat_hi, Wcat_lo;                   // [L][M*h1]
  std::vector<DevBufTc> bias;                  // [j] j=0..nh-1 : [M][dims[j+1]] fp32
  DevBufTc w_out, b_out;                       // [M][dims[nh]], [M]
};

struct TcWork {
  int chunk_rows = 0;
  std::vector<DevBufTc> H_hi, H_lo, G_hi, G_lo;   // [l] l=1..nh : [chunk_rows][M*dims[l]]
};

namespace tc {

constexpr int BM = 128;
constexpr int BK = 32;            // fp32 elements per k-block = 128 bytes = one swizzle row
constexpr int kThreads = 256;
constexpr int kMaxStages = 8;
constexpr int kAStageBytes = BM * BK * 4;   // 16 KB
enum { EPI_F = 0, EPI_B = 1, EPI_D = 2 };

struct GemmArgs {
  int R;                 // valid rows of this problem (rows >= R are masked in the epilogue)
  int N;                 // valid output columns
  int BN;                // tile width (UMMA N), multiple of 16, <= 256
  int num_n_tiles;
  int num_row_tiles;
  int K;                 // reduction length per member (multiple of 8; TMA zero-fills the ragged last k-block)
  int batch;             // members
  int a_row0;            // first row of A inside its tensor map
  int a_k_off_per_z;     // column offset of member z inside A's inner dimension
  int b_row_off_per_z;   // row offset of member z inside B
  int three_pass;
  int stages;
  const float* bias; int bias_stride_z;
  float* out_hi; float* out_lo; long long ld_out; int out_col_off_per_z;
  const float* aux_hi; const float* aux_lo; long long ld_aux; int aux_col_off_per_z;
  float scale;
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_LOOP:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra WAIT_DONE;\n"
      "bra WAIT_LOOP;\n"
      "WAIT_DONE:\n"
      "}\n" ::"r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void tma_load_2d(const CUtensorMap* tm, uint32_t bar, uint32_t dst, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(dst), "l"(tm), "r"(bar), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n"
      "}\n" ::"r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(accumulate) : "memory");
}
// K-major operand tile, rows of 128 bytes, SWIZZLE_128B: SBO = 1024 B between 8-row groups, LBO unused (=1),
// descriptor version 1 (sm_100), layout type 2 (cute/arch/mma_sm100_desc.hpp SmemDescriptor).
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFF) >> 4);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
// instruction descriptor: D=f32 (bits 4-5 = 1), A=B=TF32 (2 at bits 7-9 and 10-12), K-major both, N>>3 at 17, M>>4 at 24
__device__ __forceinline__ uint32_t umma_idesc_tf32(int M, int N) {
  return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ float rna_tf32(float f) {
  uint32_t u;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(u) : "f"(f));
  return __uint_as_float(u);
}

template <int EPI>
__global__ void __launch_bounds__(kThreads, 1)
tc_gemm_kernel(const __grid_constant__ CUtensorMap tmA_hi, const __grid_constant__ CUtensorMap tmA_lo,
               const __grid_constant__ CUtensorMap tmB_hi, const __grid_constant__ CUtensorMap tmB_lo, const GemmArgs g) {
  extern __shared__ unsigned char smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  const int planes = g.three_pass ? 2 : 1;
  const uint32_t b_bytes = (uint32_t)g.BN * 128u;
  const uint32_t stage_bytes = planes * (kAStageBytes + b_bytes);
  const uint32_t bar_base = base + g.stages * stage_bytes;       // 8-byte aligned (multiple of 1024)
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (kMaxStages + s); };
  auto tfull_bar = [&](int s) { return bar_base + 8u * (2 * kMaxStages + s); };
  auto tempty_bar = [&](int s) { return bar_base + 8u * (2 * kMaxStages + 2 + s); };
  const uint32_t tmem_slot = bar_base + 8u * (2 * kMaxStages + 4);
  volatile uint32_t* tmem_slot_ptr = (volatile uint32_t*)(smem_raw + (tmem_slot - raw));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int num_tiles = g.num_row_tiles * g.num_n_tiles * g.batch;
  const int KB = (g.K + BK - 1) / BK;

  if (warp == 1 && lane == 0) {
    for (int s = 0; s < g.stages; s++) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
    for (int s = 0; s < 2; s++) { mbar_init(tfull_bar(s), 1); mbar_init(tempty_bar(s), 4); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 2) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(512u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;

  if (warp == 0) {
    // ================================================================= TMA producer
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0;
      for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
        const int z = t % g.batch;
        const int rest = t / g.batch;
        const int nt = rest % g.num_n_tiles, rt = rest / g.num_n_tiles;
        const int a_row = g.a_row0 + rt * BM;
        const int a_col = z * g.a_k_off_per_z;
        const int b_row = z * g.b_row_off_per_z + nt * g.BN;
        for (int kb = 0; kb < KB; kb++) {
          mbar_wait(empty_bar(stage), phase ^ 1u);
          const uint32_t sA = base + stage * stage_bytes;
          const uint32_t sB = sA + planes * kAStageBytes;
          mbar_expect_tx(full_bar(stage), stage_bytes);
          tma_load_2d(&tmA_hi, full_bar(stage), sA, a_col + kb * BK, a_row);
          tma_load_2d(&tmB_hi, full_bar(stage), sB, kb * BK, b_row);
          if (g.three_pass) {
            tma_load_2d(&tmA_lo, full_bar(stage), sA + kAStageBytes, a_col + kb * BK, a_row);
            tma_load_2d(&tmB_lo, full_bar(stage), sB + b_bytes, kb * BK, b_row);
          }
          if (++stage == g.stages) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // ================================================================= MMA issuer
    if (lane == 0) {
      const uint32_t idesc = umma_idesc_tf32(BM, g.BN);
      int stage = 0; uint32_t phase = 0;
      int it = 0;
      for (int t = blockIdx.x; t < num_tiles; t += gridDim.x, it++) {
        const int as = it & 1;
        const uint32_t aphase = (uint32_t)(it >> 1) & 1u;
        mbar_wait(tempty_bar(as), aphase ^ 1u);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + (uint32_t)as * 256u;
        for (int kb = 0; kb < KB; kb++) {
          mbar_wait(full_bar(stage), phase);
          tc_fence_after();
          const uint32_t sA = base + stage * stage_bytes;
          const uint32_t sB = sA + planes * kAStageBytes;
          const uint64_t dA_hi = umma_desc_sw128(sA), dB_hi = umma_desc_sw128(sB);
          const uint64_t dA_lo = umma_desc_sw128(sA + kAStageBytes), dB_lo = umma_desc_sw128(sB + b_bytes);
#pragma unroll
          for (int k = 0; k < BK / 8; k++) {
            const uint64_t adv = (uint64_t)((k * 32) >> 4);   // 8 tf32 = 32 bytes along K inside the swizzle row
            umma_tf32(tmem_d, dA_hi + adv, dB_hi + adv, idesc, (kb | k) != 0 ? 1u : 0u);
            if (g.three_pass) {
              umma_tf32(tmem_d, dA_lo + adv, dB_hi + adv, idesc, 1u);
              umma_tf32(tmem_d, dA_hi + adv, dB_lo + adv, idesc, 1u);
            }
          }
          umma_commit(empty_bar(stage));      // frees the smem slot when these MMAs retire
          if (++stage == g.stages) { stage = 0; phase ^= 1u; }
        }
        umma_commit(tfull_bar(as));           // accumulator complete -> epilogue
      }
    }
  } else if (warp >= 4) {
    // ================================================================= epilogue (warp w owns TMEM lanes 32*(w%4)..)
    const int q = warp & 3;
    int it = 0;
    for (int t = blockIdx.x; t < num_tiles; t += gridDim.x, it++) {
      const int z = t % g.batch;
      const int rest = t / g.batch;
      const int nt = rest % g.num_n_tiles, rt = rest / g.num_n_tiles;
      const int as = it & 1;
      const uint32_t aphase = (uint32_t)(it >> 1) & 1u;
      mbar_wait(tfull_bar(as), aphase);
      tc_fence_after();
      const int r = rt * BM + q * 32 + lane;          // row inside this problem
      const bool row_ok = r < g.R;
      const uint32_t taddr0 = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)as * 256u;
      for (int c0 = 0; c0 < g.BN; c0 += 16) {
        uint32_t v[16];
        tmem_ld16(taddr0 + (uint32_t)c0, v);
        const int n0 = nt * g.BN + c0;                // column inside this member's output
        if (!row_ok || n0 >= g.N) continue;
        if (EPI == EPI_F) {
          const float* bias = g.bias + (size_t)z * g.bias_stride_z + n0;
          float* oh = g.out_hi + (size_t)r * g.ld_out + (size_t)z * g.out_col_off_per_z + n0;
          float* ol = g.out_lo + (size_t)r * g.ld_out + (size_t)z * g.out_col_off_per_z + n0;
#pragma unroll
          for (int c = 0; c < 16; c += 4) {
            float4 hi4, lo4;
            float hh[4];
#pragma unroll
            for (int e = 0; e < 4; e++) {
              float zz = __uint_as_float(v[c + e]) + __ldg(bias + c + e);
              hh[e] = zz > 0.f ? zz : 0.1f * (exp2f(zz * 14.426950408889634f) - 1.f);
            }
            hi4.x = rna_tf32(hh[0]); hi4.y = rna_tf32(hh[1]); hi4.z = rna_tf32(hh[2]); hi4.w = rna_tf32(hh[3]);
            lo4.x = rna_tf32(hh[0] - hi4.x); lo4.y = rna_tf32(hh[1] - hi4.y); lo4.z = rna_tf32(hh[2] - hi4.z);
            lo4.w = rna_tf32(hh[3] - hi4.w);
            *(float4*)(oh + c) = hi4;
            *(float4*)(ol + c) = lo4;
          }
        } else if (EPI == EPI_B) {
          const float* xh = g.aux_hi + (size_t)r * g.ld_aux + (size_t)z * g.aux_col_off_per_z + n0;
          const float* xl = g.aux_lo + (size_t)r * g.ld_aux + (size_t)z * g.aux_col_off_per_z + n0;
          float* oh = g.out_hi + (size_t)r * g.ld_out + (size_t)z * g.out_col_off_per_z + n0;
          float* ol = g.out_lo + (size_t)r * g.ld_out + (size_t)z * g.out_col_off_per_z + n0;
#pragma unroll
          for (int c = 0; c < 16; c += 4) {
            const float4 a = *(const float4*)(xh + c), b = *(const float4*)(xl + c);
            const float hsum[4] = {a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w};
            float gg[4];
#pragma unroll
            for (int e = 0; e < 4; e++) {
              const float d = hsum[e] > 0.f ? 1.f : fmaf(10.f, hsum[e], 1.f);   // CELU'(z) from h = CELU(z)
              gg[e] = __uint_as_float(v[c + e]) * d;
            }
            float4 hi4, lo4;
            hi4.x = rna_tf32(gg[0]); hi4.y = rna_tf32(gg[1]); hi4.z = rna_tf32(gg[2]); hi4.w = rna_tf32(gg[3]);
            lo4.x = rna_tf32(gg[0] - hi4.x); lo4.y = rna_tf32(gg[1] - hi4.y); lo4.z = rna_tf32(gg[2] - hi4.z);
            lo4.w = rna_tf32(gg[3] - hi4.w);
            *(float4*)(oh + c) = hi4;
            *(float4*)(ol + c) = lo4;
          }
        } else {
          float* o = g.out_hi + (size_t)r * g.ld_out + n0;
#pragma unroll
          for (int c = 0; c < 16; c += 4) {
            float4 o4;
            o4.x = __uint_as_float(v[c + 0]) * g.scale; o4.y = __uint_as_float(v[c + 1]) * g.scale;
            o4.z = __uint_as_float(v[c + 2]) * g.scale; o4.w = __uint_as_float(v[c + 3]) * g.scale;
            *(float4*)(o + c) = o4;
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(tempty_bar(as));
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
  }
}

// Output layer on planes, one warp per atom row (same math as mlp_out_kernel): e_m = b + <h, w>, G = split(w * CELU'(z)).
__global__ void __launch_bounds__(128)
tc_out_kernel(int R, int M, int h, const float* __restrict__ Hh, const float* __restrict__ Hl, const float* __restrict__ w4,
              const float* __restrict__ b4, float* __restrict__ Gh, float* __restrict__ Gl, double* __restrict__ eatom,
              double sae) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int r = blockIdx.x * 4 + warp;
  if (r >= R) return;
  double esum = 0.0;
  for (int m = 0; m < M; m++) {
    const size_t off = ((size_t)r * M + m) * h;
    const float* w = w4 + (size_t)m * h;
    float acc = 0.f;
    for (int n = lane; n < h; n += 32) {
      const float hv = Hh[off + n] + Hl[off + n];
      const float wv = w[n];
      acc = fmaf(hv, wv, acc);
      const float gv = wv * (hv > 0.f ? 1.f : fmaf(10.f, hv, 1.f));
      const float hi = rna_tf32(gv);
      Gh[off + n] = hi;
      Gl[off + n] = rna_tf32(gv - hi);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    esum += (double)(acc + b4[m]);
  }
  if (lane == 0) eatom[r] = esum / (double)M + sae;
}

// ------------------------------------------------------------------------------------------------ host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

inline EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = (EncodeTiledFn)p;
  }
  return fn;
}

// 2-D fp32 tensor [rows][cols] with row pitch ld (elements); box = 32 columns x box_rows rows; SWIZZLE_128B.
inline int make_tmap(CUtensorMap* tm, const float* ptr, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows) {
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) { set_error("cuTensorMapEncodeTiled is not available from the driver"); return 1; }
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {ld * sizeof(float)};
  cuuint32_t box[2] = {(cuuint32_t)BK, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, (void*)ptr, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled failed with code " + std::to_string((int)r)); return 1; }
  return 0;
}

inline float host_rna_tf32(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  u += 0x1000u;
  u &= 0xFFFFE000u;
  float o;
  memcpy(&o, &u, 4);
  return o;
}

inline int upload_planes(const std::vector<float>& w, DevBufTc& hi, DevBufTc& lo) {
  std::vector<float> h(w.size()), l(w.size());
  for (size_t i = 0; i < w.size(); i++) { h[i] = host_rna_tf32(w[i]); l[i] = host_rna_tf32(w[i] - h[i]); }
  if (!hi.reserve(w.size() * 4) || !lo.reserve(w.size() * 4)) { set_error("out of memory for weight planes"); return 1; }
  ANIB_CUDA_OK(cudaMemcpy(hi.p, h.data(), w.size() * 4, cudaMemcpyHostToDevice));
  ANIB_CUDA_OK(cudaMemcpy(lo.p, l.data(), w.size() * 4, cudaMemcpyHostToDevice));
  return 0;
}

inline int upload_f32(const std::vector<float>& w, DevBufTc& d) {
  if (!d.reserve(std::max<size_t>(w.size(), 1) * 4)) { set_error("out of memory for weights"); return 1; }
  ANIB_CUDA_OK(cudaMemcpy(d.p, w.data(), w.size() * 4, cudaMemcpyHostToDevice));
  return 0;
}

// largest multiple of 16 that divides n and is <= 256 (UMMA N); 0 if none
inline int pick_bn(int n) {
  for (int bn = 256; bn >= 16; bn -= 16)
    if (n % bn == 0) return bn;
  return 0;
}

inline bool tc_supported_dims(const std::vector<int>& dims) {
  const int nh = (int)dims.size() - 2;
  if (nh < 1) return false;
  for (int l = 1; l <= nh; l++)
    if (dims[l] % 32 != 0 || dims[l] > 256) return false;   // member K ranges must be whole 128-byte rows
  if (dims[0] % 8 != 0 || pick_bn(dims[0]) == 0) return false;
  return true;
}

struct Launcher {
  cudaStream_t st;
  int num_sms;
  int* launches;
  int max_smem;
};

template <int EPI>
inline int launch_gemm(const Launcher& L, const float* A_hi, const float* A_lo, uint64_t a_rows, uint64_t a_cols, uint64_t a_ld,
                       const float* B_hi, const float* B_lo, uint64_t b_rows, uint64_t b_cols, uint64_t b_ld, GemmArgs g) {
  CUtensorMap tA_hi, tA_lo, tB_hi, tB_lo;
  if (make_tmap(&tA_hi, A_hi, a_rows, a_cols, a_ld, BM)) return 1;
  if (make_tmap(&tA_lo, A_lo ? A_lo : A_hi, a_rows, a_cols, a_ld, BM)) return 1;
  if (make_tmap(&tB_hi, B_hi, b_rows, b_cols, b_ld, (uint32_t)g.BN)) return 1;
  if (make_tmap(&tB_lo, B_lo ? B_lo : B_hi, b_rows, b_cols, b_ld, (uint32_t)g.BN)) return 1;
  const int planes = g.three_pass ? 2 : 1;
  const int stage_bytes = planes * (kAStageBytes + g.BN * 128);
  int stages = (L.max_smem - 2048) / stage_bytes;
  stages = std::min(stages, kMaxStages);
  if (stages < 2) { set_error("not enough shared memory for the tensor-core GEMM pipeline"); return 1; }
  g.stages = stages;
  const int smem = stages * stage_bytes + 1024 /*alignment*/ + 512 /*barriers*/;
  ANIB_CUDA_OK(cudaFuncSetAttribute(tc_gemm_kernel<EPI>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  const int tiles = g.num_row_tiles * g.num_n_tiles * g.batch;
  const int grid = std::max(1, std::min(tiles, L.num_sms));
  tc_gemm_kernel<EPI><<<grid, kThreads, smem, L.st>>>(tA_hi, tA_lo, tB_hi, tB_lo, g);
  ANIB_CUDA_OK(cudaGetLastError());
  if (L.launches) (*L.launches)++;
  return 0;
}

}  // namespace tc

inline int64_t tc_padded_rows(int nlocal, int /*S*/) { return (int64_t)nlocal + 128; }

// Build the TF32 planes of one species' weights.  dims: in, h1..h_nh, 1;  hm weights W[m][s][l] row-major [out][in].
inline int tc_prepare_species(const HostModel& hm, int M, int s, TcSpeciesWeights& w) {
  const std::vector<int>& dims = hm.dims[s];
  if (!tc::tc_supported_dims(dims)) { w.ready = false; return 0; }
  const int nh = (int)dims.size() - 2;
  w.nh = nh;
  w.Wf_hi.resize(nh); w.Wf_lo.resize(nh); w.WT_hi.resize(nh); w.WT_lo.resize(nh); w.bias.resize(nh);
  for (int j = 0; j < nh; j++) {
    const int ni = dims[j], no = dims[j + 1];
    std::vector<float> f((size_t)M * no * ni), b((size_t)M * no);
    for (int m = 0; m < M; m++) {
      memcpy(f.data() + (size_t)m * no * ni, hm.W[m][s][j].data(), (size_t)no * ni * 4);
      memcpy(b.data() + (size_t)m * no, hm.b[m][s][j].data(), (size_t)no * 4);
    }
    if (tc::upload_planes(f, w.Wf_hi[j], w.Wf_lo[j])) return 1;
    if (tc::upload_f32(b, w.bias[j])) return 1;
    if (j >= 1) {
      std::vector<float> t((size_t)M * ni * no);
      for (int m = 0; m < M; m++)
        for (int o = 0; o < no; o++)
          for (int k = 0; k < ni; k++) t[((size_t)m * ni + k) * no + o] = hm.W[m][s][j][(size_t)o * ni + k];
      if (tc::upload_planes(t, w.WT_hi[j], w.WT_lo[j])) return 1;
    }
  }
  {
    const int L = dims[0], h1 = dims[1];
    std::vector<float> cat((size_t)L * M * h1);
    for (int m = 0; m < M; m++)
      for (int k = 0; k < h1; k++)
        for (int q = 0; q < L; q++) cat[(size_t)q * M * h1 + (size_t)m * h1 + k] = hm.W[m][s][0][(size_t)k * L + q];
    if (tc::upload_planes(cat, w.Wcat_hi, w.Wcat_lo)) return 1;
  }
  {
    const int h = dims[nh];
    std::vector<float> wo((size_t)M * h), bo(M);
    for (int m = 0; m < M; m++) {
      memcpy(wo.data() + (size_t)m * h, hm.W[m][s][nh].data(), (size_t)h * 4);
      bo[m] = hm.b[m][s][nh][0];
    }
    if (tc::upload_f32(wo, w.w_out) || tc::upload_f32(bo, w.b_out)) return 1;
  }
  w.ready = true;
  return 0;
}

// Forward + backward of one species bucket.  aev planes: [*, ld_aev] in bucket order; rows [row0, row0+R).
inline int tc_mlp_species(cudaStream_t st, bool three_pass, TcSpeciesWeights& w, TcWork& work, int M,
                          const std::vector<int>& dims, const float* aev_hi, const float* aev_lo, int64_t ld_aev,
                          int64_t aev_rows_total, int row0, int R, float* daev, double* eatom, double sae, int* launches) {
  using namespace tc;
  if (!w.ready) { set_error("this network's layer widths are not supported by the tensor-core path (need multiples of 32, <= 256)"); return 1; }
  static int num_sms = 0, max_smem = 0;
  if (!num_sms) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, dev);
    cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
  }
  Launcher L{st, num_sms, launches, max_smem};
  const int nh = w.nh;
  // chunk: a whole number of 128-row tiles, about two waves of (row tile x member) tiles
  int chunk = work.chunk_rows;
  if (chunk == 0) {
    const char* e = getenv("ANI_B200_TC_CHUNK");
    int row_tiles = std::max(1, (2 * num_sms + M - 1) / M);
    chunk = (e && atoi(e) > 0) ? (atoi(e) + BM - 1) / BM * BM : row_tiles * BM;
    work.chunk_rows = chunk;
  }
  if ((int)work.H_hi.size() < nh + 1) { work.H_hi.resize(nh + 1); work.H_lo.resize(nh + 1); work.G_hi.resize(nh + 1); work.G_lo.resize(nh + 1); }
  for (int l = 1; l <= nh; l++) {
    const size_t bytes = (size_t)chunk * M * dims[l] * 4;
    if (!work.H_hi[l].reserve(bytes) || !work.H_lo[l].reserve(bytes) || !work.G_hi[l].reserve(bytes) || !work.G_lo[l].reserve(bytes)) {
      set_error("out of memory for tensor-core MLP work buffers");
      return 1;
    }
  }
  const int Lin = dims[0];
  for (int r0 = 0; r0 < R; r0 += chunk) {
    const int Rc = std::min(chunk, R - r0);
    const int rtiles = (Rc + BM - 1) / BM;
    // ---- forward hidden layers
    for (int j = 0; j < nh; j++) {
      const int ni = dims[j], no = dims[j + 1];
      GemmArgs g{};
      g.R = Rc; g.N = no; g.BN = no; g.num_n_tiles = 1; g.num_row_tiles = rtiles; g.K = ni; g.batch = M;
      g.b_row_off_per_z = no; g.three_pass = three_pass ? 1 : 0;
      g.bias = w.bias[j].f(); g.bias_stride_z = no;
      g.out_hi = work.H_hi[j + 1].f(); g.out_lo = work.H_lo[j + 1].f(); g.ld_out = (long long)M * no; g.out_col_off_per_z = no;
      int rc;
      if (j == 0) {
        g.a_row0 = row0 + r0; g.a_k_off_per_z = 0;
        rc = launch_gemm<EPI_F>(L, aev_hi, aev_lo, (uint64_t)aev_rows_total, (uint64_t)Lin, (uint64_t)ld_aev, w.Wf_hi[0].f(),
                                w.Wf_lo[0].f(), (uint64_t)M * no, (uint64_t)ni, (uint64_t)ni, g);
      } else {
        g.a_row0 = 0; g.a_k_off_per_z = ni;
        rc = launch_gemm<EPI_F>(L, work.H_hi[j].f(), work.H_lo[j].f(), (uint64_t)chunk, (uint64_t)M * ni, (uint64_t)M * ni,
                                w.Wf_hi[j].f(), w.Wf_lo[j].f(), (uint64_t)M * no, (uint64_t)ni, (uint64_t)ni, g);
      }
      if (rc) return rc;
    }
    // ---- output layer, energies, dE/dz of the last hidden layer
    tc_out_kernel<<<(Rc + 3) / 4, 128, 0, st>>>(Rc, M, dims[nh], work.H_hi[nh].f(), work.H_lo[nh].f(), w.w_out.f(), w.b_out.f(),
                                               work.G_hi[nh].f(), work.G_lo[nh].f(), eatom + row0 + r0, sae);
    if (launches) (*launches)++;
    // ---- backward through hidden layers nh-1 .. 1
    for (int j = nh - 1; j >= 1; j--) {
      const int ni = dims[j], no = dims[j + 1];   // layer j: ni -> no ; gradient: no -> ni
      GemmArgs g{};
      g.R = Rc; g.N = ni; g.BN = ni; g.num_n_tiles = 1; g.num_row_tiles = rtiles; g.K = no; g.batch = M;
      g.a_row0 = 0; g.a_k_off_per_z = no; g.b_row_off_per_z = ni; g.three_pass = three_pass ? 1 : 0;
      g.aux_hi = work.H_hi[j].f(); g.aux_lo = work.H_lo[j].f(); g.ld_aux = (long long)M * ni; g.aux_col_off_per_z = ni;
      g.out_hi = work.G_hi[j].f(); g.out_lo = work.G_lo[j].f(); g.ld_out = (long long)M * ni; g.out_col_off_per_z = ni;
      if (launch_gemm<EPI_B>(L, work.G_hi[j + 1].f(), work.G_lo[j + 1].f(), (uint64_t)chunk, (uint64_t)M * no, (uint64_t)M * no,
                             w.WT_hi[j].f(), w.WT_lo[j].f(), (uint64_t)M * ni, (uint64_t)no, (uint64_t)no, g))
        return 1;
    }
    // ---- dE/dAEV = (1/M) * G1[Rc][M*h1] . Wcat[L][M*h1]^T   (ensemble mean folded into K)
    {
      const int h1 = dims[1];
      GemmArgs g{};
      g.BN = pick_bn(Lin);
      g.R = Rc; g.N = Lin; g.num_n_tiles = Lin / g.BN; g.num_row_tiles = rtiles; g.K = M * h1; g.batch = 1;
      g.a_row0 = 0; g.three_pass = three_pass ? 1 : 0;
      g.out_hi = daev + (size_t)(row0 + r0) * ld_aev; g.ld_out = (long long)ld_aev; g.scale = 1.0f / (float)M;
      if (launch_gemm<EPI_D>(L, work.G_hi[1].f(), work.G_lo[1].f(), (uint64_t)chunk, (uint64_t)M * h1, (uint64_t)M * h1,
                             w.Wcat_hi.f(), w.Wcat_lo.f(), (uint64_t)Lin, (uint64_t)M * h1, (uint64_t)M * h1, g))
        return 1;
    }
  }
  return 0;
}

}  // namespace anib
