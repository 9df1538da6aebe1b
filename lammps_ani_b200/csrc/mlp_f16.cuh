// mlp_f16.cuh -- the ensemble MLP on tcgen05 with fp16 hi/lo operand planes ("f16x3"): fp32-level accuracy at the
// fp16 tensor rate.
//
// Replaces BmmEnsemble's torch.bmm -> cuBLAS SGEMM chain, its bias / CELU kernels and their autograd backward
// (models/lammps_ani.py:219-257,203-214; K5/K7 of SURVEY.md 2.1).
//
// Arithmetic.  Every GEMM operand x is kept as two fp16 planes of the SCALED value s*x (s a power of two chosen per
// tensor so that the values sit in the middle of fp16's exponent range):  hi = rn16(s x), lo = rn16(s x - hi).  A product
// is issued as  A_hi.B_hi + A_lo.B_hi + A_hi.B_lo  (three kind::f16 MMAs, fp32 accumulate in TMEM); 11-bit x 11-bit
// mantissa products are exact in fp32 and the dropped lo.lo term is 2^-22 relative, the same error budget as the TF32x3
// path (mlp_tc.cuh) at twice the MMA rate and half the operand bytes.  The power-of-two scales are folded back exactly in
// the epilogues.
//
// One persistent warp-specialised kernel does every GEMM:
//   warp 0   : TMA producer (cp.async.bulk.tensor.2d, SWIZZLE_128B rows of 64 halves, mbarrier tx-count)
//   warp 1   : MMA issuer   (one lane; UMMA M=128, N = tile width <= 256, K=16 per instruction; tcgen05.commit)
//   warp 2   : TMEM allocator (512 columns = two accumulator stages, so the epilogue of tile t overlaps the MMAs of t+1)
//   warps 4-7: epilogue: tcgen05.ld 32 columns -> fused math in registers -> swizzled shared-memory staging ->
//              cp.async.bulk.tensor store (coalesced, row-clipped by the tensor map).  The backward epilogue's second
//              input (the activations, for CELU') arrives by TMA loads prefetched one column block ahead.
// Epilogues: EPI_F  h = CELU(acc + b)            -> H planes
//            EPI_FO last hidden layer + output layer: e_m = b4 + <h, w4>, g = w4 * CELU'(z)   -> e[row][m], G planes
//            EPI_B  g = acc * CELU'(z)  (z from the stored H planes)                          -> G planes
//            EPI_D  dE/dAEV = acc / M   (ensemble mean folded into K = M*h1)                   -> fp32 rows
#pragma once
#include <cuda.h>
#include <cuda_fp16.h>

#include "common.cuh"
#include "mlp_tc.cuh"

namespace anib {
namespace f16 {

using tc::mbar_arrive; using tc::mbar_expect_tx; using tc::mbar_init; using tc::mbar_wait; using tc::smem_u32;
using tc::tc_fence_after; using tc::tc_fence_before; using tc::tma_load_2d; using tc::umma_commit; using tc::umma_desc_sw128;

constexpr int BM = 128;
constexpr int kMaxThreads = 384;             // 4 control warps + up to 8 epilogue warps
constexpr int kMaxStages = 8;
constexpr int kStageBuf = 4096;              // one epilogue staging buffer (32 rows x 128 B, or 2 x 32 rows x 64 B)
constexpr int kMaxEpiWarps = 8;
enum { EPI_F = 0, EPI_FO = 1, EPI_B = 2, EPI_D = 3 };

// power-of-two operand scales (see file header)
constexpr float kScaleAev = 16.f, kScaleH = 16.f, kScaleG = 256.f;

struct Args {
  int R, N, BN, num_n_tiles, num_row_tiles, K, batch;
  int a_row0, a_k_off_per_z, b_row_off_per_z;
  int stages, swz;
  int bk;                // halves per k-block: 64 (128-byte rows, SWIZZLE_128B) or 32 (64-byte rows, SWIZZLE_64B)
  int epi_warps;         // 4 or 8: with 8, the two warps that own a TMEM lane quadrant take alternate 32-column blocks
  int out_bufs;          // staging buffers per epilogue warp for the TMA stores (1 or 2)
  int out_col_off_per_z, aux_col_off_per_z;
  float acc_scale;       // accumulator -> true value (x out scale for EPI_B / EPI_D)
  float out_scale;       // EPI_F: scale of the stored H planes; EPI_FO: scale of the stored G planes
  float aux_inv_scale;   // EPI_B: 1 / scale of the H planes
  const float* bias; int bias_stride_z;
  const float* w_out; const float* b_out;     // EPI_FO: [batch][N], [batch]
  float* e_out; int e_ld;                     // EPI_FO: e_out[row * e_ld + z]
};

__device__ __forceinline__ void umma_f16(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
      "}\n" ::"r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(accumulate) : "memory");
}
// instruction descriptor: D = f32 (1 at bits 4-5), A = B = F16 (0 at bits 7-9, 10-12), both K-major, N>>3 at 17, M>>4 at 24
__device__ __forceinline__ uint32_t umma_idesc_f16(int M, int N) {
  return (1u << 4) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
// K-major operand tile of 128-byte rows (SWIZZLE_128B, 8-row groups 1024 B apart) or 64-byte rows (SWIZZLE_64B, 512 B apart);
// descriptor version 1 (sm_100); layout type 2 / 4 (cute/arch/mma_sm100_desc.hpp)
__device__ __forceinline__ uint64_t umma_desc_k(uint32_t smem_addr, int bk) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFF) >> 4);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)((bk == 64 ? 1024 : 512) >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)(bk == 64 ? 2 : 4) << 61;
  return d;
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,"
      "%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
        "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]),
        "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]),
        "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tma_store_2d(const CUtensorMap* tm, uint32_t src, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"(tm), "r"(src), "r"(c0), "r"(c1)
               : "memory");
}
__device__ __forceinline__ void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void tma_store_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void tma_store_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void sts128(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ void lds128(uint32_t addr, uint32_t& a, uint32_t& b, uint32_t& c, uint32_t& d) {
  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(a), "=r"(b), "=r"(c), "=r"(d) : "r"(addr) : "memory");
}
// split x into fp16 hi + lo and pack two neighbouring columns per 32-bit word
// (packed conversions: F2FP.F16.F32.PACK_AB runs on the ALU, the scalar F2F.F16.F32 on the XU pipe, which the forward
//  epilogues saturated -- profiles/README.md; results are bit-identical)
__device__ __forceinline__ void split2(float x0, float x1, uint32_t& hi, uint32_t& lo) {
  const __half2 h = __floats2half2_rn(x0, x1);
  const float2 hf = __half22float2(h);
  const __half2 l = __floats2half2_rn(x0 - hf.x, x1 - hf.y);
  hi = *reinterpret_cast<const uint32_t*>(&h);
  lo = *reinterpret_cast<const uint32_t*>(&l);
}
__device__ __forceinline__ float unpack_sum(uint32_t hi, uint32_t lo, int which) {
  const unsigned short h = which ? (unsigned short)(hi >> 16) : (unsigned short)(hi & 0xFFFFu);
  const unsigned short l = which ? (unsigned short)(lo >> 16) : (unsigned short)(lo & 0xFFFFu);
  return __half2float(__ushort_as_half(h)) + __half2float(__ushort_as_half(l));
}

// byte offset of 16-byte chunk `ch` of row `r` inside a staging box whose rows are `row_bytes` (64 or 128) long;
// swz: the tensor map uses SWIZZLE_64B / SWIZZLE_128B (chunk index XOR-ed with address bits 7.. as the TMA unit does)
__device__ __forceinline__ uint32_t stage_off(int r, int ch, int row_bytes, int swz) {
  const int x = swz ? (row_bytes == 128 ? (r & 7) : ((r >> 1) & 3)) : 0;
  return (uint32_t)(r * row_bytes + ((ch ^ x) << 4));
}

template <int EPI>
__global__ void __launch_bounds__(kMaxThreads, 1)
f16_gemm_kernel(const __grid_constant__ CUtensorMap tmA_hi, const __grid_constant__ CUtensorMap tmA_lo,
                const __grid_constant__ CUtensorMap tmB_hi, const __grid_constant__ CUtensorMap tmB_lo,
                const __grid_constant__ CUtensorMap tmO_hi, const __grid_constant__ CUtensorMap tmO_lo,
                const __grid_constant__ CUtensorMap tmX_hi, const __grid_constant__ CUtensorMap tmX_lo, const Args g) {
  extern __shared__ unsigned char smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  const uint32_t row_bytes = (uint32_t)g.bk * 2u;                 // bytes of one operand row inside a k-block
  const uint32_t a_bytes = (uint32_t)BM * row_bytes;
  const uint32_t b_bytes = (uint32_t)g.BN * row_bytes;
  const uint32_t stage_bytes = 2u * (a_bytes + b_bytes);
  const uint32_t base_out = base + g.stages * stage_bytes;
  const uint32_t base_aux = base_out + g.epi_warps * g.out_bufs * kStageBuf;
  const uint32_t bar_base = base_aux + (EPI == EPI_B ? g.epi_warps * 2 * kStageBuf : 0);
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (kMaxStages + s); };
  auto tfull_bar = [&](int s) { return bar_base + 8u * (2 * kMaxStages + s); };
  auto tempty_bar = [&](int s) { return bar_base + 8u * (2 * kMaxStages + 2 + s); };
  auto aux_bar = [&](int w, int b) { return bar_base + 8u * (2 * kMaxStages + 4 + 2 * w + b); };
  const uint32_t tmem_slot = bar_base + 8u * (2 * kMaxStages + 4 + 2 * kMaxEpiWarps);
  volatile uint32_t* tmem_slot_ptr = (volatile uint32_t*)(smem_raw + (tmem_slot - raw));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int num_tiles = g.num_row_tiles * g.num_n_tiles * g.batch;
  const int KB = (g.K + g.bk - 1) / g.bk;

  if (warp == 1 && lane == 0) {
    for (int s = 0; s < g.stages; s++) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
    for (int s = 0; s < 2; s++) { mbar_init(tfull_bar(s), 1); mbar_init(tempty_bar(s), g.epi_warps); }
    for (int w = 0; w < g.epi_warps; w++) { mbar_init(aux_bar(w, 0), 1); mbar_init(aux_bar(w, 1), 1); }
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
          const uint32_t sB = sA + 2 * a_bytes;
          mbar_expect_tx(full_bar(stage), stage_bytes);
          tma_load_2d(&tmA_hi, full_bar(stage), sA, a_col + kb * g.bk, a_row);
          tma_load_2d(&tmB_hi, full_bar(stage), sB, kb * g.bk, b_row);
          tma_load_2d(&tmA_lo, full_bar(stage), sA + a_bytes, a_col + kb * g.bk, a_row);
          tma_load_2d(&tmB_lo, full_bar(stage), sB + b_bytes, kb * g.bk, b_row);
          if (++stage == g.stages) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // ================================================================= MMA issuer
    if (lane == 0) {
      const uint32_t idesc = umma_idesc_f16(BM, g.BN);
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
          const uint32_t sB = sA + 2 * a_bytes;
          const uint64_t dA_hi = umma_desc_k(sA, g.bk), dB_hi = umma_desc_k(sB, g.bk);
          const uint64_t dA_lo = umma_desc_k(sA + a_bytes, g.bk), dB_lo = umma_desc_k(sB + b_bytes, g.bk);
          const int nks = min(g.bk / 16, (g.K - kb * g.bk) / 16);   // ragged last k-block: only the valid 16-wide steps
          for (int k = 0; k < nks; k++) {
            const uint64_t adv = (uint64_t)((k * 32) >> 4);       // 16 halves = 32 bytes along K inside the swizzle row
            umma_f16(tmem_d, dA_hi + adv, dB_hi + adv, idesc, (kb | k) != 0 ? 1u : 0u);
            umma_f16(tmem_d, dA_lo + adv, dB_hi + adv, idesc, 1u);
            umma_f16(tmem_d, dA_hi + adv, dB_lo + adv, idesc, 1u);
          }
          umma_commit(empty_bar(stage));      // frees the smem slot when these MMAs retire
          if (++stage == g.stages) { stage = 0; phase ^= 1u; }
        }
        umma_commit(tfull_bar(as));           // accumulator complete -> epilogue
      }
    }
  } else if (warp >= 4 && warp - 4 < g.epi_warps) {
    // ================================================================= epilogue (warp w owns TMEM lanes 32*(w%4)..)
    const int q = warp & 3;                                  // TMEM lane quadrant this warp may read
    const int ew = warp - 4, half = ew >> 2, nsplit = g.epi_warps >> 2;
    const uint32_t my_out = base_out + (uint32_t)ew * g.out_bufs * kStageBuf;
    const uint32_t my_aux = base_aux + (uint32_t)ew * 2 * kStageBuf;
    uint32_t aux_issued = 0, aux_used = 0, store_cnt = 0;
    int it = 0;
    for (int t = blockIdx.x; t < num_tiles; t += gridDim.x, it++) {
      const int z = t % g.batch;
      const int rest = t / g.batch;
      const int nt = rest % g.num_n_tiles, rt = rest / g.num_n_tiles;
      const int as = it & 1;
      const uint32_t aphase = (uint32_t)(it >> 1) & 1u;
      const int trow = rt * BM + q * 32;                       // first row of this warp inside the problem
      const int ocol = z * g.out_col_off_per_z + nt * g.BN;
      const int xcol = z * g.aux_col_off_per_z + nt * g.BN;
      const int nvalid = min(g.BN / 32, (g.N - nt * g.BN + 31) / 32);
      auto issue_aux = [&](int cb) {
        const uint32_t buf = my_aux + (aux_issued & 1u) * kStageBuf;
        const uint32_t bar = aux_bar(ew, aux_issued & 1u);
        mbar_expect_tx(bar, kStageBuf);
        tma_load_2d(&tmX_hi, bar, buf, xcol + cb * 32, trow);
        tma_load_2d(&tmX_lo, bar, buf + kStageBuf / 2, xcol + cb * 32, trow);
      };
      if (EPI == EPI_B && half < nvalid) {
        if (lane == 0) issue_aux(half);
        aux_issued++;
      }
      mbar_wait(tfull_bar(as), aphase);
      tc_fence_after();
      const uint32_t taddr0 = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)as * 256u;
      float e_acc = 0.f;
      for (int cb = half; cb < nvalid; cb += nsplit) {
        if (EPI == EPI_B && cb + nsplit < nvalid) {
          if (lane == 0) issue_aux(cb + nsplit);
          aux_issued++;
        }
        uint32_t v[32];
        tmem_ld32(taddr0 + (uint32_t)(cb * 32), v);
        const int n0 = nt * g.BN + cb * 32;                    // column inside this member's output
        float o[32];
        if (EPI == EPI_F || EPI == EPI_FO) {
          const float4* bias4 = (const float4*)(g.bias + (size_t)z * g.bias_stride_z + n0);   // 128-byte aligned: widths are multiples of 32
          const float4* wo4 = (EPI == EPI_FO) ? (const float4*)(g.w_out + (size_t)z * g.N + n0) : nullptr;
          const float neg_a = 0.1f * g.out_scale;
#pragma unroll
          for (int c4 = 0; c4 < 8; c4++) {
            const float4 bq = __ldg(bias4 + c4);
            const float bb[4] = {bq.x, bq.y, bq.z, bq.w};
            float ww[4] = {0.f, 0.f, 0.f, 0.f};
            if (EPI == EPI_FO) { const float4 wq = __ldg(wo4 + c4); ww[0] = wq.x; ww[1] = wq.y; ww[2] = wq.z; ww[3] = wq.w; }
#pragma unroll
            for (int k = 0; k < 4; k++) {
              const int c = 4 * c4 + k;
              const float zz = fmaf(__uint_as_float(v[c]), g.acc_scale, bb[k]);
              const float ex = exp2f(zz * 14.426950408889634f);  // e^(z/alpha), alpha = 0.1
              if (EPI == EPI_F) {
                o[c] = zz > 0.f ? zz * g.out_scale : fmaf(neg_a, ex, -neg_a);
              } else {
                const float h = zz > 0.f ? zz : 0.1f * (ex - 1.f);
                e_acc = fmaf(h, ww[k], e_acc);
                o[c] = ww[k] * (zz > 0.f ? 1.f : ex) * g.out_scale;
              }
            }
          }
        } else if (EPI == EPI_B) {
          const uint32_t buf = my_aux + (aux_used & 1u) * kStageBuf;
          mbar_wait(aux_bar(ew, aux_used & 1u), (aux_used >> 1) & 1u);
          aux_used++;
#pragma unroll
          for (int ch = 0; ch < 4; ch++) {
            uint32_t hh[4], ll[4];
            const uint32_t off = stage_off(lane, ch, 64, g.swz);
            lds128(buf + off, hh[0], hh[1], hh[2], hh[3]);
            lds128(buf + kStageBuf / 2 + off, ll[0], ll[1], ll[2], ll[3]);
#pragma unroll
            for (int e = 0; e < 8; e++) {
              const float hs = unpack_sum(hh[e >> 1], ll[e >> 1], e & 1) * g.aux_inv_scale;
              const float d = hs > 0.f ? 1.f : fmaf(10.f, hs, 1.f);   // CELU'(z) from h = CELU(z)
              o[ch * 8 + e] = __uint_as_float(v[ch * 8 + e]) * g.acc_scale * d;
            }
          }
        } else {
#pragma unroll
          for (int c = 0; c < 32; c++) o[c] = __uint_as_float(v[c]) * g.acc_scale;
        }
        // ---- registers -> swizzled staging -> TMA store
        const uint32_t sbuf = my_out + (g.out_bufs == 2 ? (store_cnt & 1u) : 0u) * kStageBuf;
        if (lane == 0) {
          if (g.out_bufs == 2) { if (store_cnt >= 2) tma_store_wait_read<1>(); }
          else if (store_cnt >= 1) tma_store_wait_read<0>();
        }
        __syncwarp();
        if (EPI == EPI_D) {
#pragma unroll
          for (int ch = 0; ch < 8; ch++)
            sts128(sbuf + stage_off(lane, ch, 128, g.swz), __float_as_uint(o[4 * ch]), __float_as_uint(o[4 * ch + 1]),
                   __float_as_uint(o[4 * ch + 2]), __float_as_uint(o[4 * ch + 3]));
        } else {
#pragma unroll
          for (int ch = 0; ch < 4; ch++) {
            uint32_t hh[4], ll[4];
#pragma unroll
            for (int e = 0; e < 4; e++) split2(o[ch * 8 + 2 * e], o[ch * 8 + 2 * e + 1], hh[e], ll[e]);
            const uint32_t off = stage_off(lane, ch, 64, g.swz);
            sts128(sbuf + off, hh[0], hh[1], hh[2], hh[3]);
            sts128(sbuf + kStageBuf / 2 + off, ll[0], ll[1], ll[2], ll[3]);
          }
        }
        fence_async_smem();
        __syncwarp();
        if (lane == 0) {
          if (EPI == EPI_D) {
            tma_store_2d(&tmO_hi, sbuf, ocol + cb * 32, trow);
          } else {
            tma_store_2d(&tmO_hi, sbuf, ocol + cb * 32, trow);
            tma_store_2d(&tmO_lo, sbuf + kStageBuf / 2, ocol + cb * 32, trow);
          }
          tma_store_commit();
        }
        store_cnt++;
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(tempty_bar(as));
      if (EPI == EPI_FO) {                       // e_out is zeroed before the launch; the warps of a quadrant add their halves
        const int r = trow + lane;
        if (r < g.R) atomicAdd(g.e_out + (size_t)r * g.e_ld + z, e_acc + (half == 0 ? __ldg(g.b_out + z) : 0.f));
      }
    }
    if (lane == 0) tma_store_wait_all();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
  }
}

// ------------------------------------------------------------------------------------------------ host side
struct HalfPlanes {
  DevBufTc hi, lo;
  __half* h() const { return (__half*)hi.p; }
  __half* l() const { return (__half*)lo.p; }
  bool reserve(size_t elems) { return hi.reserve(elems * 2) && lo.reserve(elems * 2); }
};

struct SpeciesWeights {
  bool ready = false;
  int nh = 0;
  std::vector<HalfPlanes> Wf;     // [j] j=0..nh-1 : [M*dims[j+1]][dims[j]]   (scaled by sw[j])
  std::vector<HalfPlanes> WT;     // [j] j=1..nh-1 : [M*dims[j]][dims[j+1]]   (scaled by sw[j])
  HalfPlanes Wcat;                // [L][M*h1]                                 (scaled by sw[0])
  std::vector<float> sw;          // power-of-two scale of layer j's weight planes
  std::vector<DevBufTc> bias; // [j] : [M][dims[j+1]] fp32
  std::vector<DevBufTc> bias_s;   // [j] : log2(e)/alpha * bias[j]   (the fused kernel's CELU epilogue works on the scaled value)
  DevBufTc w_out, b_out;      // [M][dims[nh]], [M]
  DevBufTc w_out_s;           // kScaleG * w_out
};

struct Work {
  int chunk_rows = 0;
  std::vector<HalfPlanes> H, G;   // [l] l=1..nh : [chunk_rows][M*dims[l]]   (H[nh] is never materialised)
  DevBufTc e;                 // [chunk_rows][M] fp32 member energies
};

// generic 2-D tensor map: rows x cols elements of `esize` bytes, row pitch ld elements, box = box_cols x box_rows
inline int make_map(CUtensorMap* tm, const void* ptr, CUtensorMapDataType dt, int esize, uint64_t rows, uint64_t cols, uint64_t ld,
                    uint32_t box_cols, uint32_t box_rows, CUtensorMapSwizzle swz) {
  tc::EncodeTiledFn fn = tc::get_encode_fn();
  if (!fn) { set_error("cuTensorMapEncodeTiled is not available from the driver"); return 1; }
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {ld * (uint64_t)esize};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(tm, dt, 2, (void*)ptr, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, swz,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed with code " + std::to_string((int)r) + " (rows " + std::to_string(rows) + ", cols " +
              std::to_string(cols) + ", ld " + std::to_string(ld) + ", box " + std::to_string(box_cols) + "x" + std::to_string(box_rows) + ")");
    return 1;
  }
  return 0;
}
inline int make_operand_map(CUtensorMap* tm, const __half* p, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows, int bk) {
  return make_map(tm, p, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, rows, cols, ld, (uint32_t)bk, box_rows,
                  bk == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B);
}
inline int env_int(const char* name, int dflt) {
  const char* e = getenv(name);
  return (e && *e) ? atoi(e) : dflt;
}

inline int epi_swizzle() { return 1; }   // swizzled staging buffers (bank-conflict-free 16-byte stores), matching tensor maps

// scale w by the power of two that brings max|w| just below 2^14 and split into fp16 planes
inline float pick_scale(const std::vector<float>& w) {
  float mx = 0.f;
  for (float x : w) mx = std::max(mx, std::fabs(x));
  if (!(mx > 0.f) || !std::isfinite(mx)) return 1.f;
  int e = (int)std::floor(std::log2(16384.0 / (double)mx));
  e = std::max(-24, std::min(24, e));
  return std::ldexp(1.f, e);
}
inline size_t pad64(size_t cols) { return (cols + 63) / 64 * 64; }   // 128-byte row pitch: TMA box rows never straddle a line

// w: [rows][cols] row major -> fp16 hi / lo planes of scale * w with row pitch pad64(cols) (zero padded)
inline int upload_half_planes(const std::vector<float>& w, size_t cols, float scale, HalfPlanes& out) {
  const size_t rows = cols ? w.size() / cols : 0, ld = pad64(cols);
  std::vector<__half> h(rows * ld, __float2half_rn(0.f)), l(rows * ld, __float2half_rn(0.f));
  for (size_t r = 0; r < rows; r++)
    for (size_t c = 0; c < cols; c++) {
      const float x = w[r * cols + c] * scale;
      h[r * ld + c] = __float2half_rn(x);
      l[r * ld + c] = __float2half_rn(x - __half2float(h[r * ld + c]));
    }
  if (!out.reserve(std::max<size_t>(h.size(), 8))) { set_error("out of memory for weight planes"); return 1; }
  ANIB_CUDA_OK(cudaMemcpy(out.hi.p, h.data(), h.size() * 2, cudaMemcpyHostToDevice));
  ANIB_CUDA_OK(cudaMemcpy(out.lo.p, l.data(), l.size() * 2, cudaMemcpyHostToDevice));
  return 0;
}

inline bool supported_dims(const std::vector<int>& dims) {
  const int nh = (int)dims.size() - 2;
  if (nh < 1 || dims.back() != 1) return false;
  for (int l = 1; l <= nh; l++)
    if (dims[l] % 32 != 0 || dims[l] > 256) return false;   // 32-column epilogue blocks, one UMMA tile per member
  return dims[0] % 16 == 0;
}

// dims: in, h1..h_nh, 1;  hm weights W[m][s][l] row-major [out][in]
inline int prepare_species(const HostModel& hm, int M, int s, SpeciesWeights& w) {
  const std::vector<int>& dims = hm.dims[s];
  if (!supported_dims(dims)) { w.ready = false; return 0; }
  const int nh = (int)dims.size() - 2;
  w.nh = nh;
  w.Wf.resize(nh); w.WT.resize(nh); w.bias.resize(nh); w.bias_s.resize(nh); w.sw.assign(nh, 1.f);
  for (int j = 0; j < nh; j++) {
    const int ni = dims[j], no = dims[j + 1];
    std::vector<float> f((size_t)M * no * ni), b((size_t)M * no);
    for (int m = 0; m < M; m++) {
      memcpy(f.data() + (size_t)m * no * ni, hm.W[m][s][j].data(), (size_t)no * ni * 4);
      memcpy(b.data() + (size_t)m * no, hm.b[m][s][j].data(), (size_t)no * 4);
    }
    w.sw[j] = pick_scale(f);
    if (upload_half_planes(f, (size_t)ni, w.sw[j], w.Wf[j])) return 1;
    if (tc::upload_f32(b, w.bias[j])) return 1;
    for (float& x : b) x *= 14.426950408889634f;   // kFScaleH (mlp_f16_fused.cuh)
    if (tc::upload_f32(b, w.bias_s[j])) return 1;
    if (j >= 1) {
      std::vector<float> t((size_t)M * ni * no);
      for (int m = 0; m < M; m++)
        for (int o = 0; o < no; o++)
          for (int k = 0; k < ni; k++) t[((size_t)m * ni + k) * no + o] = hm.W[m][s][j][(size_t)o * ni + k];
      if (upload_half_planes(t, (size_t)no, w.sw[j], w.WT[j])) return 1;
    }
  }
  {
    const int L = dims[0], h1 = dims[1];
    std::vector<float> cat((size_t)L * M * h1);
    for (int m = 0; m < M; m++)
      for (int k = 0; k < h1; k++)
        for (int q = 0; q < L; q++) cat[(size_t)q * M * h1 + (size_t)m * h1 + k] = hm.W[m][s][0][(size_t)k * L + q];
    if (upload_half_planes(cat, (size_t)M * h1, w.sw[0], w.Wcat)) return 1;
  }
  {
    const int h = dims[nh];
    std::vector<float> wo((size_t)M * h), bo(M);
    for (int m = 0; m < M; m++) {
      memcpy(wo.data() + (size_t)m * h, hm.W[m][s][nh].data(), (size_t)h * 4);
      bo[m] = hm.b[m][s][nh][0];
    }
    if (tc::upload_f32(wo, w.w_out) || tc::upload_f32(bo, w.b_out)) return 1;
    for (float& x : wo) x *= kScaleG;
    if (tc::upload_f32(wo, w.w_out_s)) return 1;
  }
  w.ready = true;
  return 0;
}

struct Launcher {
  cudaStream_t st;
  int num_sms;
  int* launches;
  int max_smem;
};

struct Operand { const __half* hi; const __half* lo; uint64_t rows, cols, ld; };

// out_f32 != nullptr: EPI_D (fp32 rows);  otherwise planes.  aux only for EPI_B.
template <int EPI>
inline int launch_gemm(const Launcher& L, const Operand& A, const Operand& B, const Operand& O, float* out_f32, uint64_t o_rows,
                       uint64_t o_cols, uint64_t o_ld, const Operand& X, Args g) {
  CUtensorMap tA_hi, tA_lo, tB_hi, tB_lo, tO_hi, tO_lo, tX_hi, tX_lo;
  g.swz = epi_swizzle();
  g.bk = 32;          // measured best (profiles/README.md, round 1): 64-byte operand rows, twice the pipeline stages
  g.epi_warps = 8;
  if (make_operand_map(&tA_hi, A.hi, A.rows, A.cols, A.ld, BM, g.bk) || make_operand_map(&tA_lo, A.lo, A.rows, A.cols, A.ld, BM, g.bk) ||
      make_operand_map(&tB_hi, B.hi, B.rows, B.cols, B.ld, (uint32_t)g.BN, g.bk) ||
      make_operand_map(&tB_lo, B.lo, B.rows, B.cols, B.ld, (uint32_t)g.BN, g.bk))
    return 1;
  if (EPI == EPI_D) {
    if (make_map(&tO_hi, out_f32, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, o_rows, o_cols, o_ld, 32, 32,
                 g.swz ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE))
      return 1;
    tO_lo = tO_hi;
  } else {
    const CUtensorMapSwizzle sw = g.swz ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_NONE;
    if (make_map(&tO_hi, O.hi, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, O.rows, O.cols, O.ld, 32, 32, sw) ||
        make_map(&tO_lo, O.lo, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, O.rows, O.cols, O.ld, 32, 32, sw))
      return 1;
  }
  if (EPI == EPI_B) {
    const CUtensorMapSwizzle sw = g.swz ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_NONE;
    if (make_map(&tX_hi, X.hi, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, X.rows, X.cols, X.ld, 32, 32, sw) ||
        make_map(&tX_lo, X.lo, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, X.rows, X.cols, X.ld, 32, 32, sw))
      return 1;
  } else {
    tX_hi = tO_hi; tX_lo = tO_lo;
  }
  const int stage_bytes = 2 * (BM + g.BN) * g.bk * 2;
  const int want = (g.bk == 64) ? 2 : 4;       // k-blocks that should be in flight to cover the L2 -> smem latency
  int fixed = 0, stages = 0;
  for (g.out_bufs = 2; g.out_bufs >= 1; g.out_bufs--) {
    fixed = 1024 /*alignment*/ + g.epi_warps * kStageBuf * (g.out_bufs + (EPI == EPI_B ? 2 : 0)) + 512 /*barriers*/;
    stages = std::min((L.max_smem - fixed) / stage_bytes, kMaxStages);
    if (stages >= want || g.out_bufs == 1) break;
  }
  if (stages < 2) { set_error("not enough shared memory for the f16 tensor-core GEMM pipeline"); return 1; }
  g.stages = stages;
  const int smem = stages * stage_bytes + fixed;
  ANIB_CUDA_OK(cudaFuncSetAttribute(f16_gemm_kernel<EPI>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  const int tiles = g.num_row_tiles * g.num_n_tiles * g.batch;
  const int grid = std::max(1, std::min(tiles, L.num_sms));
  f16_gemm_kernel<EPI><<<grid, 128 + 32 * g.epi_warps, smem, L.st>>>(tA_hi, tA_lo, tB_hi, tB_lo, tO_hi, tO_lo, tX_hi, tX_lo, g);
  ANIB_CUDA_OK(cudaGetLastError());
  if (L.launches) (*L.launches)++;
  return 0;
}

// member energies [R][M] fp32 -> eatom[r] = mean_m e + sae  (fp64)
__global__ void energy_reduce_kernel(const float* __restrict__ e, int R, int M, double sae, double* __restrict__ eatom) {
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= R) return;
  double s = 0.0;
  for (int m = 0; m < M; m++) s += (double)e[(size_t)r * M + m];
  eatom[r] = s / (double)M + sae;
}

inline int64_t padded_rows(int nlocal) { return (int64_t)nlocal + BM; }

// Forward + backward of one species bucket.  aev planes: [aev_rows_total][ld_aev] halves (values x kScaleAev) in bucket
// order; this bucket = rows [row0, row0+R).  daev: fp32 rows [.., ld_daev] in bucket order.
inline int mlp_species(cudaStream_t st, SpeciesWeights& w, Work& work, int M, const std::vector<int>& dims, const __half* aev_hi,
                       const __half* aev_lo, int64_t ld_aev, int64_t aev_rows_total, int row0, int R, float* daev, int64_t ld_daev,
                       double* eatom, double sae, int* launches, int cta_cap = 0) {
  if (!w.ready) { set_error("this network's layer widths are not supported by the f16 tensor-core path (need multiples of 32, <= 256)"); return 1; }
  static int num_sms = 0, max_smem = 0;
  if (!num_sms) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, dev);
    cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
  }
  Launcher L{st, cta_cap > 0 ? std::min(num_sms, cta_cap) : num_sms, launches, max_smem};   // cta_cap: option "gemm_ctas" (leave SMs to another stream)
  const int nh = w.nh;
  int chunk = work.chunk_rows;
  if (chunk == 0) {
    const char* e = getenv("ANI_B200_TC_CHUNK");
    chunk = (e && atoi(e) > 0) ? (atoi(e) + BM - 1) / BM * BM : 262144;   // measured: fewer, larger launches win (profiles/README.md)
    work.chunk_rows = chunk;
  }
  const int alloc_rows = std::min(chunk, (R + BM - 1) / BM * BM);
  if ((int)work.H.size() < nh + 1) { work.H.resize(nh + 1); work.G.resize(nh + 1); }
  for (int l = 1; l <= nh; l++) {
    const size_t elems = (size_t)alloc_rows * M * dims[l];
    if ((l < nh && !work.H[l].reserve(elems)) || !work.G[l].reserve(elems)) { set_error("out of memory for MLP work buffers"); return 1; }
  }
  if (!work.e.reserve((size_t)alloc_rows * M * 4)) { set_error("out of memory for MLP work buffers"); return 1; }
  const int Lin = dims[0];
  const Operand none{nullptr, nullptr, 0, 0, 0};
  for (int r0 = 0; r0 < R; r0 += chunk) {
    const int Rc = std::min(chunk, R - r0);
    const int rtiles = (Rc + BM - 1) / BM;
    // ---- forward hidden layers (the last one fused with the output layer)
    for (int j = 0; j < nh; j++) {
      const int ni = dims[j], no = dims[j + 1];
      const bool last = (j == nh - 1);
      const float s_in = (j == 0) ? kScaleAev : kScaleH;
      Args g{};
      g.R = Rc; g.N = no; g.BN = no; g.num_n_tiles = 1; g.num_row_tiles = rtiles; g.K = ni; g.batch = M;
      g.b_row_off_per_z = no;
      g.acc_scale = 1.0f / (s_in * w.sw[j]);
      g.bias = w.bias[j].f(); g.bias_stride_z = no;
      g.out_col_off_per_z = no;
      Operand A = (j == 0) ? Operand{aev_hi, aev_lo, (uint64_t)aev_rows_total, (uint64_t)Lin, (uint64_t)ld_aev}
                           : Operand{work.H[j].h(), work.H[j].l(), (uint64_t)Rc, (uint64_t)M * ni, (uint64_t)M * ni};
      if (j == 0) { g.a_row0 = row0 + r0; g.a_k_off_per_z = 0; } else { g.a_row0 = 0; g.a_k_off_per_z = ni; }
      Operand B{w.Wf[j].h(), w.Wf[j].l(), (uint64_t)M * no, (uint64_t)ni, (uint64_t)pad64(ni)};
      int rc;
      if (!last) {
        g.out_scale = kScaleH;
        Operand O{work.H[j + 1].h(), work.H[j + 1].l(), (uint64_t)Rc, (uint64_t)M * no, (uint64_t)M * no};
        rc = launch_gemm<EPI_F>(L, A, B, O, nullptr, 0, 0, 0, none, g);
      } else {
        g.out_scale = kScaleG;
        ANIB_CUDA_OK(cudaMemsetAsync(work.e.p, 0, (size_t)Rc * M * 4, st));
        g.w_out = w.w_out.f(); g.b_out = w.b_out.f(); g.e_out = work.e.f(); g.e_ld = M;
        Operand O{work.G[nh].h(), work.G[nh].l(), (uint64_t)Rc, (uint64_t)M * no, (uint64_t)M * no};
        rc = launch_gemm<EPI_FO>(L, A, B, O, nullptr, 0, 0, 0, none, g);
      }
      if (rc) return rc;
    }
    energy_reduce_kernel<<<(Rc + 255) / 256, 256, 0, st>>>(work.e.f(), Rc, M, sae, eatom + row0 + r0);
    if (launches) (*launches)++;
    // ---- backward through hidden layers nh-1 .. 1:  G_j = (G_{j+1} . W_{j+1}) * CELU'(z_j)
    for (int j = nh - 1; j >= 1; j--) {
      const int ni = dims[j], no = dims[j + 1];   // layer j: ni -> no ; gradient: no -> ni
      Args g{};
      g.BN = (ni <= 192) ? ni : 128;              // two staging areas (out + aux) leave room for a 192-wide B tile at most
      g.R = Rc; g.N = ni; g.num_n_tiles = (ni + g.BN - 1) / g.BN; g.num_row_tiles = rtiles; g.K = no; g.batch = M;
      g.a_row0 = 0; g.a_k_off_per_z = no; g.b_row_off_per_z = ni;
      g.acc_scale = 1.0f / w.sw[j];               // (sG * sw) accumulator -> sG-scaled gradient
      g.aux_inv_scale = 1.0f / kScaleH;
      g.out_col_off_per_z = ni; g.aux_col_off_per_z = ni;
      Operand A{work.G[j + 1].h(), work.G[j + 1].l(), (uint64_t)Rc, (uint64_t)M * no, (uint64_t)M * no};
      Operand B{w.WT[j].h(), w.WT[j].l(), (uint64_t)M * ni, (uint64_t)no, (uint64_t)pad64(no)};
      Operand O{work.G[j].h(), work.G[j].l(), (uint64_t)Rc, (uint64_t)M * ni, (uint64_t)M * ni};
      Operand X{work.H[j].h(), work.H[j].l(), (uint64_t)Rc, (uint64_t)M * ni, (uint64_t)M * ni};
      if (launch_gemm<EPI_B>(L, A, B, O, nullptr, 0, 0, 0, X, g)) return 1;
    }
    // ---- dE/dAEV = (1/M) * G1[Rc][M*h1] . Wcat[L][M*h1]^T   (ensemble mean folded into K)
    {
      const int h1 = dims[1];
      Args g{};
      g.BN = (Lin <= 256 && Lin % 32 == 0) ? Lin : 256;
      g.R = Rc; g.N = Lin; g.num_n_tiles = (Lin + g.BN - 1) / g.BN; g.num_row_tiles = rtiles; g.K = M * h1; g.batch = 1;
      g.acc_scale = 1.0f / (kScaleG * w.sw[0] * (float)M);
      Operand A{work.G[1].h(), work.G[1].l(), (uint64_t)Rc, (uint64_t)M * h1, (uint64_t)M * h1};
      Operand B{w.Wcat.h(), w.Wcat.l(), (uint64_t)Lin, (uint64_t)M * h1, (uint64_t)pad64((size_t)M * h1)};
      if (launch_gemm<EPI_D>(L, A, B, none, daev + (size_t)(row0 + r0) * ld_daev, (uint64_t)Rc, (uint64_t)Lin, (uint64_t)ld_daev, none, g))
        return 1;
    }
  }
  return 0;
}

}  // namespace f16
}  // namespace anib
