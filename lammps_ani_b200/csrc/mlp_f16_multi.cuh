// mlp_f16_multi.cuh -- one persistent launch that interleaves the tiles of up to three INDEPENDENT f16x3 GEMMs.
//
// Why: the ensemble MLP of a chunk of atoms is six GEMMs: two are tensor bound (layer 1, K = AEV length;
// dE/dAEV, K = M*h1) and four are HBM bound (the narrow middle layers stream their activations / gradients at 5 TB/s while
// the tensor pipe idles).  Run one after the other, their times add.  Here the chunks are software-pipelined at launch
// level: launch A_t holds layer-1 tiles of chunk t, layer-2 tiles of chunk t-1 and layer-3(+output) tiles of chunk t-2;
// launch B_t holds the backward layers of chunks t-2, t-3 and dE/dAEV of chunk t-4.  No two GEMMs of one launch depend on
// each other, so every CTA simply alternates between them: while its tensor pipe works through a long-K tile, the TMA
// unit and the epilogue warps move the narrow layers' data.  Same warp roles, pipeline, epilogues and numerics as
// f16_gemm_kernel (mlp_f16.cuh); the per-problem parameters live in the __grid_constant__ block.
#pragma once
#include "mlp_f16.cuh"

namespace anib {
namespace f16 {

constexpr int kMaxProb = 3;

struct alignas(64) Prob {
  CUtensorMap tmA_hi, tmA_lo, tmB_hi, tmB_lo, tmO_hi, tmO_lo, tmX_hi, tmX_lo;
  Args g;
  int epi;       // EPI_F / EPI_FO / EPI_B / EPI_D
  int ntiles;    // num_row_tiles * num_n_tiles * batch
};

struct MultiParams {
  Prob p[kMaxProb];
  int nprob;
  int stages, stage_stride;      // shared operand ring: `stages` slots of `stage_stride` bytes (sized for the widest problem)
  int epi_warps, out_bufs, aux_bufs;   // aux_bufs == 0: no problem needs the second epilogue input
  int bk, swz;
};

// The merged tile sequence of one CTA.  CTA b owns tiles b, b+G, b+2G, ... of every problem; it walks them in proportion
// (the problem that is furthest behind goes next), which every warp role recomputes identically from the tile counts.
struct Sched {
  int n0, n1, n2, d0, d1, d2;
  __device__ __forceinline__ static int owned(int ntiles) {
    return ntiles > (int)blockIdx.x ? (ntiles - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;
  }
  __device__ __forceinline__ void init(const MultiParams& mp) {
    n0 = owned(mp.p[0].ntiles);
    n1 = mp.nprob > 1 ? owned(mp.p[1].ntiles) : 0;
    n2 = mp.nprob > 2 ? owned(mp.p[2].ntiles) : 0;
    d0 = d1 = d2 = 0;
  }
  // progress key (2d+1)/(2n), compared by cross multiplication; ties go to the lower problem index
  __device__ __forceinline__ bool next(int& p, int& t) {
    long long bn = 1, bd = 0;   // best = +inf
    p = -1;
    if (d0 < n0) { bn = 2LL * d0 + 1; bd = 2LL * n0; p = 0; }
    if (d1 < n1) { const long long a = 2LL * d1 + 1, b = 2LL * n1; if (p < 0 || a * bd < bn * b) { bn = a; bd = b; p = 1; } }
    if (d2 < n2) { const long long a = 2LL * d2 + 1, b = 2LL * n2; if (p < 0 || a * bd < bn * b) { bn = a; bd = b; p = 2; } }
    if (p < 0) return false;
    const int d = p == 0 ? d0++ : p == 1 ? d1++ : d2++;
    t = (int)blockIdx.x + d * (int)gridDim.x;
    return true;
  }
};

struct EpiState {
  uint32_t aux_issued, aux_used, store_cnt;
};

// Epilogue of one tile for one epilogue warp (identical math to f16_gemm_kernel's epilogue).
template <int EPI>
__device__ __forceinline__ void epi_tile(const Prob& P, const MultiParams& mp, int t, int it, uint32_t tmem_base, uint32_t my_out,
                                         uint32_t my_aux, uint32_t aux_bar0, uint32_t tfull, uint32_t tempty, int q, int half,
                                         int nsplit, int lane, EpiState& S) {
  const Args& g = P.g;
  const int z = t % g.batch;
  const int rest = t / g.batch;
  const int nt = rest % g.num_n_tiles, rt = rest / g.num_n_tiles;
  const int as = it & 1;
  const uint32_t aphase = (uint32_t)(it >> 1) & 1u;
  const int trow = rt * BM + q * 32;
  const int ocol = z * g.out_col_off_per_z + nt * g.BN;
  const int xcol = z * g.aux_col_off_per_z + nt * g.BN;
  const int nvalid = min(g.BN / 32, (g.N - nt * g.BN + 31) / 32);
  const uint32_t nab = (uint32_t)mp.aux_bufs;
  auto issue_aux = [&](int cb) {
    const uint32_t slot = S.aux_issued % nab;
    const uint32_t buf = my_aux + slot * kStageBuf;
    const uint32_t bar = aux_bar0 + 8u * slot;
    if (lane == 0) {
      mbar_expect_tx(bar, kStageBuf);
      tma_load_2d(&P.tmX_hi, bar, buf, xcol + cb * 32, trow);
      tma_load_2d(&P.tmX_lo, bar, buf + kStageBuf / 2, xcol + cb * 32, trow);
    }
    S.aux_issued++;
  };
  int aux_next = half;                       // next column block whose second input has not been requested yet
  if (EPI == EPI_B && aux_next < nvalid) { issue_aux(aux_next); aux_next += nsplit; }
  mbar_wait(tfull, aphase);
  tc_fence_after();
  const uint32_t taddr0 = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)as * 256u;
  float e_acc = 0.f;
  for (int cb = half; cb < nvalid; cb += nsplit) {
    if (EPI == EPI_B && aux_next < nvalid && S.aux_issued - S.aux_used < nab) { issue_aux(aux_next); aux_next += nsplit; }
    uint32_t v[32];
    tmem_ld32(taddr0 + (uint32_t)(cb * 32), v);
    const int n0 = nt * g.BN + cb * 32;
    float o[32];
    if (EPI == EPI_F || EPI == EPI_FO) {
      const float* bias = g.bias + (size_t)z * g.bias_stride_z + n0;
      const float* wo = (EPI == EPI_FO) ? g.w_out + (size_t)z * g.N + n0 : nullptr;
#pragma unroll
      for (int c = 0; c < 32; c++) {
        const float zz = fmaf(__uint_as_float(v[c]), g.acc_scale, __ldg(bias + c));
        const float ex = exp2f(zz * 14.426950408889634f);
        const float h = zz > 0.f ? zz : 0.1f * (ex - 1.f);
        if (EPI == EPI_F) {
          o[c] = h * g.out_scale;
        } else {
          const float w = __ldg(wo + c);
          e_acc = fmaf(h, w, e_acc);
          o[c] = w * (zz > 0.f ? 1.f : ex) * g.out_scale;
        }
      }
    } else if (EPI == EPI_B) {
      const uint32_t slot = S.aux_used % nab;
      const uint32_t buf = my_aux + slot * kStageBuf;
      mbar_wait(aux_bar0 + 8u * slot, (S.aux_used / nab) & 1u);
      S.aux_used++;
#pragma unroll
      for (int ch = 0; ch < 4; ch++) {
        uint32_t hh[4], ll[4];
        const uint32_t off = stage_off(lane, ch, 64, mp.swz);
        lds128(buf + off, hh[0], hh[1], hh[2], hh[3]);
        lds128(buf + kStageBuf / 2 + off, ll[0], ll[1], ll[2], ll[3]);
#pragma unroll
        for (int e = 0; e < 8; e++) {
          const float hs = unpack_sum(hh[e >> 1], ll[e >> 1], e & 1) * g.aux_inv_scale;
          const float d = hs > 0.f ? 1.f : fmaf(10.f, hs, 1.f);
          o[ch * 8 + e] = __uint_as_float(v[ch * 8 + e]) * g.acc_scale * d;
        }
      }
    } else {
#pragma unroll
      for (int c = 0; c < 32; c++) o[c] = __uint_as_float(v[c]) * g.acc_scale;
    }
    const uint32_t sbuf = my_out + (mp.out_bufs == 2 ? (S.store_cnt & 1u) : 0u) * kStageBuf;
    if (lane == 0) {
      if (mp.out_bufs == 2) { if (S.store_cnt >= 2) tma_store_wait_read<1>(); }
      else if (S.store_cnt >= 1) tma_store_wait_read<0>();
    }
    __syncwarp();      // also orders this block's reads of the aux buffer before the next TMA load into it
    if (EPI == EPI_B && aux_next < nvalid && S.aux_issued - S.aux_used < nab) { issue_aux(aux_next); aux_next += nsplit; }
    if (EPI == EPI_D) {
#pragma unroll
      for (int ch = 0; ch < 8; ch++)
        sts128(sbuf + stage_off(lane, ch, 128, mp.swz), __float_as_uint(o[4 * ch]), __float_as_uint(o[4 * ch + 1]),
               __float_as_uint(o[4 * ch + 2]), __float_as_uint(o[4 * ch + 3]));
    } else {
#pragma unroll
      for (int ch = 0; ch < 4; ch++) {
        uint32_t hh[4], ll[4];
#pragma unroll
        for (int e = 0; e < 4; e++) split2(o[ch * 8 + 2 * e], o[ch * 8 + 2 * e + 1], hh[e], ll[e]);
        const uint32_t off = stage_off(lane, ch, 64, mp.swz);
        sts128(sbuf + off, hh[0], hh[1], hh[2], hh[3]);
        sts128(sbuf + kStageBuf / 2 + off, ll[0], ll[1], ll[2], ll[3]);
      }
    }
    fence_async_smem();
    __syncwarp();
    if (lane == 0) {
      tma_store_2d(&P.tmO_hi, sbuf, ocol + cb * 32, trow);
      if (EPI != EPI_D) tma_store_2d(&P.tmO_lo, sbuf + kStageBuf / 2, ocol + cb * 32, trow);
      tma_store_commit();
    }
    S.store_cnt++;
  }
  tc_fence_before();
  __syncwarp();
  if (lane == 0) mbar_arrive(tempty);
  if (EPI == EPI_FO) {
    const int r = trow + lane;
    if (r < g.R) atomicAdd(g.e_out + (size_t)r * g.e_ld + z, e_acc + (half == 0 ? __ldg(g.b_out + z) : 0.f));
  }
}

__global__ void __launch_bounds__(kMaxThreads, 1) f16_multi_kernel(const __grid_constant__ MultiParams mp) {
  extern __shared__ unsigned char smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  const uint32_t row_bytes = (uint32_t)mp.bk * 2u;
  const uint32_t a_bytes = (uint32_t)BM * row_bytes;
  const uint32_t base_out = base + mp.stages * mp.stage_stride;
  const uint32_t base_aux = base_out + mp.epi_warps * mp.out_bufs * kStageBuf;
  const uint32_t bar_base = base_aux + mp.epi_warps * mp.aux_bufs * kStageBuf;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (kMaxStages + s); };
  auto tfull_bar = [&](int s) { return bar_base + 8u * (2 * kMaxStages + s); };
  auto tempty_bar = [&](int s) { return bar_base + 8u * (2 * kMaxStages + 2 + s); };
  auto aux_bar = [&](int w, int b) { return bar_base + 8u * (2 * kMaxStages + 4 + 2 * w + b); };
  const uint32_t tmem_slot = bar_base + 8u * (2 * kMaxStages + 4 + 2 * kMaxEpiWarps);
  volatile uint32_t* tmem_slot_ptr = (volatile uint32_t*)(smem_raw + (tmem_slot - raw));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp == 1 && lane == 0) {
    for (int s = 0; s < mp.stages; s++) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
    for (int s = 0; s < 2; s++) { mbar_init(tfull_bar(s), 1); mbar_init(tempty_bar(s), mp.epi_warps); }
    for (int w = 0; w < mp.epi_warps; w++) { mbar_init(aux_bar(w, 0), 1); mbar_init(aux_bar(w, 1), 1); }
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
      Sched sc; sc.init(mp);
      int stage = 0; uint32_t phase = 0;
      int p, t;
      while (sc.next(p, t)) {
        const Prob& P = mp.p[p];
        const Args& g = P.g;
        const uint32_t b_bytes = (uint32_t)g.BN * row_bytes;
        const uint32_t tx = 2u * (a_bytes + b_bytes);
        const int KB = (g.K + mp.bk - 1) / mp.bk;
        const int z = t % g.batch;
        const int rest = t / g.batch;
        const int nt = rest % g.num_n_tiles, rt = rest / g.num_n_tiles;
        const int a_row = g.a_row0 + rt * BM;
        const int a_col = z * g.a_k_off_per_z;
        const int b_row = z * g.b_row_off_per_z + nt * g.BN;
        for (int kb = 0; kb < KB; kb++) {
          mbar_wait(empty_bar(stage), phase ^ 1u);
          const uint32_t sA = base + stage * mp.stage_stride;
          const uint32_t sB = sA + 2 * a_bytes;
          mbar_expect_tx(full_bar(stage), tx);
          tma_load_2d(&P.tmA_hi, full_bar(stage), sA, a_col + kb * mp.bk, a_row);
          tma_load_2d(&P.tmB_hi, full_bar(stage), sB, kb * mp.bk, b_row);
          tma_load_2d(&P.tmA_lo, full_bar(stage), sA + a_bytes, a_col + kb * mp.bk, a_row);
          tma_load_2d(&P.tmB_lo, full_bar(stage), sB + b_bytes, kb * mp.bk, b_row);
          if (++stage == mp.stages) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // ================================================================= MMA issuer
    if (lane == 0) {
      Sched sc; sc.init(mp);
      int stage = 0; uint32_t phase = 0;
      int it = 0, p, t;
      while (sc.next(p, t)) {
        const Args& g = mp.p[p].g;
        const uint32_t b_bytes = (uint32_t)g.BN * row_bytes;
        const int KB = (g.K + mp.bk - 1) / mp.bk;
        const uint32_t idesc = umma_idesc_f16(BM, g.BN);
        const int as = it & 1;
        const uint32_t aphase = (uint32_t)(it >> 1) & 1u;
        mbar_wait(tempty_bar(as), aphase ^ 1u);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + (uint32_t)as * 256u;
        for (int kb = 0; kb < KB; kb++) {
          mbar_wait(full_bar(stage), phase);
          tc_fence_after();
          const uint32_t sA = base + stage * mp.stage_stride;
          const uint32_t sB = sA + 2 * a_bytes;
          const uint64_t dA_hi = umma_desc_k(sA, mp.bk), dB_hi = umma_desc_k(sB, mp.bk);
          const uint64_t dA_lo = umma_desc_k(sA + a_bytes, mp.bk), dB_lo = umma_desc_k(sB + b_bytes, mp.bk);
          const int nks = min(mp.bk / 16, (g.K - kb * mp.bk) / 16);
          for (int k = 0; k < nks; k++) {
            const uint64_t adv = (uint64_t)((k * 32) >> 4);
            umma_f16(tmem_d, dA_hi + adv, dB_hi + adv, idesc, (kb | k) != 0 ? 1u : 0u);
            umma_f16(tmem_d, dA_lo + adv, dB_hi + adv, idesc, 1u);
            umma_f16(tmem_d, dA_hi + adv, dB_lo + adv, idesc, 1u);
          }
          umma_commit(empty_bar(stage));
          if (++stage == mp.stages) { stage = 0; phase ^= 1u; }
        }
        umma_commit(tfull_bar(as));
        it++;
      }
    }
  } else if (warp >= 4 && warp - 4 < mp.epi_warps) {
    // ================================================================= epilogue
    const int q = warp & 3;
    const int ew = warp - 4, half = ew >> 2, nsplit = mp.epi_warps >> 2;
    const uint32_t my_out = base_out + (uint32_t)ew * mp.out_bufs * kStageBuf;
    const uint32_t my_aux = base_aux + (uint32_t)ew * mp.aux_bufs * kStageBuf;
    EpiState S{0u, 0u, 0u};
    Sched sc; sc.init(mp);
    int it = 0, p, t;
    while (sc.next(p, t)) {
      const Prob& P = mp.p[p];
      const int as = it & 1;
      switch (P.epi) {
        case EPI_F:
          epi_tile<EPI_F>(P, mp, t, it, tmem_base, my_out, my_aux, aux_bar(ew, 0), tfull_bar(as), tempty_bar(as), q, half, nsplit, lane, S);
          break;
        case EPI_FO:
          epi_tile<EPI_FO>(P, mp, t, it, tmem_base, my_out, my_aux, aux_bar(ew, 0), tfull_bar(as), tempty_bar(as), q, half, nsplit, lane, S);
          break;
        case EPI_B:
          epi_tile<EPI_B>(P, mp, t, it, tmem_base, my_out, my_aux, aux_bar(ew, 0), tfull_bar(as), tempty_bar(as), q, half, nsplit, lane, S);
          break;
        default:
          epi_tile<EPI_D>(P, mp, t, it, tmem_base, my_out, my_aux, aux_bar(ew, 0), tfull_bar(as), tempty_bar(as), q, half, nsplit, lane, S);
          break;
      }
      it++;
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
struct Unit { int s, row0, R; };   // one chunk of one species bucket: rows [row0, row0 + R) of the bucket-ordered AEV

struct PipeCtx {
  cudaStream_t st;
  int M;
  const __half* aev_hi; const __half* aev_lo; int64_t ld_aev, aev_rows_total;
  float* daev; int64_t ld_daev;
  double* eatom;
  int* launches;
  int num_sms, max_smem, bk, swz;
};

inline int encode_prob_maps(Prob& P, const Operand& A, const Operand& B, const Operand& O, float* out_f32, uint64_t o_rows,
                            uint64_t o_cols, uint64_t o_ld, const Operand& X, int bk, int swz) {
  const int BN = P.g.BN;
  if (make_operand_map(&P.tmA_hi, A.hi, A.rows, A.cols, A.ld, BM, bk) || make_operand_map(&P.tmA_lo, A.lo, A.rows, A.cols, A.ld, BM, bk) ||
      make_operand_map(&P.tmB_hi, B.hi, B.rows, B.cols, B.ld, (uint32_t)BN, bk) ||
      make_operand_map(&P.tmB_lo, B.lo, B.rows, B.cols, B.ld, (uint32_t)BN, bk))
    return 1;
  const CUtensorMapSwizzle sw64 = swz ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_NONE;
  if (P.epi == EPI_D) {
    if (make_map(&P.tmO_hi, out_f32, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, o_rows, o_cols, o_ld, 32, 32,
                 swz ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE))
      return 1;
    P.tmO_lo = P.tmO_hi;
  } else if (make_map(&P.tmO_hi, O.hi, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, O.rows, O.cols, O.ld, 32, 32, sw64) ||
             make_map(&P.tmO_lo, O.lo, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, O.rows, O.cols, O.ld, 32, 32, sw64)) {
    return 1;
  }
  if (P.epi == EPI_B) {
    if (make_map(&P.tmX_hi, X.hi, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, X.rows, X.cols, X.ld, 32, 32, sw64) ||
        make_map(&P.tmX_lo, X.lo, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, X.rows, X.cols, X.ld, 32, 32, sw64))
      return 1;
  } else {
    P.tmX_hi = P.tmO_hi; P.tmX_lo = P.tmO_lo;
  }
  P.ntiles = P.g.num_row_tiles * P.g.num_n_tiles * P.g.batch;
  return 0;
}

// stage 0..nh-1: forward layers (the last one fused with the output layer); nh..2nh-2: backward layers nh-1..1; 2nh-1: dE/dAEV
inline int make_stage_prob(Prob& P, int stage, const Unit& u, SpeciesWeights& w, Work& wk, const std::vector<int>& dims,
                           const PipeCtx& c) {
  const int nh = w.nh, M = c.M, Rc = u.R;
  const int rtiles = (Rc + BM - 1) / BM;
  const Operand none{nullptr, nullptr, 0, 0, 0};
  Args g{};
  g.R = Rc; g.num_row_tiles = rtiles;
  P = Prob{};
  if (stage < nh) {
    const int j = stage, ni = dims[j], no = dims[j + 1];
    const bool last = (j == nh - 1);
    const float s_in = (j == 0) ? kScaleAev : kScaleH;
    g.N = no; g.BN = no; g.num_n_tiles = 1; g.K = ni; g.batch = M;
    g.b_row_off_per_z = no;
    g.acc_scale = 1.0f / (s_in * w.sw[j]);
    g.bias = w.bias[j].f(); g.bias_stride_z = no;
    g.out_col_off_per_z = no;
    Operand A = (j == 0) ? Operand{c.aev_hi, c.aev_lo, (uint64_t)c.aev_rows_total, (uint64_t)dims[0], (uint64_t)c.ld_aev}
                         : Operand{wk.H[j].h(), wk.H[j].l(), (uint64_t)Rc, (uint64_t)M * ni, (uint64_t)M * ni};
    if (j == 0) { g.a_row0 = u.row0; g.a_k_off_per_z = 0; } else { g.a_row0 = 0; g.a_k_off_per_z = ni; }
    Operand B{w.Wf[j].h(), w.Wf[j].l(), (uint64_t)M * no, (uint64_t)ni, (uint64_t)pad64(ni)};
    if (!last) {
      g.out_scale = kScaleH;
      P.g = g; P.epi = EPI_F;
      Operand O{wk.H[j + 1].h(), wk.H[j + 1].l(), (uint64_t)Rc, (uint64_t)M * no, (uint64_t)M * no};
      return encode_prob_maps(P, A, B, O, nullptr, 0, 0, 0, none, c.bk, c.swz);
    }
    g.out_scale = kScaleG;
    g.w_out = w.w_out.f(); g.b_out = w.b_out.f(); g.e_out = wk.e.f(); g.e_ld = M;
    P.g = g; P.epi = EPI_FO;
    Operand O{wk.G[nh].h(), wk.G[nh].l(), (uint64_t)Rc, (uint64_t)M * no, (uint64_t)M * no};
    return encode_prob_maps(P, A, B, O, nullptr, 0, 0, 0, none, c.bk, c.swz);
  }
  if (stage < 2 * nh - 1) {
    const int j = nh - 1 - (stage - nh);        // layer j: ni -> no ; gradient: no -> ni
    const int ni = dims[j], no = dims[j + 1];
    g.BN = (ni <= 192) ? ni : 128;
    g.N = ni; g.num_n_tiles = (ni + g.BN - 1) / g.BN; g.K = no; g.batch = M;
    g.a_row0 = 0; g.a_k_off_per_z = no; g.b_row_off_per_z = ni;
    g.acc_scale = 1.0f / w.sw[j];
    g.aux_inv_scale = 1.0f / kScaleH;
    g.out_col_off_per_z = ni; g.aux_col_off_per_z = ni;
    P.g = g; P.epi = EPI_B;
    Operand A{wk.G[j + 1].h(), wk.G[j + 1].l(), (uint64_t)Rc, (uint64_t)M * no, (uint64_t)M * no};
    Operand B{w.WT[j].h(), w.WT[j].l(), (uint64_t)M * ni, (uint64_t)no, (uint64_t)pad64(no)};
    Operand O{wk.G[j].h(), wk.G[j].l(), (uint64_t)Rc, (uint64_t)M * ni, (uint64_t)M * ni};
    Operand X{wk.H[j].h(), wk.H[j].l(), (uint64_t)Rc, (uint64_t)M * ni, (uint64_t)M * ni};
    return encode_prob_maps(P, A, B, O, nullptr, 0, 0, 0, X, c.bk, c.swz);
  }
  {
    const int Lin = dims[0], h1 = dims[1];
    g.BN = (Lin <= 256 && Lin % 32 == 0) ? Lin : 256;
    g.N = Lin; g.num_n_tiles = (Lin + g.BN - 1) / g.BN; g.K = M * h1; g.batch = 1;
    g.acc_scale = 1.0f / (kScaleG * w.sw[0] * (float)M);
    P.g = g; P.epi = EPI_D;
    Operand A{wk.G[1].h(), wk.G[1].l(), (uint64_t)Rc, (uint64_t)M * h1, (uint64_t)M * h1};
    Operand B{w.Wcat.h(), w.Wcat.l(), (uint64_t)Lin, (uint64_t)M * h1, (uint64_t)pad64((size_t)M * h1)};
    return encode_prob_maps(P, A, B, none, c.daev + (size_t)u.row0 * c.ld_daev, (uint64_t)Rc, (uint64_t)Lin, (uint64_t)c.ld_daev,
                            none, c.bk, c.swz);
  }
}

inline int launch_multi(const PipeCtx& c, MultiParams& mp) {
  if (mp.nprob < 1) return 0;
  int maxbn = 0, tiles = 0;
  bool aux = false;
  for (int p = 0; p < mp.nprob; p++) {
    maxbn = std::max(maxbn, mp.p[p].g.BN);
    tiles += mp.p[p].ntiles;
    aux = aux || mp.p[p].epi == EPI_B;
  }
  mp.bk = c.bk; mp.swz = c.swz; mp.epi_warps = 8;
  mp.stage_stride = 2 * (BM + maxbn) * mp.bk * 2;
  const int want = (mp.bk == 64) ? 2 : 4;
  int best_stages = 0, fixed = 0;
  // prefer more operand stages over deeper epilogue buffering
  const int cand[4][2] = {{2, 2}, {1, 2}, {1, 1}, {1, 1}};
  for (int k = 0; k < 3; k++) {
    const int ob = cand[k][0], ab = aux ? cand[k][1] : 0;
    const int fx = 1024 + mp.epi_warps * kStageBuf * (ob + ab) + 512;
    const int stg = std::min((c.max_smem - fx) / mp.stage_stride, kMaxStages);
    if (stg > best_stages) { best_stages = stg; mp.out_bufs = ob; mp.aux_bufs = ab; fixed = fx; }
    if (best_stages >= want) break;
  }
  if (best_stages < 2) { set_error("not enough shared memory for the f16 tensor-core GEMM pipeline"); return 1; }
  mp.stages = best_stages;
  const int smem = mp.stages * mp.stage_stride + fixed;
  static int attr_set = 0;
  if (attr_set < smem) {
    ANIB_CUDA_OK(cudaFuncSetAttribute(f16_multi_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, c.max_smem));
    attr_set = c.max_smem;
  }
  const int grid = std::max(1, std::min(tiles, c.num_sms));
  f16_multi_kernel<<<grid, 128 + 32 * mp.epi_warps, smem, c.st>>>(mp);
  ANIB_CUDA_OK(cudaGetLastError());
  if (c.launches) (*c.launches)++;
  return 0;
}

// Software pipeline over the units (see the file header).  W / dims / sae are indexed by species; `ring` holds 2*nh - 1
// work sets (one per unit in flight), each sized for `chunk_rows` rows of the widest network.
inline int mlp_pipeline(const PipeCtx& c, const std::vector<Unit>& units, std::vector<SpeciesWeights*>& W,
                        std::vector<const std::vector<int>*>& dims, const std::vector<double>& sae, std::vector<Work>& ring, int nh) {
  const int U = (int)units.size();
  const int depth = 2 * nh - 1;
  if ((int)ring.size() < depth) { set_error("internal error: work ring too small"); return 1; }
  for (int t = 0; t < U + depth - 1; t++) {
    MultiParams a{};   // forward stages: F_i of unit t - i
    for (int i = 0; i < nh; i++) {
      const int u = t - i;
      if (u < 0 || u >= U) continue;
      const Unit& un = units[u];
      Work& wk = ring[u % depth];
      if (i == nh - 1) ANIB_CUDA_OK(cudaMemsetAsync(wk.e.p, 0, (size_t)un.R * c.M * 4, c.st));
      if (make_stage_prob(a.p[a.nprob], i, un, *W[un.s], wk, *dims[un.s], c)) return 1;
      a.nprob++;
    }
    if (launch_multi(c, a)) return 1;
    {
      const int u = t - (nh - 1);
      if (u >= 0 && u < U) {
        const Unit& un = units[u];
        energy_reduce_kernel<<<(un.R + 255) / 256, 256, 0, c.st>>>(ring[u % depth].e.f(), un.R, c.M, sae[un.s], c.eatom + un.row0);
        if (c.launches) (*c.launches)++;
      }
    }
    MultiParams b{};   // backward stages: stage nh + i of unit t - (nh - 1) - i
    for (int i = 0; i < nh; i++) {
      const int u = t - (nh - 1) - i;
      if (u < 0 || u >= U) continue;
      const Unit& un = units[u];
      if (make_stage_prob(b.p[b.nprob], nh + i, un, *W[un.s], ring[u % depth], *dims[un.s], c)) return 1;
      b.nprob++;
    }
    if (launch_multi(c, b)) return 1;
  }
  return 0;
}

}  // namespace f16
}  // namespace anib
