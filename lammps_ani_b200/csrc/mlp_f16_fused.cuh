// mlp_f16_fused.cuh -- the whole ensemble MLP of one species bucket, forward and dE/dAEV backward, in ONE tcgen05 kernel.
//
// Replaces BmmEnsemble's torch.bmm chain, its bias / CELU kernels and their autograd backward
// (models/lammps_ani.py:219-257,203-214; K5/K7 of SURVEY.md 2.1), like mlp_f16.cuh, whose f16x3 arithmetic it shares
// (scaled fp16 hi/lo operand planes, A_hi.B_hi + A_lo.B_hi + A_hi.B_lo in fp32 TMEM accumulators).
//
// Why: the layer-by-layer kernels of mlp_f16.cuh write every activation to HBM as hi/lo planes and read it back two or
// three times (23 GB per 300 k-atom step, 80 % of HBM peak, tensor pipe 20-48 %).  Here a work item = 128 atom rows x one
// ensemble member runs all six GEMMs back to back on one SM:
//
//   S0  AEV.W1^T  -> h1 = CELU      S1  h1.W2^T -> h2 = CELU       S2  h2.W3^T -> h3, e = <h3,w4>+b4, g3 = w4*CELU'
//   S3  g3.W3 * CELU'(z2) -> g2     S4  g2.W2 * CELU'(z1) -> g1    S5  g1.W1 -> dE/dAEV (summed over members)
//
// The epilogue of stage s converts the accumulator IN PLACE in tensor memory: 16 fp32 columns become 8 columns of packed fp16
// hi values and 8 of lo values (tcgen05.st), which is exactly the K = 16 slice of an A operand held in TMEM, and stage s+1's
// MMAs (tcgen05.mma with A from TMEM, B = the weights from shared memory) consume each 16-column granule as soon as it is
// complete -- the activations never leave the SM, never even touch shared memory.  (An earlier version staged them in shared
// memory: slot hand-offs, proxy fences and st.shared made the per-block chain three times longer.)  Only the last stage's output
// goes through shared memory, as the source of a TMA store / reduce.  Only three things touch global memory besides the weights (streamed from L2 through a
// 4 x 32 KB TMA ring) and the AEV tile: CELU'(z1), CELU'(z2) parked in a per-CTA fp32 scratch (written and read back by
// the SAME thread, so no synchronisation; 224 KB per CTA, L2-resident), the member energies, and the dE/dAEV tile, which
// is added into the output rows with a TMA reduce (cp.reduce.async.bulk.tensor .add) because the eight members of a row
// tile are different work items.  Accumulators alternate between TMEM columns 0 and 256.
//
// Roles (640 threads): warp 0 TMA producer (weight ring, AEV ring), warp 1 MMA issuer, warp 2 TMEM allocator + TMA
// store / reduce warp, warp 3 idle, warps 4-19 epilogue: warp (q, j) serves TMEM lane quadrant q (rows 32q..32q+31) and the
// granules gi = j, j+4, ... of every stage.
// fp16 range guard: every scaled value written to an operand plane is checked against the fp16 maximum; an overflow raises
// a flag the host turns into an error (nothing silent).
#pragma once
#include "mlp_f16.cuh"

namespace anib {
namespace f16 {

constexpr int kFusedMaxStages = 6;
constexpr int kFG = 4;                          // epilogue groups (four warps each)
constexpr int kFH = 2;                          // output-block slots of the last stage (a power of two <= kFG)
constexpr int kFThreads = 128 + 128 * kFG;
constexpr int kFNB = 4;                         // weight ring depth
constexpr int kFNA = 4;                         // AEV ring depth: the next item's whole tile (K <= 128) is prefetched during the current
                                                // item's last stage; with two blocks stage 0 waited ~1.3 k cycles for each of the others
constexpr int kFBSlot = 32768;                  // one weight k-block: 256 rows x 32 halves x (hi + lo)
constexpr int kFBlk = 16384;                    // 128 rows x 32 halves x (hi + lo), or 128 rows x 32 floats
constexpr int kFGran = 16;                       // 16-column granules of a stage (N <= 256)
constexpr int kFBars = 2 * kFNB + 2 * kFNA + 3 * kFG + 4 + 2 + kFGran;
constexpr float kF16Max = 65504.f;
// Scale of the activation planes: log2(e) / alpha, so that u = scale * z feeds ex2 directly (e^(z/alpha) = 2^u) and
// scale * CELU(z) = max(u, c (min(2^u, 1) - 1)) with c = scale * alpha: five instructions per element.  Not a power of two,
// which only means the next layer's accumulator scale 1 / (scale * sw) is a rounded fp32 constant like any other factor.
constexpr float kFScaleH = 14.426950408889634f;
constexpr float kFCeluC = 1.4426950408889634f;

struct FusedArgs {
  int R, num_row_tiles, batch, a_row0;
  int nstages;                                  // 6: dE/dAEV fused (TMA reduce); 5: g1 planes stored for a separate dE/dAEV GEMM
  int K[kFusedMaxStages], N[kFusedMaxStages];
  int b_row_off_per_z[kFusedMaxStages], b_k_off_per_z[kFusedMaxStages];
  float acc_scale[kFusedMaxStages];
  const float* bias0; const float* bias1; const float* bias2;   // [M][n]: kFScaleH * b1, b2, b3
  const float* w_out_s; const float* b_out;                      // kScaleG*w4 [M][n3], b4 [M]
  int cslot_bytes;                              // one slot of per-stage constants in shared memory (bias [+ w4])
  float* e_out; int e_ld;
  float* scratch; long long scratch_per_cta;    // floats
  int g1_col_off_per_z;
  int* flag;                                    // |= 1 when a plane value left the fp16 range
  int* dbg;                                     // host-mapped words of the wait watchdog
  long long* prof;                              // optional cycle counters, 8 per CTA, then 64 trace stamps of CTA 0's third item
  int prof_ctas;
};

struct alignas(64) FusedMaps {
  CUtensorMap a_hi, a_lo;
  CUtensorMap b_hi[kFusedMaxStages], b_lo[kFusedMaxStages];
  CUtensorMap out0, out1;                       // dE/dAEV rows (fp32, reduce-add)  |  g1 hi / lo planes
};

// byte offset of 16-byte chunk ch of row r (0..127) in a block of 64-byte rows under SWIZZLE_64B / 128-byte rows under SWIZZLE_128B
__device__ __forceinline__ uint32_t blk_off64(int r, int ch) { return (uint32_t)(r * 64 + ((ch ^ ((r >> 1) & 3)) << 4)); }
__device__ __forceinline__ uint32_t blk_off128(int r, int ch) { return (uint32_t)(r * 128 + ((ch ^ (r & 7)) << 4)); }

// One lane of a fully converged warp.  The MMA / TMA warps keep their whole loop warp-uniform and only predicate the issuing
// instructions with this: descriptors and addresses then live in uniform registers.  (Under an `if (lane == 0)` around the
// loop the compiler wraps every tcgen05.mma in an ELECT + 7 x R2UR.BROADCAST "waterfall", ~125 cycles per instruction --
// more than a 128 x 128 x 16 MMA takes to execute.)
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "elect.sync _|p, 0xffffffff;\n"
      "selp.u32 %0, 1, 0, p;\n"
      "}\n" : "=r"(pred));
  return pred != 0;
}

// 8 consecutive accumulator columns of this thread's TMEM lane.  The _nowait form only issues the load; tmem_ld_wait8 waits
// for it and hands the values over (the "+r" operands keep the compiler from reading them earlier).
__device__ __forceinline__ void tmem_ld8_nowait(uint32_t taddr, uint32_t (&v)[8]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]) : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld_wait8(uint32_t (&vn)[8], uint32_t (&v)[8]) {
  asm volatile("tcgen05.wait::ld.sync.aligned;" : "+r"(vn[0]), "+r"(vn[1]), "+r"(vn[2]), "+r"(vn[3]), "+r"(vn[4]), "+r"(vn[5]), "+r"(vn[6]),
               "+r"(vn[7]) :: "memory");
#pragma unroll
  for (int i = 0; i < 8; i++) v[i] = vn[i];
}
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, uint32_t (&v)[8]) {
  tmem_ld8_nowait(taddr, v);
  asm volatile("tcgen05.wait::ld.sync.aligned;" : "+r"(v[0]), "+r"(v[1]), "+r"(v[2]), "+r"(v[3]), "+r"(v[4]), "+r"(v[5]), "+r"(v[6]), "+r"(v[7])
               :: "memory");
}

// D[tmem] (+)= A[tmem] . B[smem]: the A operand is read from tensor memory (128 lanes x 8 columns of packed fp16 pairs per K = 16)
__device__ __forceinline__ void umma_f16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t db, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n"
      "}\n" ::"r"(tmem_d), "r"(tmem_a), "l"(db), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void tmem_st8(uint32_t taddr, const uint32_t (&v)[8]) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]),
               "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]) : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_ld16w(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
        "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" : "+r"(v[0]), "+r"(v[1]), "+r"(v[2]), "+r"(v[3]), "+r"(v[4]), "+r"(v[5]), "+r"(v[6]), "+r"(v[7]),
               "+r"(v[8]), "+r"(v[9]), "+r"(v[10]), "+r"(v[11]), "+r"(v[12]), "+r"(v[13]), "+r"(v[14]), "+r"(v[15]) :: "memory");
}

// plain (non-tensor) bulk copy global -> shared, completion on an mbarrier
__device__ __forceinline__ void bulk_load(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src), "r"(bytes),
               "r"(bar) : "memory");
}
__device__ __forceinline__ float4 lds_f4(uint32_t addr) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
  return v;
}

// mbarrier wait with a watchdog: a protocol error becomes a trap with a code in host-mapped memory instead of a hung GPU
__device__ __forceinline__ void mbar_wait_wd(uint32_t bar, uint32_t parity, int* dbg, uint32_t code) {
  uint32_t done = 0;
  long long t0 = 0;
  for (uint32_t spins = 0;; spins++) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n" : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    if (done) return;
    if ((spins & 0xfffu) == 0xfffu) {
      const long long t = clock64();
      if (t0 == 0) t0 = t;
      else if (t - t0 > 4000000000ll) {      // ~2 s: record who waits where (one word per warp), trap a little later
        if (dbg) { atomicCAS(dbg + ((code >> 24) & 31u), 0, (int)code); __threadfence_system(); }
        if (t - t0 > 6000000000ll) __trap();
      }
    }
  }
}

__device__ __forceinline__ void tma_reduce_add_2d(const CUtensorMap* tm, uint32_t src, int c0, int c1) {
  asm volatile("cp.reduce.async.bulk.tensor.2d.global.shared::cta.add.tile.bulk_group [%0, {%2, %3}], [%1];" ::"l"(tm), "r"(src),
               "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ int fused_uses(int nb, int g) { return (nb - g + kFH - 1) / kFH; }   // blocks cb < nb with cb % kFH == g

// ---- epilogue pieces: 8 columns at a time (one 16-byte chunk of the hi plane and one of the lo plane)
__device__ __forceinline__ void put_planes8(uint32_t blk, int r, int ch, const float (&x)[8], float& amax) {
  uint32_t hh[4], ll[4];
#pragma unroll
  for (int e = 0; e < 4; e++) {
    amax = fmaxf(fmaxf(fabsf(x[2 * e]), fabsf(x[2 * e + 1])), amax);
    split2(x[2 * e], x[2 * e + 1], hh[e], ll[e]);
  }
  const uint32_t off = blk_off64(r, ch);
  sts128(blk + off, hh[0], hh[1], hh[2], hh[3]);
  sts128(blk + kFBlk / 2 + off, ll[0], ll[1], ll[2], ll[3]);
}

template <bool PROF>
__global__ void __launch_bounds__(kFThreads, 1)
f16_fused_kernel(const __grid_constant__ FusedMaps maps, const __grid_constant__ FusedArgs g) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  if ((raw & 1023u) != 0u) __trap();             // the operand rings need 1024-byte alignment and there is no room for slack
  const uint32_t baseB = raw;
  const uint32_t baseA = baseB + (uint32_t)kFNB * kFBSlot;
  const uint32_t baseH = baseA + (uint32_t)kFNA * kFBlk;
  const uint32_t baseC = baseH + (uint32_t)kFH * kFBlk;           // two slots of per-stage constants
  const uint32_t bar_base = baseC + 2u * (uint32_t)g.cslot_bytes;
  auto bfull = [&](int i) { return bar_base + 8u * (uint32_t)i; };
  auto bempty = [&](int i) { return bar_base + 8u * (uint32_t)(kFNB + i); };
  auto afull = [&](int i) { return bar_base + 8u * (uint32_t)(2 * kFNB + i); };
  auto aempty = [&](int i) { return bar_base + 8u * (uint32_t)(2 * kFNB + kFNA + i); };
  auto hfull = [&](int i) { return bar_base + 8u * (uint32_t)(2 * kFNB + 2 * kFNA + i); };
  auto hfree = [&](int i) { return bar_base + 8u * (uint32_t)(2 * kFNB + 2 * kFNA + kFG + i); };
  auto accfull = [&](int i) { return bar_base + 8u * (uint32_t)(2 * kFNB + 2 * kFNA + 2 * kFG + i); };
  auto accfree = [&](int i) { return bar_base + 8u * (uint32_t)(2 * kFNB + 2 * kFNA + 2 * kFG + 2 + i); };
  // the last stage's blocks are announced on their own barriers: a waiter must see EVERY phase of a barrier it waits on
  // (parity waits alias two phases apart), and the store warp sees only those blocks, the MMA issuer only the others
  auto dfull = [&](int i) { return bar_base + 8u * (uint32_t)(2 * kFNB + 2 * kFNA + 2 * kFG + 4 + i); };
  auto cfull = [&](int i) { return bar_base + 8u * (uint32_t)(2 * kFNB + 2 * kFNA + 3 * kFG + 4 + i); };
  auto gfull = [&](int i) { return bar_base + 8u * (uint32_t)(2 * kFNB + 2 * kFNA + 3 * kFG + 4 + 2 + i); };   // granules 2i, 2i+1 of the current stage converted
  const uint32_t tmem_slot = bar_base + 8u * (uint32_t)kFBars;
  volatile uint32_t* tmem_slot_ptr = (volatile uint32_t*)(smem_raw + (tmem_slot - raw));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#define WD(grp_, idx_) (0x80000000u | ((uint32_t)warp << 24) | ((uint32_t)(grp_) << 20) | (((uint32_t)(idx_) & 15u) << 16) | ((uint32_t)__LINE__ & 0xffffu))
  const int num_items = g.num_row_tiles * g.batch;
  const int ns = g.nstages;

  if (warp == 1 && lane == 0) {
    for (int i = 0; i < kFNB; i++) { mbar_init(bfull(i), 1); mbar_init(bempty(i), 1); }
    for (int i = 0; i < kFNA; i++) { mbar_init(afull(i), 1); mbar_init(aempty(i), 1); }
    // an output block is announced by the warps that wrote it: all sixteen for an fp32 dE/dAEV block, eight for a g1 block
    for (int i = 0; i < kFG; i++) { mbar_init(hfull(i), 1); mbar_init(hfree(i), 1); mbar_init(dfull(i), g.nstages == kFusedMaxStages ? 4 * kFG : 8); }
    for (int i = 0; i < 2; i++) { mbar_init(accfull(i), 1); mbar_init(accfree(i), 4 * kFG); mbar_init(cfull(i), 1); }
    for (int i = 0; i < kFGran; i++) mbar_init(gfull(i), 8);       // a pair of granules: two warps of each quadrant
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
    // ================================================================= TMA producer: weight ring + AEV ring
    if (lane == 0) {
      uint32_t bq = 0, aq = 0;
      for (int it = blockIdx.x; it < num_items; it += gridDim.x) {
        const int rt = it / g.batch, z = it % g.batch;
        const int a_row = g.a_row0 + rt * BM;
        for (int s = 0; s < ns; s++) {
          const int KB = (g.K[s] + 31) / 32;
          const uint32_t bplane = (uint32_t)g.N[s] * 64u;
          const int bcol = z * g.b_k_off_per_z[s], brow = z * g.b_row_off_per_z[s];
          for (int kb = 0; kb < KB; kb++) {
            if (s == 0) {
              const int sa = (int)(aq % (uint32_t)kFNA);
              mbar_wait_wd(aempty(sa), ((aq / (uint32_t)kFNA) & 1u) ^ 1u, g.dbg, WD(3, sa));
              const uint32_t da = baseA + (uint32_t)sa * kFBlk;
              mbar_expect_tx(afull(sa), kFBlk);
              tma_load_2d(&maps.a_hi, afull(sa), da, kb * 32, a_row);
              tma_load_2d(&maps.a_lo, afull(sa), da + kFBlk / 2, kb * 32, a_row);
              aq++;
            }
            const int sb = (int)(bq % (uint32_t)kFNB);
            mbar_wait_wd(bempty(sb), ((bq / (uint32_t)kFNB) & 1u) ^ 1u, g.dbg, WD(1, sb));
            const uint32_t dst = baseB + (uint32_t)sb * kFBSlot;
            mbar_expect_tx(bfull(sb), 2u * bplane);
            tma_load_2d(&maps.b_hi[s], bfull(sb), dst, bcol + kb * 32, brow);
            tma_load_2d(&maps.b_lo[s], bfull(sb), dst + bplane, bcol + kb * 32, brow);
            bq++;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ================================================================= MMA issuer (warp-uniform loop, one elected lane issues)
    {
      uint32_t bq = 0, aq = 0, t = 0, gpar = 0;     // gpar bit i: parity of the uses of granule barrier i so far
      long long w_acc = 0, w_b = 0, w_a = 0, w_i = 0, t_begin = 0;
      if (PROF) t_begin = clock64();
      // weight k-blocks this CTA consumes in total: the wait for the NEXT one is taken off the critical path (it is issued right
      // after the current k-block's MMAs, while they run; the block has usually landed long before)
      uint32_t total_kb = 0;
      {
        uint32_t per_item = 0;
        for (int s = 0; s < ns; s++) per_item += (uint32_t)((g.K[s] + 31) / 32);
        const int my_items = blockIdx.x < (unsigned)num_items ? (num_items - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;
        total_kb = per_item * (uint32_t)my_items;
      }
      if (total_kb > 0) mbar_wait_wd(bfull(0), 0u, g.dbg, WD(0, 0));
      for (int it = blockIdx.x; it < num_items; it += gridDim.x) {
        const int z = it % g.batch;
        for (int s = 0; s < ns; s++, t++) {
          const uint32_t r = t & 1u;
          long long c0 = 0;
          if (PROF) c0 = clock64();
          if (t >= 2) {                            // the epilogue that last read these TMEM columns (global stage t-2) is done
            mbar_wait_wd(accfree(r), ((t >> 1) - 1u) & 1u, g.dbg, WD(9, r));
            tc_fence_after();
          }
          if (PROF) w_acc += clock64() - c0;
          if (s <= 2) {
            // this stage's epilogue constants (bias, and w4 for S2) into constants slot r: its previous reader was the epilogue of
            // global stage t-2, the one the wait above covers
            const uint32_t nbytes = (uint32_t)g.N[s] * 4u, dstc = baseC + r * (uint32_t)g.cslot_bytes;
            if (elect_one()) {
              mbar_expect_tx(cfull(r), s == 2 ? 2u * nbytes : nbytes);
              bulk_load(dstc, (s == 0 ? g.bias0 : s == 1 ? g.bias1 : g.bias2) + (size_t)z * g.N[s], nbytes, cfull(r));
              if (s == 2) bulk_load(dstc + nbytes, g.w_out_s + (size_t)z * g.N[2], nbytes, cfull(r));
            }
          }
          const uint32_t tmem_d = tmem_base + r * 256u;
          const uint32_t idesc = umma_idesc_f16(BM, g.N[s]);
          const int K = g.K[s], KB = (K + 31) / 32;
          const uint32_t bplane = (uint32_t)g.N[s] * 64u;
          const uint32_t tmem_a0 = tmem_base + (r ^ 1u) * 256u;      // stage s-1's accumulator, converted in place (s >= 1)
          const bool tr = PROF && g.prof && blockIdx.x == 0 && it == (int)(2 * gridDim.x) && lane == 0;
          long long* const trp = g.prof + (size_t)g.prof_ctas * 8 + s * 8;
          if (tr) trp[0] = clock64();                                 // stage start: TMEM region free
          for (int kb = 0; kb < KB; kb++) {
            const int sb = (int)(bq % (uint32_t)kFNB);      // (its weights were waited for one k-block ago)
            if (PROF) c0 = clock64();
            const uint32_t sB = baseB + (uint32_t)sb * kFBSlot;
            const uint64_t dB_hi = umma_desc_k(sB, 32), dB_lo = umma_desc_k(sB + bplane, 32);
            const int nks = min(2, (K - kb * 32) / 16);
            if (s == 0) {
              // A = the AEV tile from shared memory (TMA)
              const int sa = (int)(aq % (uint32_t)kFNA);
              mbar_wait_wd(afull(sa), (aq / (uint32_t)kFNA) & 1u, g.dbg, WD(2, sa));
              if (PROF) { const long long c1 = clock64(); w_a += c1 - c0; c0 = c1; }
              tc_fence_after();
              const uint32_t sA = baseA + (uint32_t)sa * kFBlk;
              const uint64_t dA_hi = umma_desc_k(sA, 32), dA_lo = umma_desc_k(sA + kFBlk / 2, 32);
              if (elect_one()) {
                for (int k = 0; k < nks; k++) {
                  const uint64_t adv = (uint64_t)(2 * k);          // 16 halves = 32 bytes along K, in 16-byte units
                  umma_f16(tmem_d, dA_hi + adv, dB_hi + adv, idesc, (kb | k) != 0 ? 1u : 0u);
                  umma_f16(tmem_d, dA_lo + adv, dB_hi + adv, idesc, 1u);
                  umma_f16(tmem_d, dA_hi + adv, dB_lo + adv, idesc, 1u);
                }
                umma_commit(bempty(sb));
                umma_commit(aempty(sa));
              }
              __syncwarp();
              aq++;
            } else {
              // A = granules 2 kb, 2 kb + 1 of the previous stage, fp16 hi / lo pairs in tensor memory; one barrier per pair
              // (the two granules belong to neighbouring warps of each quadrant and land together)
              mbar_wait_wd(gfull(kb), (gpar >> kb) & 1u, g.dbg, WD(4, kb));
              gpar ^= 1u << kb;
              if (tr && kb == 0) trp[1] = clock64();                  // first granule pair there
              if (tr && kb == KB - 1) trp[2] = clock64();             // last granule pair there
              if (PROF) { const long long c1 = clock64(); w_a += c1 - c0; c0 = c1; }
              tc_fence_after();
              if (elect_one()) {
                const uint32_t ta = tmem_a0 + (uint32_t)(32 * kb);
                umma_f16_ts(tmem_d, ta, dB_hi, idesc, kb != 0 ? 1u : 0u);
                umma_f16_ts(tmem_d, ta + 8u, dB_hi, idesc, 1u);
                umma_f16_ts(tmem_d, ta, dB_lo, idesc, 1u);
                umma_f16_ts(tmem_d, ta + 16u, dB_hi + 2u, idesc, 1u);
                umma_f16_ts(tmem_d, ta + 24u, dB_hi + 2u, idesc, 1u);
                umma_f16_ts(tmem_d, ta + 16u, dB_lo + 2u, idesc, 1u);
                umma_commit(bempty(sb));
              }
              __syncwarp();
            }
            bq++;
            if (PROF) { const long long c1 = clock64(); w_i += c1 - c0; c0 = c1; }
            if (bq < total_kb) {
              const int sn = (int)(bq % (uint32_t)kFNB);
              mbar_wait_wd(bfull(sn), (bq / (uint32_t)kFNB) & 1u, g.dbg, WD(0, sn));
            }
            if (PROF) w_b += clock64() - c0;
          }
          if (tr) trp[3] = clock64();                                 // all MMAs of the stage issued
          if (elect_one()) umma_commit(accfull(r));
          __syncwarp();
        }
      }                                           // (the last stage's blocks go to the store warp, on dfull)
      if (PROF && g.prof && lane == 0) {
        long long* p = g.prof + (size_t)blockIdx.x * 8;
        p[0] = clock64() - t_begin; p[1] = w_acc; p[2] = w_b; p[3] = w_a; p[7] = w_i;
      }
    }
  } else if (warp == 2) {
    // ================================================================= store warp: the last stage's blocks -> global
    if (lane == 0) {
      uint32_t dcnt[kFG];
#pragma unroll
      for (int i = 0; i < kFG; i++) dcnt[i] = 0;
      const int nb = g.N[ns - 1] >> 5;
      for (int it = blockIdx.x; it < num_items; it += gridDim.x) {
        const int rt = it / g.batch, z = it % g.batch;
        for (int cb = 0; cb < nb; cb++) {
          const int slot = cb % kFH;
          mbar_wait_wd(dfull(slot), dcnt[slot] & 1u, g.dbg, WD(10, slot));
          dcnt[slot]++;
          const uint32_t src = baseH + (uint32_t)slot * kFBlk;
          if (ns == kFusedMaxStages) {
            tma_reduce_add_2d(&maps.out0, src, cb * 32, rt * BM);
          } else {
            const int col = z * g.g1_col_off_per_z + cb * 32;
            tma_store_2d(&maps.out0, src, col, rt * BM);
            tma_store_2d(&maps.out1, src + kFBlk / 2, col, rt * BM);
          }
          tma_store_commit();
          tma_store_wait_read<0>();
          mbar_arrive(hfree(slot));
        }
      }
      tma_store_wait_all();
    }
  } else if (warp >= 4) {
    // ================================================================= epilogue warps
    // Warp (q, j) owns TMEM lanes 32q..32q+31 and the 16-column granules gi = j, j+4, ... of every stage.  A granule of the
    // accumulator is read (tcgen05.ld), turned into the next stage's operand (bias + CELU / * CELU', fp16 hi + lo split) and
    // written back over itself (tcgen05.st: 8 columns hi, 8 columns lo); the MMA warp is told per granule.  The four warps of a
    // quadrant work on four different granules at once, so granules complete in bursts of four, 64 columns = two weight
    // k-blocks, and the next stage's MMAs trail by at most that.
    // The warp scheduler favours the highest warp ids of a sub-partition: they take the EARLIEST granules, so that the first
    // granule pair of a stage -- what the next stage's MMAs wait for -- is finished first instead of last.
    const int q = warp & 3, j8 = 3 - ((warp - 4) >> 2);           // lane quadrant, granule phase (and 8-column slice in S5)
    const int r = q * 32 + lane;                                  // row inside the 128-row tile
    float4* const scr = (float4*)(g.scratch + (size_t)blockIdx.x * g.scratch_per_cta);
    const int n1 = g.N[0];
    uint32_t hpar = 0, t = 0, cuse[2] = {0, 0};                   // hpar bit i: parity of this warp's uses of output-block slot i
    long long w_acc = 0, w_h = 0, t_begin = 0;
    const bool prof_me = PROF && warp == 16 && lane == 0;     // (the warp that takes granule 0)
    if (prof_me) t_begin = clock64();
    // float4 index of (granule gi, quadrant q, vector 0, lane) in a CELU' scratch area (vector v is 32 float4 further on): a
    // warp's store / load of one vector is 512 contiguous bytes, and a value is read back by the thread that wrote it
    auto dgran0 = [&](int gi) { return ((gi * 4 + q) * 4) * 32 + lane; };
    const uint32_t quad_taddr = tmem_base + ((uint32_t)(q * 32) << 16);

    for (int it = blockIdx.x; it < num_items; it += gridDim.x) {
      const int rt = it / g.batch, z = it % g.batch;
      const int row = rt * BM + r;
      float amax = 0.f, e_acc = 0.f;

      const bool etr = prof_me && g.prof && blockIdx.x == 0 && it == (int)(2 * gridDim.x);
      int estage = 0;
      auto wait_acc = [&]() {
        long long c0 = 0;
        if (prof_me) c0 = clock64();
        mbar_wait_wd(accfull(t & 1u), (t >> 1) & 1u, g.dbg, WD(8, t & 1u));
        if (prof_me) w_acc += clock64() - c0;
        tc_fence_after();
        if (etr) g.prof[(size_t)g.prof_ctas * 8 + estage * 8 + 4] = clock64();     // accumulator seen complete
        estage++;
      };
      auto wait_consts = [&]() {                 // returns the shared-memory address of this stage's constants
        const uint32_t rr = t & 1u;
        mbar_wait_wd(cfull(rr), cuse[rr] & 1u, g.dbg, WD(11, rr));
        cuse[rr]++;
        return baseC + rr * (uint32_t)g.cslot_bytes;
      };
      auto wait_slot = [&](int slot) {           // the previous block in this output slot has left (TMA store / reduce read it)
        long long c0 = 0;
        if (prof_me) c0 = clock64();
        mbar_wait_wd(hfree(slot), ((hpar >> slot) & 1u) ^ 1u, g.dbg, WD(5, slot));
        hpar ^= 1u << slot;
        if (prof_me) w_h += clock64() - c0;
      };
      auto end_stage = [&]() {
        if (etr) g.prof[(size_t)g.prof_ctas * 8 + (estage - 1) * 8 + 6] = clock64();   // this warp's last granule of the stage written
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(accfree(t & 1u));
        t++;
      };
      auto gaddr = [&](int gi) { return quad_taddr + (t & 1u) * 256u + (uint32_t)(16 * gi); };
      // 16 values -> fp16 hi / lo pairs -> back over the granule they came from; tell the MMA warp
      auto emit_tmem = [&](int gi, const float (&x)[16]) {
        uint32_t hh[8], ll[8];
#pragma unroll
        for (int e = 0; e < 8; e++) {
          amax = fmaxf(fmaxf(fabsf(x[2 * e]), fabsf(x[2 * e + 1])), amax);
          split2(x[2 * e], x[2 * e + 1], hh[e], ll[e]);
        }
        const uint32_t ta = gaddr(gi);
        tmem_st8(ta, hh);
        tmem_st8(ta + 8u, ll);
        tmem_st_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(gfull(gi >> 1));
        if (etr && gi == j8) g.prof[(size_t)g.prof_ctas * 8 + (estage - 1) * 8 + 5] = clock64();   // this warp's first granule written
      };
      // the same 16 values into an output block in shared memory (the K-major SWIZZLE_64B layout of the g1 planes' TMA store);
      // granules 2 cb and 2 cb + 1 make block cb
      auto emit_smem = [&](int gi, const float (&x)[16]) {
        const int slot = (gi >> 1) & (kFH - 1);
        wait_slot(slot);
        const uint32_t dstb = baseH + (uint32_t)slot * kFBlk;
#pragma unroll
        for (int hchunk = 0; hchunk < 2; hchunk++) {
          uint32_t hh[4], ll[4];
#pragma unroll
          for (int e = 0; e < 4; e++) {
            const float x0 = x[hchunk * 8 + 2 * e], x1 = x[hchunk * 8 + 2 * e + 1];
            amax = fmaxf(fmaxf(fabsf(x0), fabsf(x1)), amax);
            split2(x0, x1, hh[e], ll[e]);
          }
          const uint32_t off = blk_off64(r, 2 * (gi & 1) + hchunk);
          sts128(dstb + off, hh[0], hh[1], hh[2], hh[3]);
          sts128(dstb + kFBlk / 2 + off, ll[0], ll[1], ll[2], ll[3]);
        }
        fence_async_smem();
        __syncwarp();
        if (lane == 0) mbar_arrive(dfull(slot));
      };

      // ---------------- S0, S1: u = kFScaleH * z = acc*a + b';  h = CELU -> operand (x kFScaleH);  CELU'(z) -> scratch
#pragma unroll 1
      for (int s = 0; s < 2; s++) {
        const int ng = g.N[s] >> 4;
        const float a = kFScaleH * g.acc_scale[s];
        float4* const dst = scr + (s == 0 ? 0 : n1 * 32);       // n1*128 floats = n1*32 float4
        wait_acc();
        const uint32_t cb_s = wait_consts();
#pragma unroll 1
        for (int gi = j8; gi < ng; gi += 4) {
          uint32_t v[16];
          tmem_ld16w(gaddr(gi), v);
          float h[16];
          float4* const dp = dst + dgran0(gi);
#pragma unroll
          for (int c4 = 0; c4 < 4; c4++) {
            const float4 bq = lds_f4(cb_s + (uint32_t)gi * 64u + c4 * 16u);
            const float bb[4] = {bq.x, bq.y, bq.z, bq.w};
            float d[4];
#pragma unroll
            for (int e = 0; e < 4; e++) {
              const float u = fmaf(__uint_as_float(v[c4 * 4 + e]), a, bb[e]);   // log2(e)/alpha * z
              d[e] = fminf(exp2f(u), 1.f);                                      // CELU'(z) = min(e^(z/alpha), 1)
              h[c4 * 4 + e] = fmaxf(u, fmaf(kFCeluC, d[e], -kFCeluC));          // kFScaleH * CELU(z)   (e^x - 1 >= x)
            }
            dp[c4 * 32] = make_float4(d[0], d[1], d[2], d[3]);      // plain (weak) store: read back by this same thread only
          }
          emit_tmem(gi, h);
        }
        end_stage();
      }
      // ---------------- S2: h3 = CELU(z3); e += <h3, w4>; g3 = w4 * CELU'(z3) -> operand (x kScaleG)
      {
        const int ng = g.N[2] >> 4;
        const float a = kFScaleH * g.acc_scale[2];
        wait_acc();
        const uint32_t cb_s = wait_consts();
        const uint32_t w_s = cb_s + (uint32_t)g.N[2] * 4u;
#pragma unroll 1
        for (int gi = j8; gi < ng; gi += 4) {
          uint32_t v[16];
          tmem_ld16w(gaddr(gi), v);
          float gg[16];
#pragma unroll
          for (int c4 = 0; c4 < 4; c4++) {
            const float4 bq = lds_f4(cb_s + (uint32_t)gi * 64u + c4 * 16u), wq = lds_f4(w_s + (uint32_t)gi * 64u + c4 * 16u);
            const float bb[4] = {bq.x, bq.y, bq.z, bq.w}, ww[4] = {wq.x, wq.y, wq.z, wq.w};
#pragma unroll
            for (int e = 0; e < 4; e++) {
              const float u = fmaf(__uint_as_float(v[c4 * 4 + e]), a, bb[e]);
              const float d = fminf(exp2f(u), 1.f);
              const float h = fmaxf(u, fmaf(kFCeluC, d, -kFCeluC));
              e_acc = fmaf(h, ww[e], e_acc);                    // x kFScaleH * kScaleG, folded back below
              gg[c4 * 4 + e] = ww[e] * d;
            }
          }
          emit_tmem(gi, gg);
        }
        end_stage();
        if (row < g.R)                                          // e_out is zeroed before the launch; the warps of a row add their parts
          atomicAdd(g.e_out + (size_t)row * g.e_ld + z, e_acc * (1.0f / (kFScaleH * kScaleG)) + (j8 == 0 ? __ldg(g.b_out + z) : 0.f));
      }
      // ---------------- S3, S4: g = acc * s * CELU'(z) (from scratch) -> operand; S4's output leaves through shared memory when the
      // dE/dAEV GEMM is not fused (ns == 5)
#pragma unroll 1
      for (int s = 3; s < 5; s++) {
        const int ng = g.N[s] >> 4;
        const float a = g.acc_scale[s];
        const float4* const src = scr + (s == 3 ? n1 * 32 : 0);
        const bool to_smem = (s == ns - 1);
        float4 dd[4];
        if (j8 < ng) {                                          // the first granule's CELU' does not need the accumulator
          const float4* const dp = src + dgran0(j8);
#pragma unroll
          for (int c4 = 0; c4 < 4; c4++) dd[c4] = __ldcg(dp + c4 * 32);
        }
        wait_acc();
#pragma unroll 1
        for (int gi = j8; gi < ng; gi += 4) {
          uint32_t v[16];
          tmem_ld16w(gaddr(gi), v);
          float gg[16];
#pragma unroll
          for (int c4 = 0; c4 < 4; c4++) {
            const float d4[4] = {dd[c4].x, dd[c4].y, dd[c4].z, dd[c4].w};
#pragma unroll
            for (int e = 0; e < 4; e++) gg[c4 * 4 + e] = __uint_as_float(v[c4 * 4 + e]) * a * d4[e];
          }
          if (gi + 4 < ng) {                                    // next granule's CELU' while this one is written back
            const float4* const dp = src + dgran0(gi + 4);
#pragma unroll
            for (int c4 = 0; c4 < 4; c4++) dd[c4] = __ldcg(dp + c4 * 32);
          }
          if (to_smem) emit_smem(gi, gg); else emit_tmem(gi, gg);
        }
        end_stage();
      }
      // 8-column slices of a 32-column block for S5 (all sixteen warps on one fp32 output block)
      auto tslice = [&](int cb) { return quad_taddr + (uint32_t)(8 * j8) + (t & 1u) * 256u + (uint32_t)(cb * 32); };
      auto publish = [&](int slot, bool) {
        fence_async_smem();
        __syncwarp();
        if (lane == 0) mbar_arrive(dfull(slot));
      };
      // ---------------- S5: dE/dAEV tile of this member -> fp32 block (SWIZZLE_128B) -> TMA reduce-add
      if (ns == kFusedMaxStages) {
        const int nb = g.N[5] >> 5;
        const float a = g.acc_scale[5];
        wait_acc();
#pragma unroll 1
        for (int cb = 0; cb < nb; cb++) {
          const int slot = cb & (kFH - 1);
          uint32_t v[8];
          tmem_ld8(tslice(cb), v);
          wait_slot(slot);
          const uint32_t dst = baseH + (uint32_t)slot * kFBlk;
          sts128(dst + blk_off128(r, 2 * j8), __float_as_uint(__uint_as_float(v[0]) * a), __float_as_uint(__uint_as_float(v[1]) * a),
                 __float_as_uint(__uint_as_float(v[2]) * a), __float_as_uint(__uint_as_float(v[3]) * a));
          sts128(dst + blk_off128(r, 2 * j8 + 1), __float_as_uint(__uint_as_float(v[4]) * a), __float_as_uint(__uint_as_float(v[5]) * a),
                 __float_as_uint(__uint_as_float(v[6]) * a), __float_as_uint(__uint_as_float(v[7]) * a));
          publish(slot, true);
        }
        end_stage();
      }
      if (row < g.R && !(amax < kF16Max)) atomicOr(g.flag, 1);
    }
    if (prof_me && g.prof) {
      long long* p = g.prof + (size_t)blockIdx.x * 8;
      p[4] = clock64() - t_begin; p[5] = w_acc; p[6] = w_h;
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
  }
#undef WD
}

// ------------------------------------------------------------------------------------------------ host side
inline bool fused_supported(const std::vector<int>& dims) {
  return supported_dims(dims) && (int)dims.size() == 5;   // three hidden layers (every ANI-1x / 2x / 1xnr network)
}

inline int* fused_dbg_word(bool device_ptr) {
  static int* h = nullptr; static int* d = nullptr;
  if (!h) {
    if (cudaHostAlloc((void**)&h, 128, cudaHostAllocMapped) != cudaSuccess) { h = nullptr; cudaGetLastError(); return nullptr; }
    for (int i = 0; i < 32; i++) h[i] = 0;
    if (cudaHostGetDevicePointer((void**)&d, h, 0) != cudaSuccess) { d = nullptr; cudaGetLastError(); }
  }
  return device_ptr ? d : h;
}

struct FusedWork {
  DevBufTc scratch, e, prof;
  HalfPlanes G1;                 // only when the AEV is wider than 256 columns (dE/dAEV as a separate GEMM)
};

struct FusedOptions {
  int* flag = nullptr;           // device word for the fp16 range guard
  int prof = 0;                  // 1: cycle counters of the MMA issuer and one epilogue warp, printed to stderr
  int cta_cap = 0;               // > 0: fewer CTAs than SMs (experiments)
};

inline int plane_block_map(CUtensorMap* hi, CUtensorMap* lo, const HalfPlanes& p, uint64_t rows, uint64_t cols) {
  return make_map(hi, p.h(), CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, rows, cols, cols, 32, BM, CU_TENSOR_MAP_SWIZZLE_64B) ||
         make_map(lo, p.l(), CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, rows, cols, cols, 32, BM, CU_TENSOR_MAP_SWIZZLE_64B);
}
inline int weight_map(CUtensorMap* hi, CUtensorMap* lo, const HalfPlanes& p, uint64_t rows, uint64_t cols, uint32_t box_rows) {
  return make_operand_map(hi, p.h(), rows, cols, pad64(cols), box_rows, 32) ||
         make_operand_map(lo, p.l(), rows, cols, pad64(cols), box_rows, 32);
}

// Forward + backward of one species bucket.  Same contract as mlp_species(); daev rows of the bucket must be ZERO on entry
// (the members' dE/dAEV tiles are added into them).
inline int mlp_species_fused(cudaStream_t st, SpeciesWeights& w, FusedWork& work, const FusedOptions& opt, int M, const std::vector<int>& dims,
                             const __half* aev_hi, const __half* aev_lo, int64_t ld_aev, int64_t aev_rows_total, int row0, int R, float* daev,
                             int64_t ld_daev, double* eatom, double sae, int* launches) {
  if (!w.ready || !fused_supported(dims)) { set_error("the fused f16 MLP needs three hidden layers with widths that are multiples of 32, <= 256"); return 1; }
  static int num_sms = 0, max_smem = 0;
  if (!num_sms) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, dev);
    cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
  }
  const int Lin = dims[0], n1 = dims[1], n2 = dims[2], n3 = dims[3];
  const int cslot = (std::max(std::max(n1, n2), 2 * n3) * 4 + 127) / 128 * 128;
  const int smem = kFNB * kFBSlot + (kFNA + kFH) * kFBlk + 2 * cslot + 8 * kFBars + 16;
  if (smem > max_smem) { set_error("not enough shared memory for the fused f16 MLP kernel"); return 1; }
  const int Npad = (Lin + 31) / 32 * 32;
  const bool fuse_d = Npad <= 256;
  const int chunk = fuse_d ? (1 << 30) : 262144;              // only the separate dE/dAEV GEMM needs a work buffer per row
  const int ctas = opt.cta_cap > 0 ? std::min(num_sms, opt.cta_cap) : num_sms;
  const long long scratch_per_cta = (long long)(n1 + n2) * BM;
  if (!work.scratch.reserve((size_t)ctas * scratch_per_cta * 4) || !work.e.reserve((size_t)std::max(R, 1) * M * 4) ||
      (opt.prof && !work.prof.reserve((size_t)(ctas * 8 + 64) * sizeof(long long))) ||
      (!fuse_d && !work.G1.reserve((size_t)std::min(chunk, (R + BM - 1) / BM * BM) * M * n1))) {
    set_error("out of memory for MLP work buffers");
    return 1;
  }
  ANIB_CUDA_OK(cudaMemsetAsync(work.e.p, 0, (size_t)R * M * 4, st));
  for (int r0 = 0; r0 < R; r0 += chunk) {
    const int Rc = std::min(chunk, R - r0);
    const int rtiles = (Rc + BM - 1) / BM;
    float* daev_chunk = daev + (size_t)(row0 + r0) * ld_daev;
    FusedMaps mp;
    memset(&mp, 0, sizeof(mp));
    FusedArgs g{};
    g.R = Rc; g.num_row_tiles = rtiles; g.batch = M; g.a_row0 = row0 + r0; g.nstages = fuse_d ? 6 : 5;
    const int Ks[6] = {Lin, n1, n2, n3, n2, n1}, Ns[6] = {n1, n2, n3, n2, n1, Npad};
    for (int s = 0; s < 6; s++) { g.K[s] = Ks[s]; g.N[s] = Ns[s]; g.b_row_off_per_z[s] = Ns[s]; }
    g.b_row_off_per_z[5] = 0; g.b_k_off_per_z[5] = n1;        // Wcat[Lin][M*n1]: member z = columns z*n1 ..
    g.acc_scale[0] = 1.0f / (kScaleAev * w.sw[0]);
    g.acc_scale[1] = 1.0f / (kFScaleH * w.sw[1]);
    g.acc_scale[2] = 1.0f / (kFScaleH * w.sw[2]);
    g.cslot_bytes = cslot;
    g.acc_scale[3] = 1.0f / w.sw[2];
    g.acc_scale[4] = 1.0f / w.sw[1];
    g.acc_scale[5] = 1.0f / (kScaleG * w.sw[0] * (float)M);
    g.bias0 = w.bias_s[0].f(); g.bias1 = w.bias_s[1].f(); g.bias2 = w.bias_s[2].f();
    g.w_out_s = w.w_out_s.f(); g.b_out = w.b_out.f();
    g.e_out = work.e.f() + (size_t)r0 * M; g.e_ld = M;
    g.scratch = work.scratch.f(); g.scratch_per_cta = scratch_per_cta;
    g.g1_col_off_per_z = n1;
    g.flag = opt.flag;
    g.dbg = fused_dbg_word(true);
    g.prof = opt.prof ? (long long*)work.prof.p : nullptr;
    g.prof_ctas = ctas;
    if (make_operand_map(&mp.a_hi, aev_hi, (uint64_t)aev_rows_total, (uint64_t)Lin, (uint64_t)ld_aev, BM, 32) ||
        make_operand_map(&mp.a_lo, aev_lo, (uint64_t)aev_rows_total, (uint64_t)Lin, (uint64_t)ld_aev, BM, 32) ||
        weight_map(&mp.b_hi[0], &mp.b_lo[0], w.Wf[0], (uint64_t)M * n1, (uint64_t)Lin, (uint32_t)n1) ||
        weight_map(&mp.b_hi[1], &mp.b_lo[1], w.Wf[1], (uint64_t)M * n2, (uint64_t)n1, (uint32_t)n2) ||
        weight_map(&mp.b_hi[2], &mp.b_lo[2], w.Wf[2], (uint64_t)M * n3, (uint64_t)n2, (uint32_t)n3) ||
        weight_map(&mp.b_hi[3], &mp.b_lo[3], w.WT[2], (uint64_t)M * n2, (uint64_t)n3, (uint32_t)n2) ||
        weight_map(&mp.b_hi[4], &mp.b_lo[4], w.WT[1], (uint64_t)M * n1, (uint64_t)n2, (uint32_t)n1))
      return 1;
    if (fuse_d) {
      if (weight_map(&mp.b_hi[5], &mp.b_lo[5], w.Wcat, (uint64_t)Lin, (uint64_t)M * n1, (uint32_t)Npad) ||
          make_map(&mp.out0, daev_chunk, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, (uint64_t)Rc, (uint64_t)Lin, (uint64_t)ld_daev, 32, BM,
                   CU_TENSOR_MAP_SWIZZLE_128B))
        return 1;
      mp.out1 = mp.out0;
    } else {
      mp.b_hi[5] = mp.b_hi[4]; mp.b_lo[5] = mp.b_lo[4];
      if (plane_block_map(&mp.out0, &mp.out1, work.G1, (uint64_t)Rc, (uint64_t)M * n1)) return 1;
    }
    const int items = rtiles * M;
    const int grid = std::max(1, std::min(items, ctas));
    if (opt.prof) {
      ANIB_CUDA_OK(cudaFuncSetAttribute(f16_fused_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
      ANIB_CUDA_OK(cudaMemsetAsync(work.prof.p, 0, (size_t)(ctas * 8 + 64) * sizeof(long long), st));
      f16_fused_kernel<true><<<grid, kFThreads, smem, st>>>(mp, g);
    } else {
      ANIB_CUDA_OK(cudaFuncSetAttribute(f16_fused_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
      f16_fused_kernel<false><<<grid, kFThreads, smem, st>>>(mp, g);
    }
    ANIB_CUDA_OK(cudaGetLastError());
    if (launches) (*launches)++;
    if (opt.prof) {
      std::vector<long long> h((size_t)grid * 8), tr(64);
      ANIB_CUDA_OK(cudaStreamSynchronize(st));
      ANIB_CUDA_OK(cudaMemcpy(h.data(), work.prof.p, h.size() * sizeof(long long), cudaMemcpyDeviceToHost));
      ANIB_CUDA_OK(cudaMemcpy(tr.data(), (long long*)work.prof.p + (size_t)ctas * 8, 64 * sizeof(long long), cudaMemcpyDeviceToHost));
      if (tr[0] != 0) {
        fprintf(stderr, "[fused trace] CTA 0, third item (cycles from its stage-0 start): MMA warp: start / first granule / last granule / all issued | "
                        "epilogue warp 16 (granules 0, 4, ..): accumulator seen / first granule written / last written\n");
        for (int q = 0; q < g.nstages; q++)
          fprintf(stderr, "  S%d (K %3d N %3d): %6lld %6lld %6lld %6lld | %6lld %6lld %6lld\n", q, g.K[q], g.N[q], tr[q * 8] - tr[0], tr[q * 8 + 1] ? tr[q * 8 + 1] - tr[0] : 0,
                  tr[q * 8 + 2] ? tr[q * 8 + 2] - tr[0] : 0, tr[q * 8 + 3] - tr[0], tr[q * 8 + 4] - tr[0], tr[q * 8 + 5] ? tr[q * 8 + 5] - tr[0] : 0, tr[q * 8 + 6] - tr[0]);
      }
      double s[8] = {0, 0, 0, 0, 0, 0, 0, 0};
      for (int c = 0; c < grid; c++)
        for (int k = 0; k < 8; k++) s[k] += (double)h[(size_t)c * 8 + k] / grid;
      const double per = (double)items / grid;
      fprintf(stderr, "[fused prof] K0 %d widths %d-%d-%d rows %d items %d grid %d | per item (cycles): MMA thread total %.0f, waits: acc-free %.0f weights %.0f A-block %.0f, issue+commit %.0f"
              " | epilogue warp 16 total %.0f, waits: acc-full %.0f slot %.0f\n",
              Lin, n1, n2, n3, Rc, items, grid, s[0] / per, s[1] / per, s[2] / per, s[3] / per, s[7] / per, s[4] / per, s[5] / per, s[6] / per);
    }
    if (!fuse_d) {
      // wide AEV (more than 256 active columns): dE/dAEV = (1/M) G1 . Wcat^T with the layer-by-layer kernel (ensemble mean folded into K)
      static int sms = 0, msm = 0;
      if (!sms) { sms = num_sms; msm = max_smem; }
      Launcher L{st, sms, launches, msm};
      const Operand none{nullptr, nullptr, 0, 0, 0};
      Args a{};
      a.BN = (Lin <= 256 && Lin % 32 == 0) ? Lin : 256;
      a.R = Rc; a.N = Lin; a.num_n_tiles = (Lin + a.BN - 1) / a.BN; a.num_row_tiles = rtiles; a.K = M * n1; a.batch = 1;
      a.acc_scale = 1.0f / (kScaleG * w.sw[0] * (float)M);
      Operand A{work.G1.h(), work.G1.l(), (uint64_t)Rc, (uint64_t)M * n1, (uint64_t)M * n1};
      Operand B{w.Wcat.h(), w.Wcat.l(), (uint64_t)Lin, (uint64_t)M * n1, (uint64_t)pad64((size_t)M * n1)};
      if (launch_gemm<EPI_D>(L, A, B, none, daev_chunk, (uint64_t)Rc, (uint64_t)Lin, (uint64_t)ld_daev, none, a)) return 1;
    }
  }
  energy_reduce_kernel<<<(R + 255) / 256, 256, 0, st>>>(work.e.f(), R, M, sae, eatom + row0);
  if (launches) (*launches)++;
  return 0;
}

}  // namespace f16
}  // namespace anib
