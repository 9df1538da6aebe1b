// mlp_f16_chain.cuh -- chained f16x3 GEMMs: consecutive layers of the ensemble MLP inside ONE tcgen05 kernel.
//
// Why: with the AEV compressed to the active species (ani_b200.cu, Engine::set_active) the layer-by-layer kernels of
// mlp_f16.cuh are bound by L2 / HBM traffic, not by the tensor pipe (profiles/: 13-15 TB/s of L2 traffic, 4-5 TB/s of
// DRAM traffic per GEMM launch): every activation is written once and read two or three times as fp16 hi/lo planes.
// Here the epilogue of layer l writes its output tile straight into shared memory in the K-major SWIZZLE_64B layout a
// UMMA A operand needs, 128 rows x 32 columns (hi + lo plane = 16 KB) at a time, and layer l+1's MMAs consume the block
// as soon as it is complete; the same shared-memory block is also the source of the TMA store for the tensors that the
// other direction of the pass still needs (H1, H2 for CELU', G2 for the second kernel).
//
//   forward kernel  (item = row tile x member):  AEV -L1-> h1 -L2-> h2 -L3+out-> g3 -L3^T (x CELU'(z2))-> g2
//        reads AEV, weights; writes H1, H2, G2 planes and the member energies.  h3 / g3 never leave the SM.
//   backward kernel (item = row tile, members inside):  G2 -L2^T (x CELU'(z1))-> g1 -L1^T-> dE/dAEV accumulated over the
//        members in TMEM, one fp32 tile store at the end.  g1 never leaves the SM.
//
// Roles (384 threads): warp 0 TMA producer (A ring for a global A operand, B ring for the weights), warp 1 MMA issuer,
// warp 2 TMEM allocator + TMA-store warp (ships completed ring blocks to global), warp 3 aux loader (the activations a
// CELU' epilogue needs, 128 x 32 blocks), warps 4-11 epilogue (two warps per TMEM lane quadrant, alternate column blocks).
// Replaces BmmEnsemble's bmm chain and its autograd backward (models/lammps_ani.py:219-257,203-214), like mlp_f16.cuh.
#pragma once
#include "mlp_f16.cuh"

namespace anib {
namespace f16 {

constexpr int kChainMax = 4;
constexpr int kEpiGroups = 4;                 // epilogue warps per TMEM lane quadrant
constexpr int kChainThreads = 128 + 128 * kEpiGroups;
constexpr int kNA = 4;                        // ring depth of the global A operand
constexpr int kNH = kEpiGroups;               // output-block slots: epilogue group g owns slot g (block cb lives in slot cb % kNH), so
                                              // every waiter of a slot's mbarriers sees each of its phases (no parity aliasing)
constexpr int kNX = 4;                        // aux loads in flight (they land in the H-ring slot their block will be written to)
constexpr int kBlkBytes = 16384;   // 128 rows x 32 halves x (hi + lo), or 128 rows x 32 floats
constexpr int kMaxRing = 8;

struct ChainStage {
  int K, N;                  // contraction length (multiple of 16), output width (multiple of 32, <= 256)
  int a_ring;                // A operand: 0 = TMA from global (maps.a_*), 1 = the blocks the previous stage wrote to the H ring
  int epi;                   // EPI_F / EPI_FO / EPI_B / EPI_D
  int to_ring, to_global;    // output blocks feed the next stage's MMAs / are TMA-stored
  int tmem_col;              // accumulator column base
  int dep_stage, dep_prev;   // epilogue that must have drained the overlapping TMEM columns (dep_stage < 0: none)
  int acc_over_z;            // accumulate over the inner member loop; epilogue after the last member only
  int aux_wait_pub;          // EPI_B: the aux tensor is stored by an earlier stage of the same item (to_global == 2)
  int a_k_off_per_z, b_row_off_per_z, b_k_off_per_z, out_col_off_per_z, aux_col_off_per_z;
  float acc_scale, out_scale, aux_inv_scale;
  const float* bias; int bias_stride_z;
  const float* w_out; const float* b_out; float* e_out; int e_ld;
};

struct ChainArgs {
  int R, num_row_tiles, batch, inner_z, nstages, a_row0;
  int NB;                    // depth of the weight ring (what is left of shared memory)
  int b_slot_bytes;
  int* dbg;                  // host-mapped word for the wait watchdog (may be null)
  long long* trace;          // optional clock64 stamps of CTA 0's third item (ANI_B200_CHAIN_TRACE=1): [stage][8]
  ChainStage st[kChainMax];
};

struct alignas(64) ChainMaps {
  CUtensorMap a_hi, a_lo;
  CUtensorMap b_hi[kChainMax], b_lo[kChainMax], o_hi[kChainMax], o_lo[kChainMax], x_hi[kChainMax], x_lo[kChainMax];
};

// byte offset of 16-byte chunk ch of row r (0..127) in a block of 64-byte rows under SWIZZLE_64B
__device__ __forceinline__ uint32_t blk_off64(int r, int ch) { return (uint32_t)(r * 64 + ((ch ^ ((r >> 1) & 3)) << 4)); }
// same for 128-byte rows under SWIZZLE_128B
__device__ __forceinline__ uint32_t blk_off128(int r, int ch) { return (uint32_t)(r * 128 + ((ch ^ (r & 7)) << 4)); }

__device__ __forceinline__ bool mbar_test(uint32_t bar, uint32_t parity) {
  uint32_t done;
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
      "selp.u32 %0, 1, 0, p;\n"
      "}\n" : "=r"(done) : "r"(bar), "r"(parity) : "memory");
  return done != 0;
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

__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
        "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

__global__ void __launch_bounds__(kChainThreads, 1)
f16_chain_kernel(const __grid_constant__ ChainMaps maps, const __grid_constant__ ChainArgs g) {
  extern __shared__ unsigned char smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  const uint32_t baseB = base;
  const uint32_t baseA = baseB + (uint32_t)g.NB * g.b_slot_bytes;
  const uint32_t baseH = baseA + (uint32_t)kNA * kBlkBytes;
  const uint32_t bar_base = baseH + (uint32_t)kNH * kBlkBytes;
  auto bar = [&](int group, int i) { return bar_base + 8u * (uint32_t)(group * kMaxRing + i); };
  // groups: 0 bfull, 1 bempty, 2 afull, 3 aempty, 4 hfull, 5 hfree, 6 xfull, 8 accfull[stage], 9 accfree[stage]
  const uint32_t tmem_slot = bar_base + 8u * 10u * kMaxRing;
  volatile uint32_t* tmem_slot_ptr = (volatile uint32_t*)(smem_raw + (tmem_slot - raw));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#define WD(grp_, idx_) (0x80000000u | ((uint32_t)warp << 24) | ((uint32_t)(grp_) << 20) | (((uint32_t)(idx_) & 15u) << 16) | ((uint32_t)__LINE__ & 0xffffu))
  const int nz = g.inner_z ? g.batch : 1;
  const int num_items = g.inner_z ? g.num_row_tiles : g.num_row_tiles * g.batch;

  if (warp == 1 && lane == 0) {
    for (int i = 0; i < g.NB; i++) { mbar_init(bar(0, i), 1); mbar_init(bar(1, i), 1); }
    for (int i = 0; i < kNA; i++) { mbar_init(bar(2, i), 1); mbar_init(bar(3, i), 1); }
    for (int i = 0; i < kNH; i++) { mbar_init(bar(4, i), 4); mbar_init(bar(5, i), 2); }
    for (int i = 0; i < kNX; i++) mbar_init(bar(6, i), 1);
    for (int i = 0; i < kChainMax; i++) { mbar_init(bar(8, i), 1); mbar_init(bar(9, i), 4 * kEpiGroups); }
    mbar_init(bar(8, kChainMax), 1);   // "published": stores of a to_global == 2 stage are visible
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
    // ================================================================= TMA producer: weights (B ring) and a global A operand (A ring)
    if (lane == 0) {
      uint32_t bq = 0, aq = 0;
      for (int it = blockIdx.x; it < num_items; it += gridDim.x) {
        const int rt = g.inner_z ? it : it / g.batch;
        const int z0 = g.inner_z ? 0 : it % g.batch;
        const int a_row = g.a_row0 + rt * BM;
        for (int zi = 0; zi < nz; zi++) {
          const int z = z0 + zi;
          for (int s = 0; s < g.nstages; s++) {
            const ChainStage& S = g.st[s];
            const int KB = (S.K + 31) / 32;
            const uint32_t bplane = (uint32_t)S.N * 64u;
            for (int kb = 0; kb < KB; kb++) {
              const int sb = (int)(bq % (uint32_t)g.NB);
              mbar_wait_wd(bar(1, sb), ((bq / (uint32_t)g.NB) & 1u) ^ 1u, g.dbg, WD(1, sb));
              const uint32_t dst = baseB + (uint32_t)sb * g.b_slot_bytes;
              mbar_expect_tx(bar(0, sb), 2u * bplane);
              tma_load_2d(&maps.b_hi[s], bar(0, sb), dst, z * S.b_k_off_per_z + kb * 32, z * S.b_row_off_per_z);
              tma_load_2d(&maps.b_lo[s], bar(0, sb), dst + bplane, z * S.b_k_off_per_z + kb * 32, z * S.b_row_off_per_z);
              bq++;
              if (!S.a_ring) {
                const int sa = (int)(aq % (uint32_t)kNA);
                mbar_wait_wd(bar(3, sa), ((aq / (uint32_t)kNA) & 1u) ^ 1u, g.dbg, WD(3, sa));
                const uint32_t da = baseA + (uint32_t)sa * kBlkBytes;
                mbar_expect_tx(bar(2, sa), kBlkBytes);
                tma_load_2d(&maps.a_hi, bar(2, sa), da, z * S.a_k_off_per_z + kb * 32, a_row);
                tma_load_2d(&maps.a_lo, bar(2, sa), da + kBlkBytes / 2, z * S.a_k_off_per_z + kb * 32, a_row);
                aq++;
              }
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ================================================================= MMA issuer
    if (lane == 0) {
      uint32_t bq = 0, aq = 0;
      uint32_t huse[kNH] = {0, 0, 0, 0}, pbase[kNH] = {0, 0, 0, 0};   // uses of each H slot so far / before the previous stage
      uint32_t occ[kChainMax] = {0, 0, 0, 0};     // accumulator-complete commits issued so far, per stage
      for (int it = blockIdx.x; it < num_items; it += gridDim.x) {
        for (int zi = 0; zi < nz; zi++) {
          for (int s = 0; s < g.nstages; s++) {
            const ChainStage& S = g.st[s];
            const bool first = !S.acc_over_z || zi == 0, last = !S.acc_over_z || zi == nz - 1;
            if (S.dep_stage >= 0 && first && occ[S.dep_stage] > 0) {   // the columns must have been drained by that epilogue
              mbar_wait_wd(bar(9, S.dep_stage), (occ[S.dep_stage] - 1u) & 1u, g.dbg, WD(9, S.dep_stage));
              tc_fence_after();
            }
            const bool tr = g.trace != nullptr && blockIdx.x == 0 && it == blockIdx.x + 2 * (int)gridDim.x && zi == 0;
            if (tr) g.trace[s * 8 + 0] = clock64();                   // stage start (TMEM columns free)
            const uint32_t tmem_d = tmem_base + (uint32_t)S.tmem_col;
            const uint32_t idesc = umma_idesc_f16(BM, S.N);
            const int KB = (S.K + 31) / 32;
            const uint32_t bplane = (uint32_t)S.N * 64u;
            for (int kb = 0; kb < KB; kb++) {
              const int sb = (int)(bq % (uint32_t)g.NB);
              mbar_wait_wd(bar(0, sb), (bq / (uint32_t)g.NB) & 1u, g.dbg, WD(0, sb));
              if (tr && kb == 0) g.trace[s * 8 + 1] = clock64();       // first weight block there
              if (tr && kb == KB - 1) g.trace[s * 8 + 3] = clock64();  // last weight block there
              uint32_t sA;
              int sa = 0;
              if (S.a_ring) {
                sa = kb % kNH;
                mbar_wait_wd(bar(4, sa), (pbase[sa] + (uint32_t)(kb / kNH)) & 1u, g.dbg, WD(4, sa));
                sA = baseH + (uint32_t)sa * kBlkBytes;
              } else {
                sa = (int)(aq % (uint32_t)kNA);
                mbar_wait_wd(bar(2, sa), (aq / (uint32_t)kNA) & 1u, g.dbg, WD(2, sa));
                sA = baseA + (uint32_t)sa * kBlkBytes;
              }
              tc_fence_after();
              if (tr && kb == 0) g.trace[s * 8 + 2] = clock64();       // first A block there
              if (tr && kb == KB - 1) g.trace[s * 8 + 4] = clock64();  // last A block there (all MMAs issued right after)
              const uint32_t sB = baseB + (uint32_t)sb * g.b_slot_bytes;
              const uint64_t dA_hi = umma_desc_k(sA, 32), dA_lo = umma_desc_k(sA + kBlkBytes / 2, 32);
              const uint64_t dB_hi = umma_desc_k(sB, 32), dB_lo = umma_desc_k(sB + bplane, 32);
              const int nks = min(2, (S.K - kb * 32) / 16);
              for (int k = 0; k < nks; k++) {
                const uint64_t adv = (uint64_t)(2 * k);            // 16 halves = 32 bytes along K, in 16-byte units
                umma_f16(tmem_d, dA_hi + adv, dB_hi + adv, idesc, ((kb | k) != 0 || !first) ? 1u : 0u);
                umma_f16(tmem_d, dA_lo + adv, dB_hi + adv, idesc, 1u);
                umma_f16(tmem_d, dA_hi + adv, dB_lo + adv, idesc, 1u);
              }
              umma_commit(bar(1, sb));
              bq++;
              if (S.a_ring) umma_commit(bar(5, sa));
              else { umma_commit(bar(3, sa)); aq++; }
            }
            if (last) { umma_commit(bar(8, s)); occ[s]++; }
            for (int i = 0; i < kNH; i++) {
              pbase[i] = huse[i];
              if (last) huse[i] += (uint32_t)((S.N / 32 - i + kNH - 1) / kNH);
            }
          }
        }
      }
    }
  } else if (warp == 2) {
    // ================================================================= store warp: completed ring blocks -> global (TMA)
    if (lane == 0) {
      uint32_t huse[kNH] = {0, 0, 0, 0};
      int pend_slot = -1, pend_arrivals = 0;
      for (int it = blockIdx.x; it < num_items; it += gridDim.x) {
        const int rt = g.inner_z ? it : it / g.batch;
        const int z0 = g.inner_z ? 0 : it % g.batch;
        for (int zi = 0; zi < nz; zi++) {
          const int z = z0 + zi;
          for (int s = 0; s < g.nstages; s++) {
            const ChainStage& S = g.st[s];
            if (S.acc_over_z && zi != nz - 1) continue;
            const int nbo = S.N / 32;
            for (int cb = 0; cb < nbo; cb++) {
              const int slot = cb % kNH;
              if (pend_slot >= 0 && !mbar_test(bar(4, slot), huse[slot] & 1u)) {
                // never hold a stored block's release while waiting for the next block: with one slot per epilogue group
                // the next block may need exactly that slot (last block of a stage -> first block of the next)
                tma_store_wait_read<0>();
                for (int a = 0; a < pend_arrivals; a++) mbar_arrive(bar(5, pend_slot));
                pend_slot = -1;
              }
              mbar_wait_wd(bar(4, slot), huse[slot] & 1u, g.dbg, WD(4, slot));
              huse[slot]++;
              if (S.to_global) {
                const uint32_t src = baseH + (uint32_t)slot * kBlkBytes;
                const int col = z * S.out_col_off_per_z + cb * 32;
                tma_store_2d(&maps.o_hi[s], src, col, rt * BM);
                if (S.epi != EPI_D) tma_store_2d(&maps.o_lo[s], src + kBlkBytes / 2, col, rt * BM);
                tma_store_commit();
                if (pend_slot >= 0) {
                  tma_store_wait_read<1>();
                  for (int a = 0; a < pend_arrivals; a++) mbar_arrive(bar(5, pend_slot));
                }
                pend_slot = slot; pend_arrivals = S.to_ring ? 1 : 2;
              } else {
                if (pend_slot >= 0) {           // do not sit on a stored block while no further store is coming
                  tma_store_wait_read<0>();
                  for (int a = 0; a < pend_arrivals; a++) mbar_arrive(bar(5, pend_slot));
                  pend_slot = -1;
                }
                mbar_arrive(bar(5, slot));      // the MMA commit is the other arrival
              }
            }
            if (S.to_global == 2) {
              // a later stage of this item reads these blocks back (aux): make the writes visible, then tell the aux loader
              tma_store_wait_all();
              if (pend_slot >= 0) { for (int a = 0; a < pend_arrivals; a++) mbar_arrive(bar(5, pend_slot)); pend_slot = -1; }
              mbar_arrive(bar(8, kChainMax));   // "published" barrier (group 8, index kChainMax)
            }
          }
        }
      }
      if (pend_slot >= 0) {
        tma_store_wait_read<0>();
        for (int a = 0; a < pend_arrivals; a++) mbar_arrive(bar(5, pend_slot));
      }
      tma_store_wait_all();
    }
  } else if (warp == 3) {
    // idle (the aux loads are issued by the epilogue warps themselves, in step with the ring sequence)
  } else {
    // ================================================================= epilogue warps (warp w reads TMEM lanes 32*(w%4)..)
    // kEpiGroups warps per lane quadrant take alternate 32-column blocks; a block is done as two 16-column halves to keep
    // the register count under the 640-thread budget.
    const int q = warp & 3, grp = (warp - 4) >> 2;
    const int r = q * 32 + lane;                                 // row inside the 128-row tile
    uint32_t huse = 0, xuse = 0;   // blocks / aux blocks this warp's group has put through its slot
    uint32_t eocc[kChainMax] = {0, 0, 0, 0};
    for (int it = blockIdx.x; it < num_items; it += gridDim.x) {
      const int rt = g.inner_z ? it : it / g.batch;
      const int z0 = g.inner_z ? 0 : it % g.batch;
      for (int zi = 0; zi < nz; zi++) {
        const int z = z0 + zi;
#pragma unroll 1
        for (int s = 0; s < g.nstages; s++) {
          const ChainStage& S = g.st[s];
          if (S.acc_over_z && zi != nz - 1) continue;
          const int epi = S.epi, nbo = S.N / 32;
          const float acc_scale = S.acc_scale, out_scale = S.out_scale, aux_inv = S.aux_inv_scale;
          const float neg_a = 0.1f * out_scale;
          const float* bias_z = (epi == EPI_F || epi == EPI_FO) ? S.bias + (size_t)z * S.bias_stride_z : nullptr;
          const float* wo_z = (epi == EPI_FO) ? S.w_out + (size_t)z * S.N : nullptr;
          if (S.aux_wait_pub) mbar_wait_wd(bar(8, kChainMax), eocc[s] & 1u, g.dbg, WD(8, kChainMax));   // the aux tensor was stored by an earlier stage of this item
          auto issue_aux = [&](int cb) {
            // aux block (activations for CELU') straight into the slot this block will be written to: the epilogue works in
            // place.  Issued by the group itself, in step with its slot's phases (a free-running loader would alias them).
            mbar_wait_wd(bar(5, grp), (huse & 1u) ^ 1u, g.dbg, WD(5, grp));
            const uint32_t xbar = bar(6, grp), dst = baseH + (uint32_t)grp * kBlkBytes;
            mbar_expect_tx(xbar, kBlkBytes);
            tma_load_2d(&maps.x_hi[s], xbar, dst, z * S.aux_col_off_per_z + cb * 32, rt * BM);
            tma_load_2d(&maps.x_lo[s], xbar, dst + kBlkBytes / 2, z * S.aux_col_off_per_z + cb * 32, rt * BM);
          };
          // the first block's aux load does not need the accumulator: start it while this stage's MMAs are still running
          if (epi == EPI_B && grp < nbo && q == 0 && lane == 0) issue_aux(grp);
          mbar_wait_wd(bar(8, s), eocc[s] & 1u, g.dbg, WD(8, s));
          eocc[s]++;
          tc_fence_after();
          const bool tr = g.trace != nullptr && blockIdx.x == 0 && it == blockIdx.x + 2 * (int)gridDim.x && zi == nz - 1 && warp == 4 && lane == 0;
          if (tr) g.trace[s * 8 + 5] = clock64();                     // accumulator complete seen by the epilogue
          const uint32_t taddr0 = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)S.tmem_col;
          float e_acc = 0.f;
#pragma unroll 1
          for (int cb = grp; cb < nbo; cb += kEpiGroups) {
            const int slot = grp;
            const uint32_t blk = baseH + (uint32_t)slot * kBlkBytes;
            const int xslot = grp;
            const uint32_t xb = blk;             // the aux block was loaded into this very slot
            if (epi == EPI_B && cb != grp && q == 0 && lane == 0) issue_aux(cb);
#pragma unroll 1
            for (int sub = 0; sub < 2; sub++) {
              uint32_t v[16];
              tmem_ld16(taddr0 + (uint32_t)(cb * 32 + sub * 16), v);
              const int n0 = cb * 32 + sub * 16;
              float o[16];
              if (epi == EPI_F || epi == EPI_FO) {
                float b[16], w[16];
#pragma unroll
                for (int k = 0; k < 4; k++) {
                  const float4 t = __ldg((const float4*)(bias_z + n0) + k);
                  b[4 * k] = t.x; b[4 * k + 1] = t.y; b[4 * k + 2] = t.z; b[4 * k + 3] = t.w;
                }
                if (epi == EPI_FO) {
#pragma unroll
                  for (int k = 0; k < 4; k++) {
                    const float4 t = __ldg((const float4*)(wo_z + n0) + k);
                    w[4 * k] = t.x; w[4 * k + 1] = t.y; w[4 * k + 2] = t.z; w[4 * k + 3] = t.w;
                  }
#pragma unroll
                  for (int c = 0; c < 16; c++) {
                    const float zz = fmaf(__uint_as_float(v[c]), acc_scale, b[c]);
                    const float ex = exp2f(zz * 14.426950408889634f);  // e^(z/alpha), alpha = 0.1
                    const float h = zz > 0.f ? zz : fmaf(0.1f, ex, -0.1f);
                    e_acc = fmaf(h, w[c], e_acc);
                    o[c] = w[c] * out_scale * (zz > 0.f ? 1.f : ex);
                  }
                } else {
#pragma unroll
                  for (int c = 0; c < 16; c++) {
                    const float zz = fmaf(__uint_as_float(v[c]), acc_scale, b[c]);
                    const float ex = exp2f(zz * 14.426950408889634f);
                    o[c] = zz > 0.f ? zz * out_scale : fmaf(neg_a, ex, -neg_a);
                  }
                }
              } else if (epi == EPI_B) {
                if (sub == 0) { mbar_wait_wd(bar(6, xslot), xuse & 1u, g.dbg, WD(6, xslot)); xuse++; }
#pragma unroll
                for (int c2 = 0; c2 < 2; c2++) {
                  uint32_t hh[4], ll[4];
                  const uint32_t off = blk_off64(r, 2 * sub + c2);
                  lds128(xb + off, hh[0], hh[1], hh[2], hh[3]);
                  lds128(xb + kBlkBytes / 2 + off, ll[0], ll[1], ll[2], ll[3]);
#pragma unroll
                  for (int e = 0; e < 4; e++) {
                    const float2 fh = __half22float2(*reinterpret_cast<const __half2*>(&hh[e]));
                    const float2 fl = __half22float2(*reinterpret_cast<const __half2*>(&ll[e]));
                    const float h0 = (fh.x + fl.x) * aux_inv, h1 = (fh.y + fl.y) * aux_inv;
                    const float d0 = h0 > 0.f ? 1.f : fmaf(10.f, h0, 1.f);   // CELU'(z) from h = CELU(z)
                    const float d1 = h1 > 0.f ? 1.f : fmaf(10.f, h1, 1.f);
                    o[c2 * 8 + 2 * e] = __uint_as_float(v[c2 * 8 + 2 * e]) * acc_scale * d0;
                    o[c2 * 8 + 2 * e + 1] = __uint_as_float(v[c2 * 8 + 2 * e + 1]) * acc_scale * d1;
                  }
                }
              } else {
#pragma unroll
                for (int c = 0; c < 16; c++) o[c] = __uint_as_float(v[c]) * acc_scale;
              }
              // ---- registers -> ring block (UMMA A-operand layout == TMA-store source layout)
              if (sub == 0 && epi != EPI_B) mbar_wait_wd(bar(5, slot), (huse & 1u) ^ 1u, g.dbg, WD(5, slot));   // EPI_B: the aux issue above waited
              if (epi == EPI_D) {
#pragma unroll
                for (int c4 = 0; c4 < 4; c4++)
                  sts128(blk + blk_off128(r, 4 * sub + c4), __float_as_uint(o[4 * c4]), __float_as_uint(o[4 * c4 + 1]),
                         __float_as_uint(o[4 * c4 + 2]), __float_as_uint(o[4 * c4 + 3]));
              } else {
#pragma unroll
                for (int c2 = 0; c2 < 2; c2++) {
                  uint32_t hh[4], ll[4];
#pragma unroll
                  for (int e = 0; e < 4; e++) {
                    const float x0 = o[c2 * 8 + 2 * e], x1 = o[c2 * 8 + 2 * e + 1];
                    const __half2 h2 = __floats2half2_rn(x0, x1);
                    const float2 hf = __half22float2(h2);
                    const __half2 l2 = __floats2half2_rn(x0 - hf.x, x1 - hf.y);
                    hh[e] = *reinterpret_cast<const uint32_t*>(&h2);
                    ll[e] = *reinterpret_cast<const uint32_t*>(&l2);
                  }
                  const uint32_t off = blk_off64(r, 2 * sub + c2);
                  sts128(blk + off, hh[0], hh[1], hh[2], hh[3]);
                  sts128(blk + kBlkBytes / 2 + off, ll[0], ll[1], ll[2], ll[3]);
                }
              }
            }
            fence_async_smem();
            __syncwarp();
            if (lane == 0) mbar_arrive(bar(4, slot));
            huse++;
          }
          if (tr) g.trace[s * 8 + 6] = clock64();                     // this warp's blocks of the stage written
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(bar(9, s));
          if (epi == EPI_FO) {                   // e_out is zeroed before the launch; the warps of a quadrant add their parts
            const int row = rt * BM + r;
            if (row < g.R) atomicAdd(S.e_out + (size_t)row * S.e_ld + z, e_acc + (grp == 0 ? __ldg(S.b_out + z) : 0.f));
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
  }
}

#undef WD

// ------------------------------------------------------------------------------------------------ host side
inline bool chain_supported(const std::vector<int>& dims) {
  return supported_dims(dims) && (int)dims.size() == 5;   // three hidden layers (every ANI-1x/2x/1xnr network)
}
inline bool chain_bwd_supported(const std::vector<int>& dims) { return chain_supported(dims) && (dims[0] + 31) / 32 * 32 <= 256; }

// TMEM dependency of every stage: the latest earlier epilogue (in issue order, wrapping to the previous item / member)
// whose accumulator columns overlap this stage's.
inline void chain_fill_deps(ChainArgs& g) {
  auto overlap = [&](int a, int b) {
    const int a0 = g.st[a].tmem_col, a1 = a0 + g.st[a].N, b0 = g.st[b].tmem_col, b1 = b0 + g.st[b].N;
    return a0 < b1 && b0 < a1;
  };
  for (int s = 0; s < g.nstages; s++) {
    g.st[s].dep_stage = -1; g.st[s].dep_prev = 0;
    for (int k = 1; k <= g.nstages; k++) {
      const int d = ((s - k) % g.nstages + g.nstages) % g.nstages;
      if (overlap(s, d)) { g.st[s].dep_stage = d; g.st[s].dep_prev = (s - k) < 0; break; }
    }
  }
}

inline int* chain_dbg_word(bool device_ptr) {
  static int* h = nullptr; static int* d = nullptr;
  if (!h) {
    if (cudaHostAlloc((void**)&h, 128, cudaHostAllocMapped) != cudaSuccess) { h = nullptr; cudaGetLastError(); return nullptr; }
    for (int i = 0; i < 32; i++) h[i] = 0;
    if (cudaHostGetDevicePointer((void**)&d, h, 0) != cudaSuccess) { d = nullptr; cudaGetLastError(); }
  }
  return device_ptr ? d : h;
}

inline int launch_chain(const Launcher& L, const ChainMaps& maps, ChainArgs g) {
  g.dbg = chain_dbg_word(true);
  static const int trace_on = env_int("ANI_B200_CHAIN_TRACE", 0);
  static long long* trace_dev = nullptr;
  static int traced = 0;
  g.trace = nullptr;
  if (trace_on && traced < 4) {
    if (!trace_dev) cudaMalloc((void**)&trace_dev, 64 * sizeof(long long));
    cudaMemsetAsync(trace_dev, 0, 64 * sizeof(long long), L.st);
    g.trace = trace_dev;
  }
  int maxN = 0;
  for (int s = 0; s < g.nstages; s++) maxN = std::max(maxN, g.st[s].N);
  g.b_slot_bytes = 2 * maxN * 64;
  const int fixed = 1024 + (kNA + kNH) * kBlkBytes + 8 * 10 * kMaxRing + 64;
  g.NB = std::min((L.max_smem - fixed) / g.b_slot_bytes, kMaxRing);
  if (g.NB < 2) { set_error("not enough shared memory for the chained f16 GEMM kernel"); return 1; }
  const int smem = fixed + g.NB * g.b_slot_bytes;
  chain_fill_deps(g);
  ANIB_CUDA_OK(cudaFuncSetAttribute(f16_chain_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  const int items = g.inner_z ? g.num_row_tiles : g.num_row_tiles * g.batch;
  const int grid = std::max(1, std::min(items, L.num_sms));
  f16_chain_kernel<<<grid, kChainThreads, smem, L.st>>>(maps, g);
  ANIB_CUDA_OK(cudaGetLastError());
  if (g.trace) {
    long long h[64];
    cudaStreamSynchronize(L.st);
    cudaMemcpy(h, trace_dev, sizeof(h), cudaMemcpyDeviceToHost);
    long long t0 = h[0];
    fprintf(stderr, "[chain trace] launch %d: %d stages, inner_z %d, items %d (cycles relative to stage 0 start of CTA 0's third item)\n", traced, g.nstages, g.inner_z, items);
    for (int s = 0; s < g.nstages; s++)
      fprintf(stderr, "  stage %d (K %d N %d): start %lld  B0 %lld  A0 %lld  Blast %lld  Alast %lld | epi: acc seen %lld  done %lld\n", s, g.st[s].K, g.st[s].N,
              h[s * 8] - t0, h[s * 8 + 1] - t0, h[s * 8 + 2] - t0, h[s * 8 + 3] - t0, h[s * 8 + 4] - t0, h[s * 8 + 5] - t0, h[s * 8 + 6] - t0);
    traced++;
  }
  if (L.launches) (*L.launches)++;
  return 0;
}

inline int plane_block_map(CUtensorMap* hi, CUtensorMap* lo, const HalfPlanes& p, uint64_t rows, uint64_t cols) {
  return make_map(hi, p.h(), CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, rows, cols, cols, 32, BM, CU_TENSOR_MAP_SWIZZLE_64B) ||
         make_map(lo, p.l(), CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, rows, cols, cols, 32, BM, CU_TENSOR_MAP_SWIZZLE_64B);
}
inline int weight_map(CUtensorMap* hi, CUtensorMap* lo, const HalfPlanes& p, uint64_t rows, uint64_t cols, uint32_t box_rows) {
  return make_operand_map(hi, p.h(), rows, cols, pad64(cols), box_rows, 32) ||
         make_operand_map(lo, p.l(), rows, cols, pad64(cols), box_rows, 32);
}

// Forward + backward of one species bucket with the chained kernels.  Same contract as mlp_species().
inline int mlp_species_chain(cudaStream_t st, SpeciesWeights& w, Work& work, int M, const std::vector<int>& dims, const __half* aev_hi,
                             const __half* aev_lo, int64_t ld_aev, int64_t aev_rows_total, int row0, int R, float* daev,
                             int64_t ld_daev, double* eatom, double sae, int* launches) {
  if (!w.ready || !chain_supported(dims)) { set_error("the chained f16 MLP needs three hidden layers with widths that are multiples of 32, <= 256"); return 1; }
  static int num_sms = 0, max_smem = 0;
  if (!num_sms) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, dev);
    cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
  }
  Launcher L{st, num_sms, launches, max_smem};
  static const int chunk = (std::max(env_int("ANI_B200_CHAIN_CHUNK", 65536), BM) + BM - 1) / BM * BM;
  const int alloc_rows = std::min(chunk, (R + BM - 1) / BM * BM);
  const int Lin = dims[0], n1 = dims[1], n2 = dims[2], n3 = dims[3];
  static const int bwd_chain = env_int("ANI_B200_CHAIN_BWD", 1);
  const bool fused_bwd = bwd_chain && chain_bwd_supported(dims);
  if ((int)work.H.size() < 4) { work.H.resize(4); work.G.resize(4); }
  if (!work.H[1].reserve((size_t)alloc_rows * M * n1) || !work.H[2].reserve((size_t)alloc_rows * M * n2) ||
      !work.G[2].reserve((size_t)alloc_rows * M * n2) || (!fused_bwd && !work.G[1].reserve((size_t)alloc_rows * M * n1)) ||
      !work.e.reserve((size_t)alloc_rows * M * 4)) {
    set_error("out of memory for MLP work buffers");
    return 1;
  }
  for (int r0 = 0; r0 < R; r0 += chunk) {
    const int Rc = std::min(chunk, R - r0);
    const int rtiles = (Rc + BM - 1) / BM;
    // ================= forward chain: AEV -> h1 -> h2 -> (h3, e, g3) -> g2
    {
      ChainMaps mp;
      memset(&mp, 0, sizeof(mp));
      ChainArgs g{};
      g.R = Rc; g.num_row_tiles = rtiles; g.batch = M; g.inner_z = 0; g.nstages = 4; g.a_row0 = row0 + r0;
      if (make_operand_map(&mp.a_hi, aev_hi, (uint64_t)aev_rows_total, (uint64_t)Lin, (uint64_t)ld_aev, BM, 32) ||
          make_operand_map(&mp.a_lo, aev_lo, (uint64_t)aev_rows_total, (uint64_t)Lin, (uint64_t)ld_aev, BM, 32))
        return 1;
      ChainStage& s0 = g.st[0];
      s0.K = Lin; s0.N = n1; s0.a_ring = 0; s0.epi = EPI_F; s0.to_ring = 1; s0.to_global = 1; s0.tmem_col = 0;
      s0.b_row_off_per_z = n1; s0.acc_scale = 1.0f / (kScaleAev * w.sw[0]); s0.out_scale = kScaleH;
      s0.bias = w.bias[0].f(); s0.bias_stride_z = n1; s0.out_col_off_per_z = n1;
      if (weight_map(&mp.b_hi[0], &mp.b_lo[0], w.Wf[0], (uint64_t)M * n1, (uint64_t)Lin, (uint32_t)n1) ||
          plane_block_map(&mp.o_hi[0], &mp.o_lo[0], work.H[1], (uint64_t)Rc, (uint64_t)M * n1))
        return 1;
      ChainStage& s1 = g.st[1];
      s1.K = n1; s1.N = n2; s1.a_ring = 1; s1.epi = EPI_F; s1.to_ring = 1; s1.to_global = 2; s1.tmem_col = 256;
      s1.b_row_off_per_z = n2; s1.acc_scale = 1.0f / (kScaleH * w.sw[1]); s1.out_scale = kScaleH;
      s1.bias = w.bias[1].f(); s1.bias_stride_z = n2; s1.out_col_off_per_z = n2;
      if (weight_map(&mp.b_hi[1], &mp.b_lo[1], w.Wf[1], (uint64_t)M * n2, (uint64_t)n1, (uint32_t)n2) ||
          plane_block_map(&mp.o_hi[1], &mp.o_lo[1], work.H[2], (uint64_t)Rc, (uint64_t)M * n2))
        return 1;
      ChainStage& s2 = g.st[2];
      s2.K = n2; s2.N = n3; s2.a_ring = 1; s2.epi = EPI_FO; s2.to_ring = 1; s2.to_global = 0; s2.tmem_col = 0;
      s2.b_row_off_per_z = n3; s2.acc_scale = 1.0f / (kScaleH * w.sw[2]); s2.out_scale = kScaleG;
      s2.bias = w.bias[2].f(); s2.bias_stride_z = n3;
      s2.w_out = w.w_out.f(); s2.b_out = w.b_out.f(); s2.e_out = work.e.f(); s2.e_ld = M;
      if (weight_map(&mp.b_hi[2], &mp.b_lo[2], w.Wf[2], (uint64_t)M * n3, (uint64_t)n2, (uint32_t)n3)) return 1;
      ChainStage& s3 = g.st[3];
      s3.K = n3; s3.N = n2; s3.a_ring = 1; s3.epi = EPI_B; s3.to_ring = 0; s3.to_global = 1; s3.tmem_col = 256;
      s3.b_row_off_per_z = n2; s3.acc_scale = 1.0f / w.sw[2]; s3.aux_inv_scale = 1.0f / kScaleH; s3.aux_wait_pub = 1;
      s3.out_col_off_per_z = n2; s3.aux_col_off_per_z = n2;
      if (weight_map(&mp.b_hi[3], &mp.b_lo[3], w.WT[2], (uint64_t)M * n2, (uint64_t)n3, (uint32_t)n2) ||
          plane_block_map(&mp.o_hi[3], &mp.o_lo[3], work.G[2], (uint64_t)Rc, (uint64_t)M * n2) ||
          plane_block_map(&mp.x_hi[3], &mp.x_lo[3], work.H[2], (uint64_t)Rc, (uint64_t)M * n2))
        return 1;
      ANIB_CUDA_OK(cudaMemsetAsync(work.e.p, 0, (size_t)Rc * M * 4, st));
      if (launch_chain(L, mp, g)) return 1;
    }
    energy_reduce_kernel<<<(Rc + 255) / 256, 256, 0, st>>>(work.e.f(), Rc, M, sae, eatom + row0 + r0);
    if (launches) (*launches)++;
    float* daev_chunk = daev + (size_t)(row0 + r0) * ld_daev;
    if (fused_bwd) {
      // ================= backward chain: per row tile, members inside: G2 -> g1 -> dE/dAEV (accumulated in TMEM)
      ChainMaps mp;
      memset(&mp, 0, sizeof(mp));
      ChainArgs g{};
      g.R = Rc; g.num_row_tiles = rtiles; g.batch = M; g.inner_z = 1; g.nstages = 2; g.a_row0 = 0;
      if (plane_block_map(&mp.a_hi, &mp.a_lo, work.G[2], (uint64_t)Rc, (uint64_t)M * n2)) return 1;
      ChainStage& s0 = g.st[0];
      s0.K = n2; s0.N = n1; s0.a_ring = 0; s0.epi = EPI_B; s0.to_ring = 1; s0.to_global = 0; s0.tmem_col = 0;
      s0.a_k_off_per_z = n2; s0.b_row_off_per_z = n1; s0.acc_scale = 1.0f / w.sw[1]; s0.aux_inv_scale = 1.0f / kScaleH;
      s0.aux_col_off_per_z = n1;
      if (weight_map(&mp.b_hi[0], &mp.b_lo[0], w.WT[1], (uint64_t)M * n1, (uint64_t)n2, (uint32_t)n1) ||
          plane_block_map(&mp.x_hi[0], &mp.x_lo[0], work.H[1], (uint64_t)Rc, (uint64_t)M * n1))
        return 1;
      const int Npad = (Lin + 31) / 32 * 32;
      ChainStage& s1 = g.st[1];
      s1.K = n1; s1.N = Npad; s1.a_ring = 1; s1.epi = EPI_D; s1.to_ring = 0; s1.to_global = 1; s1.tmem_col = 256;
      s1.acc_over_z = 1; s1.b_k_off_per_z = n1; s1.acc_scale = 1.0f / (kScaleG * w.sw[0] * (float)M);
      if (weight_map(&mp.b_hi[1], &mp.b_lo[1], w.Wcat, (uint64_t)Lin, (uint64_t)M * n1, (uint32_t)Npad) ||
          make_map(&mp.o_hi[1], daev_chunk, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, (uint64_t)Rc, (uint64_t)Lin, (uint64_t)ld_daev, 32, BM,
                   CU_TENSOR_MAP_SWIZZLE_128B))
        return 1;
      if (launch_chain(L, mp, g)) return 1;
    } else {
      // wide AEV (more than 256 active columns): G1 and dE/dAEV with the layer-by-layer kernels
      const Operand none{nullptr, nullptr, 0, 0, 0};
      {
        Args a{};
        a.BN = (n1 <= 192) ? n1 : 128;
        a.R = Rc; a.N = n1; a.num_n_tiles = (n1 + a.BN - 1) / a.BN; a.num_row_tiles = rtiles; a.K = n2; a.batch = M;
        a.a_row0 = 0; a.a_k_off_per_z = n2; a.b_row_off_per_z = n1; a.acc_scale = 1.0f / w.sw[1]; a.aux_inv_scale = 1.0f / kScaleH;
        a.out_col_off_per_z = n1; a.aux_col_off_per_z = n1;
        Operand A{work.G[2].h(), work.G[2].l(), (uint64_t)Rc, (uint64_t)M * n2, (uint64_t)M * n2};
        Operand B{w.WT[1].h(), w.WT[1].l(), (uint64_t)M * n1, (uint64_t)n2, (uint64_t)pad64(n2)};
        Operand O{work.G[1].h(), work.G[1].l(), (uint64_t)Rc, (uint64_t)M * n1, (uint64_t)M * n1};
        Operand X{work.H[1].h(), work.H[1].l(), (uint64_t)Rc, (uint64_t)M * n1, (uint64_t)M * n1};
        if (launch_gemm<EPI_B>(L, A, B, O, nullptr, 0, 0, 0, X, a)) return 1;
      }
      {
        Args a{};
        a.BN = (Lin <= 256 && Lin % 32 == 0) ? Lin : 256;
        a.R = Rc; a.N = Lin; a.num_n_tiles = (Lin + a.BN - 1) / a.BN; a.num_row_tiles = rtiles; a.K = M * n1; a.batch = 1;
        a.acc_scale = 1.0f / (kScaleG * w.sw[0] * (float)M);
        Operand A{work.G[1].h(), work.G[1].l(), (uint64_t)Rc, (uint64_t)M * n1, (uint64_t)M * n1};
        Operand B{w.Wcat.h(), w.Wcat.l(), (uint64_t)Lin, (uint64_t)M * n1, (uint64_t)pad64((size_t)M * n1)};
        if (launch_gemm<EPI_D>(L, A, B, none, daev_chunk, (uint64_t)Rc, (uint64_t)Lin, (uint64_t)ld_daev, none, a)) return 1;
      }
    }
  }
  return 0;
}

}  // namespace f16
}  // namespace anib
