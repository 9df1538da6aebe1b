// ani_b200.cu -- context, per-timestep orchestration and the C ABI (include/ani_b200.h) of libani_b200.so.
//
// Replaces the reference's libtorch runner `class ANI` (src/ani_csrc/ani.{h,cpp}) and the TorchScript step
// graph `LammpsANI.forward` (models/lammps_ani.py:131-216).  One context = one GPU, one stream, weights and all
// work buffers resident in HBM; per step: pack coordinates -> AEV forward -> per-species ensemble MLP forward
// + dE/dAEV backward -> AEV backward / force + virial scatter -> energy reduction.
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <mutex>

#include "../../include/ani_b200.h"
#include "aev.cuh"
#include "aev2.cuh"
#include "common.cuh"
#include "misc.cuh"
#include "mlp_simt.cuh"
#include "mlp_f16.cuh"
#include "mlp_f16_fused.cuh"
#include "mlp_tc.cuh"
#include "nbuild.cuh"
#include "md.cuh"
#include "repulsion.cuh"
#include "halo.cuh"

namespace anib {

static thread_local std::string g_last_error;
static int* (*g_dbg_word_fn)(bool) = nullptr;   // set below: the chained-GEMM wait watchdog's host-mapped word
void set_error(const std::string& msg) {
  g_last_error = msg;
  if (g_dbg_word_fn) {
    int* w = g_dbg_word_fn(false);
    for (int i = 0; w && i < 32; i++)
      if (w[i] != 0) {
        char buf[96];
        const unsigned c = (unsigned)w[i];
        snprintf(buf, sizeof(buf), " [wd warp %u grp %u idx %u line %u]", (c >> 24) & 0x7f, (c >> 20) & 15, (c >> 16) & 15, c & 0xffff);
        g_last_error += buf;
      }
  }
}

struct DevBuf {
  void* p = nullptr;
  size_t cap = 0;
  ~DevBuf() { if (p) cudaFree(p); }
  // grow-only; returns false on allocation failure
  bool reserve(size_t bytes) {
    if (bytes <= cap) return true;
    if (p) { cudaFree(p); p = nullptr; cap = 0; }
    size_t want = bytes + bytes / 8 + 256;
    if (cudaMalloc(&p, want) != cudaSuccess) { p = nullptr; return false; }
    cap = want;
    return true;
  }
  template <typename U> U* as() const { return (U*)p; }
};

struct PinnedBuf {
  void* p = nullptr;
  size_t cap = 0;
  ~PinnedBuf() { if (p) cudaFreeHost(p); }
  bool reserve(size_t bytes) {
    if (bytes <= cap) return true;
    if (p) { cudaFreeHost(p); p = nullptr; cap = 0; }
    size_t want = bytes + bytes / 8 + 256;
    if (cudaMallocHost(&p, want) != cudaSuccess) { p = nullptr; return false; }
    cap = want;
    return true;
  }
  template <typename U> U* as() const { return (U*)p; }
};

#define ANIB_RESERVE(buf, bytes)                                                                              \
  do {                                                                                                        \
    if (!(buf).reserve(bytes)) {                                                                              \
      set_error(std::string("out of memory reserving ") + std::to_string((size_t)(bytes)) + " bytes for " #buf); \
      return 1;                                                                                               \
    }                                                                                                         \
  } while (0)

// OPT-IN (option "pin_user" 1; the pair style and bench.py set it): a caller buffer (LAMMPS's x / the pair style's force
// vector) that was seen at the same address on consecutive calls is page-locked in place (cudaHostRegister) so the H2D / D2H
// copies go straight from / to it: no staging memcpy, no host conversion loop.  The caller then owes the library an
// ani_ctx_release_host_buffers() before it frees or reallocates such a buffer (a new allocation at the same address would
// otherwise inherit a stale registration).  Off by default: generic callers get the pinned-staging path, which has no such
// contract.
struct UserPin {
  const void* p = nullptr; size_t bytes = 0; bool registered = false, foreign = false; int seen = 0;
  bool ensure(const void* q, size_t n, bool enabled) {
    if (!enabled) { if (p) release(); return false; }
    if (q == nullptr || n == 0) return false;
    if (q == p && (registered || foreign) && n <= bytes) return true;   // ghost counts shrink and grow between rebuilds
    if (q == p && n == bytes) {
      if (++seen < 2) return false;
      const cudaError_t rc = cudaHostRegister((void*)q, n, cudaHostRegisterDefault);
      registered = rc == cudaSuccess;
      if (rc == cudaErrorHostMemoryAlreadyRegistered) { cudaGetLastError(); foreign = true; return true; }   // the caller pinned it: DMA as is
      if (!registered) { cudaGetLastError(); seen = -1000000; }   // do not retry every step
      return registered;
    }
    release();
    p = q; bytes = n; seen = 1;
    return false;
  }
  void release() {
    if (registered) { cudaHostUnregister((void*)p); cudaGetLastError(); }
    registered = false; foreign = false; p = nullptr; bytes = 0; seen = 0;
  }
  ~UserPin() { release(); }
};

// ---- ghost exchange pieces (ani_halo_forward / ani_halo_reverse, and the overlapped exchange inside Engine::compute), on a given stream
static int halo_forward_on(HaloPlan& p, double* x_dev, cudaStream_t st) {
  if (p.nsend > 0) {
    halo_pack_kernel<double><<<(unsigned)((p.nsend * 3 + 255) / 256), 256, 0, st>>>(x_dev, p.d_send_index, p.d_send_shift, p.nsend, p.d_send_buf);
    ANIB_CUDA_OK(cudaGetLastError());
  }
  double* ghosts = x_dev + (size_t)p.nlocal * 3;
  const int64_t self = p.send_off[p.rank + 1] - p.send_off[p.rank];
  if (self > 0)       // periodic self images: a device copy, no communication
    ANIB_CUDA_OK(cudaMemcpyAsync(ghosts + p.recv_off[p.rank] * 3, p.d_send_buf + p.send_off[p.rank] * 3, (size_t)self * 3 * sizeof(double),
                                 cudaMemcpyDeviceToDevice, st));
  if (p.nranks > 1) {
    NcclApi& n = nccl_api();
    ANIB_NCCL_OK(n.GroupStart());
    for (int r = 0; r < p.nranks; r++) {
      if (r == p.rank) continue;
      const int64_t ns = p.send_off[r + 1] - p.send_off[r], nr = p.recv_off[r + 1] - p.recv_off[r];
      if (ns > 0) ANIB_NCCL_OK(n.Send(p.d_send_buf + p.send_off[r] * 3, (size_t)ns * 3, ncclDouble, r, p.comm, st));
      if (nr > 0) ANIB_NCCL_OK(n.Recv(ghosts + p.recv_off[r] * 3, (size_t)nr * 3, ncclDouble, r, p.comm, st));
    }
    ANIB_NCCL_OK(n.GroupEnd());
  }
  return 0;
}
// ghost forces back to their owners' send buffers ...
static int halo_reverse_exchange_on(HaloPlan& p, double* f_dev, cudaStream_t st) {
  double* ghosts = f_dev + (size_t)p.nlocal * 3;
  const int64_t self = p.send_off[p.rank + 1] - p.send_off[p.rank];
  if (self > 0)
    ANIB_CUDA_OK(cudaMemcpyAsync(p.d_send_buf + p.send_off[p.rank] * 3, ghosts + p.recv_off[p.rank] * 3, (size_t)self * 3 * sizeof(double),
                                 cudaMemcpyDeviceToDevice, st));
  if (p.nranks > 1) {
    NcclApi& n = nccl_api();
    ANIB_NCCL_OK(n.GroupStart());
    for (int r = 0; r < p.nranks; r++) {
      if (r == p.rank) continue;
      const int64_t ns = p.send_off[r + 1] - p.send_off[r], nr = p.recv_off[r + 1] - p.recv_off[r];
      if (nr > 0) ANIB_NCCL_OK(n.Send(ghosts + p.recv_off[r] * 3, (size_t)nr * 3, ncclDouble, r, p.comm, st));
      if (ns > 0) ANIB_NCCL_OK(n.Recv(p.d_send_buf + p.send_off[r] * 3, (size_t)ns * 3, ncclDouble, r, p.comm, st));
    }
    ANIB_NCCL_OK(n.GroupEnd());
  }
  return 0;
}
// ... and f[owner] += what came back (newton off: the pair style folds its own ghost forces, pair_ani.cpp:197-201)
static int halo_reverse_unpack_on(HaloPlan& p, double* f_dev, cudaStream_t st) {
  if (p.nsend > 0) {
    halo_unpack_add_kernel<double><<<(unsigned)((p.nsend * 3 + 255) / 256), 256, 0, st>>>(f_dev, p.d_send_index, p.d_send_buf, p.nsend);
    ANIB_CUDA_OK(cudaGetLastError());
  }
  return 0;
}

struct EngineBase {
  virtual ~EngineBase() {}
  // csr_dev_maxnn >= 0: beg_h / cnt_h are DEVICE arrays (already the engine's own d_beg / d_cnt or to be copied device to device)
  // and the value is the longest row; < 0: host arrays
  virtual int set_neighbors(int nlocal, int ntotal, const int* species_h, const int64_t* beg_h, const int* cnt_h,
                            const int* jlist, int64_t npairs, bool jlist_on_device, int csr_dev_maxnn = -1,
                            bool bucketed_on_device = false) = 0;
  virtual int compute(const void* x, bool x_on_device, bool x_is_float, double* out_energy, double* out_force_host,
                      void* f_dev, bool f_is_float, double* out_eatom_host, double* eatom_dev, double* out_virial9,
                      HaloPlan* halo = nullptr) = 0;
  virtual int get_aev(double* aev, double* daev) = 0;
  virtual int build_neighbors(int nlocal, int ntotal, const int* species_h, const double* x, bool x_on_device, double cutoff,
                              const double* lo, const double* hi, int64_t* npairs_out) = 0;
  virtual int get_neighbors(int* numneigh, int* jlist) = 0;
  virtual void release_pins() = 0;
  virtual int check_errors() = 0;
};

template <typename T>
struct SpeciesNet {
  int nl = 0;                 // number of linear layers (hidden + output)
  std::vector<int> dims;      // in, h1, ..., 1
  std::vector<DevBuf> W, b;   // [l]: [M][out][in], [M][out]   (l < nl-1: hidden; l == nl-1: output row w4[M][h], b4[M])
  std::vector<DevBuf> WT;     // [l] (1 <= l < nl-1): [M][in][out]  (operand of the backward GEMMs)
  DevBuf W1cat;               // [L][M*h1]: W1cat[n][m*h1+k] = W1_m[k][n]
  TcSpeciesWeights tc;        // TF32 hi/lo planes for the tcgen05 path (fp32 only)
  f16::SpeciesWeights hw;     // scaled fp16 hi/lo planes for the f16x3 tcgen05 path (fp32 only)
};

}  // namespace anib

using namespace anib;

struct ani_model {
  HostModel hm;
};

struct ani_halo_plan {
  anib::HaloPlan p;
  ani_ctx* ctx = nullptr;
};

struct ani_ctx {
  int device = 0;
  bool use_single = true, use_cuaev = true, use_fullnbr = true;
  int num_models = 0;
  int mlp_mode = ANI_MLP_FP32_SIMT;
  cudaStream_t stream = nullptr, own_stream = nullptr;
  bool timing = false;
  // per-context switches (ani_ctx_set_option; the environment only provides their defaults at creation)
  int opt_mlp_fused = 1, opt_fused_prof = 0, opt_gemm_ctas = 0, opt_aev_two_pass = 1, opt_pin_user = 0;
  cudaEvent_t ev[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  float stage_ms[5] = {0, 0, 0, 0, 0};
  int launches = 0;
  const HostModel* hm = nullptr;
  std::unique_ptr<EngineBase> eng;
  // host scratch for list conversion
  std::vector<int64_t> h_beg;
  std::vector<int> h_cnt, h_species, h_tmp;
  std::vector<int> h_jl;
  // caller arrays page-locked on request (ani_ctx_pin_host), released by ani_ctx_release_host_buffers / destroy
  std::vector<void*> pinned_host;
  cudaStream_t comm_stream = nullptr;        // second stream + events of the overlapped ghost exchange (Engine::compute)
  cudaEvent_t evh[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
  int opt_halo_overlap = 1;
  unsigned long long* d_word = nullptr;      // one device word + its pinned host mirror (ani_md_max_disp2)
  unsigned long long* h_word = nullptr;
};

namespace anib {

template <typename T>
struct Engine : EngineBase {
  ani_ctx* c;
  const HostModel& hm;
  int M;   // members in use
  AevConst<T> P;
  std::vector<SpeciesNet<T>> nets;
  // per-rebuild state
  int nlocal = 0, ntotal = 0, cap = 32, capA = 64;
  int cap_fit = 0, capA_fit = 0;      // buffers sized from the survivor counts of recent steps (0 = unknown yet); see Aev2Rows
  bool stats_pending = false;
  DevBuf d_ovf, d_ovf_rows;           // [maxn, maxna, deferred rows fwd, deferred rows bwd]; row lists [2][nlocal]
  int64_t npairs = 0;
  std::vector<int> boff;   // bucket offsets [S+1] (rows of species s = [boff[s], boff[s+1]))
  std::vector<int> h_order, h_cspecies;
  DevBuf d_species, d_beg, d_cnt, d_jlist, d_order;
  const int* jlist_ext = nullptr;     // the device-built list in nb_jlist, used in place
  const int* jl() const { return jlist_ext ? jlist_ext : d_jlist.as<int>(); }
  // per-step state
  DevBuf d_xraw, d_x4, d_aev, d_daev, d_force, d_eatom, d_eatom_atom, d_partial, d_scalars, d_virial, d_err;
  DevBuf d_aev_hi, d_aev_lo;
  DevBuf d_H, d_D, d_G;   // MLP intermediates of one chunk
  TcWork tcw;
  f16::Work hwork;
  f16::FusedWork fwork;
  int64_t ld_planes = 0;   // row pitch of the fp16 AEV planes (multiple of 8 halves = 16 bytes for TMA)
  PinnedBuf h_x, h_f, h_small;
  UserPin pin_x, pin_f;
  DevBuf d_fout;
  bool f_direct = false;
  int64_t ld_aev = 0;
  int chunk_rows = 2048;

  // Active-species AEV columns.  An AEV column belongs to one neighbour species (radial) or one unordered pair of
  // neighbour species (angular); columns of species that occur nowhere in the subdomain (local + ghost) are structurally
  // zero for every atom, so they -- and the matching columns of every first-layer weight matrix -- are dropped: the AEV
  // kernels run on the compact species indices cmap[s], rows have Lc = Sp*nR + Sp(Sp+1)/2*nA*nZ columns (water: 128 of
  // ANI-2x's 1008) and layer 1 / dE/dAEV contract over Lc.  Exact (only zero products are skipped).  The active set only
  // grows (a superset stays exact); ANI_B200_DENSE_AEV=1 keeps all columns.
  AevConst<T> Pfull;
  RepConst<T> rep;
  HostModel hmc;                // hm with the first-layer columns gathered (dims[s][0] = Lc); what the MLP paths upload
  uint32_t active_mask = 0;
  std::vector<int> cmap;        // species -> compact index, -1 when absent
  std::vector<int> col_full;    // compact column -> column of the full AEV
  bool dense_aev = false;

  int aev_gen = 2;   // 2: triple-per-lane kernels (aev2.cuh); 1: the first-generation kernels (aev.cuh), env ANI_B200_AEV=1
  Engine(ani_ctx* ctx, const HostModel& m, int M_) : c(ctx), hm(m), M(M_) {
    if (const char* g = getenv("ANI_B200_AEV")) aev_gen = atoi(g) == 1 ? 1 : 2;
    if (const char* g = getenv("ANI_B200_DENSE_AEV")) dense_aev = atoi(g) != 0;
    const char* e = getenv("ANI_B200_MLP_CHUNK");
    if (e && atoi(e) > 0) chunk_rows = atoi(e);
  }

  int init() {
    memset(&P, 0, sizeof(P));
    P.Rcr = (T)hm.Rcr; P.Rca = (T)hm.Rca; P.Rcr2 = (T)(hm.Rcr * hm.Rcr); P.Rca2 = (T)(hm.Rca * hm.Rca);
    P.EtaR = (T)hm.EtaR; P.EtaA = (T)hm.EtaA; P.Zeta = (T)hm.Zeta;
    P.pi_over_Rcr = (T)(M_PI / hm.Rcr); P.pi_over_Rca = (T)(M_PI / hm.Rca);
    for (int k = 0; k < hm.nR; k++) P.ShfR[k] = (T)hm.ShfR[k];
    for (int k = 0; k < hm.nA; k++) P.ShfA[k] = (T)hm.ShfA[k];
    for (int k = 0; k < hm.nZ; k++) { P.cosZ[k] = (T)cos(hm.ShfZ[k]); P.sinZ[k] = (T)sin(hm.ShfZ[k]); }
    P.S = hm.S; P.nR = hm.nR; P.nA = hm.nA; P.nZ = hm.nZ;
    P.rad_len = hm.rad_len(); P.ang_sub = hm.ang_sub(); P.aev_len = hm.aev_len(); P.npairs = hm.S * (hm.S + 1) / 2;
    if (P.ang_sub > 32) { set_error("nShfA*nShfZ > 32 is not supported by the angular kernels"); return 1; }
    if (hm.S > 31) { set_error("more than 31 species"); return 1; }
    int p = 0;
    for (int s = 0; s < hm.S; s++)
      for (int t = s; t < hm.S; t++) { P.triu[s][t] = p; P.triu[t][s] = p; p++; }
    Pfull = P;
    nets.resize(hm.S);
    ANIB_RESERVE(d_err, 2 * sizeof(int));   // [0] angular-neighbour overflow, [1] bit 0 fp16 operand range (mlp_f16_fused.cuh), bit 1 fixed-point range (aev2.cuh backward)
    ANIB_CUDA_OK(cudaMemset(d_err.p, 0, 2 * sizeof(int)));
    ANIB_RESERVE(d_scalars, 16 * sizeof(double));
    ANIB_RESERVE(d_virial, kVirialSlots * 9 * sizeof(double));
    ANIB_RESERVE(d_partial, 1024 * sizeof(double));
    return 0;
  }

  // ---------------------------------------------------------------------------------------------- active species
  // (Re)build the compact AEV layout for the species in `mask` and upload every network with its first layer gathered.
  int set_active(uint32_t mask) {
    active_mask = mask;
    cmap.assign(hm.S, -1);
    std::vector<int> act;
    for (int s = 0; s < hm.S; s++)
      if (mask >> s & 1u) { cmap[s] = (int)act.size(); act.push_back(s); }
    const int Sp = (int)act.size();
    P = Pfull;
    P.S = Sp; P.rad_len = Sp * P.nR; P.npairs = Sp * (Sp + 1) / 2; P.aev_len = P.rad_len + P.npairs * P.ang_sub;
    col_full.assign(P.aev_len, 0);
    for (int a = 0; a < Sp; a++)
      for (int k = 0; k < P.nR; k++) col_full[a * P.nR + k] = act[a] * Pfull.nR + k;
    int p = 0;
    for (int a = 0; a < Sp; a++)
      for (int b = a; b < Sp; b++) {
        P.triu[a][b] = p; P.triu[b][a] = p;
        for (int q = 0; q < P.ang_sub; q++)
          col_full[P.rad_len + p * P.ang_sub + q] = Pfull.rad_len + Pfull.triu[act[a]][act[b]] * P.ang_sub + q;
        p++;
      }
    if (hm.use_repulsion) {     // RepulsionXTB tables on the compact species indices the kernels see in x4.w
      memset(&rep, 0, sizeof(rep));
      for (int a = 0; a < Sp; a++)
        for (int b = 0; b < Sp; b++) {
          rep.y[a][b] = (T)hm.rep_y[(size_t)act[a] * hm.S + act[b]];
          rep.sa[a][b] = (T)hm.rep_sa[(size_t)act[a] * hm.S + act[b]];
          rep.k[a][b] = (T)hm.rep_k[(size_t)act[a] * hm.S + act[b]];
        }
      rep.rc = (T)hm.rep_cutoff; rep.rc2 = (T)(hm.rep_cutoff * hm.rep_cutoff); rep.fn = hm.rep_cutoff_fn;
    }
    ld_aev = (P.aev_len + 3) / 4 * 4;
    ld_planes = (P.aev_len + 63) / 64 * 64;   // 128-byte row pitch: every TMA box row starts on a cache-line boundary
    // gathered copy of the model: only dims / W / b are read from it
    const int Lc = P.aev_len, Lf = Pfull.aev_len;
    hmc = hm;
    for (int s = 0; s < hm.S; s++) {
      if (hm.dims[s].empty() || hm.dims[s][0] != Lf) { set_error("network input width does not match the AEV length"); return 1; }
      hmc.dims[s][0] = Lc;
      const int h1 = hm.dims[s][1];
      for (size_t m = 0; m < hm.W.size(); m++) {
        const std::vector<float>& src = hm.W[m][s][0];
        std::vector<float>& dst = hmc.W[m][s][0];
        dst.assign((size_t)h1 * Lc, 0.f);
        for (int k = 0; k < h1; k++)
          for (int q = 0; q < Lc; q++) dst[(size_t)k * Lc + q] = src[(size_t)k * Lf + col_full[q]];
      }
    }
    return upload_weights(hmc);
  }

  int upload_tensor_core_planes(const HostModel& src);   // fp32 engine only

  int upload_weights(const HostModel& src) {
    for (int s = 0; s < hm.S; s++) {
      SpeciesNet<T>& n = nets[s];
      n.dims = src.dims[s];
      n.nl = (int)n.dims.size() - 1;
      n.W.resize(n.nl); n.b.resize(n.nl); n.WT.resize(n.nl);
      for (int l = 0; l < n.nl; l++) {
        const int ni = n.dims[l], no = n.dims[l + 1];
        std::vector<T> w((size_t)M * no * ni), bb((size_t)M * no);
        for (int m = 0; m < M; m++) {
          for (size_t q = 0; q < (size_t)no * ni; q++) w[(size_t)m * no * ni + q] = (T)src.W[m][s][l][q];
          for (int q = 0; q < no; q++) bb[(size_t)m * no + q] = (T)src.b[m][s][l][q];
        }
        ANIB_RESERVE(n.W[l], w.size() * sizeof(T));
        ANIB_RESERVE(n.b[l], bb.size() * sizeof(T));
        ANIB_CUDA_OK(cudaMemcpy(n.W[l].p, w.data(), w.size() * sizeof(T), cudaMemcpyHostToDevice));
        ANIB_CUDA_OK(cudaMemcpy(n.b[l].p, bb.data(), bb.size() * sizeof(T), cudaMemcpyHostToDevice));
        if (l >= 1 && l < n.nl - 1) {
          std::vector<T> wt((size_t)M * ni * no);
          for (int m = 0; m < M; m++)
            for (int o = 0; o < no; o++)
              for (int k = 0; k < ni; k++)
                wt[((size_t)m * ni + k) * no + o] = (T)src.W[m][s][l][(size_t)o * ni + k];
          ANIB_RESERVE(n.WT[l], wt.size() * sizeof(T));
          ANIB_CUDA_OK(cudaMemcpy(n.WT[l].p, wt.data(), wt.size() * sizeof(T), cudaMemcpyHostToDevice));
        }
      }
      {
        const int L = n.dims[0], h1 = n.dims[1];
        std::vector<T> cat((size_t)L * M * h1);
        for (int m = 0; m < M; m++)
          for (int k = 0; k < h1; k++)
            for (int q = 0; q < L; q++) cat[(size_t)q * M * h1 + (size_t)m * h1 + k] = (T)src.W[m][s][0][(size_t)k * L + q];
        ANIB_RESERVE(n.W1cat, cat.size() * sizeof(T));
        ANIB_CUDA_OK(cudaMemcpy(n.W1cat.p, cat.data(), cat.size() * sizeof(T), cudaMemcpyHostToDevice));
      }
    }
    return upload_tensor_core_planes(src);
  }

  // ---------------------------------------------------------------------------------------------- rebuild
  int set_neighbors(int nlocal_, int ntotal_, const int* species_h, const int64_t* beg_h, const int* cnt_h,
                    const int* jlist, int64_t npairs_, bool jlist_on_device, int csr_dev_maxnn = -1,
                    bool bucketed_on_device = false) override {
    nlocal = nlocal_; ntotal = ntotal_; npairs = npairs_;
    cudaStream_t st = c->stream;
    // species bucketing (stable counting sort over local atoms) -- K4 of SURVEY.md 2.1, on rebuild only.  The device build has
    // done it already (boff set, d_species / d_order filled: build_neighbors); a host list is bucketed here.
    if (!bucketed_on_device) {
    rows_int = -1;                                   // a host list comes without the interior / boundary split
    boff.assign(hm.S + 1, 0);
    for (int i = 0; i < nlocal; i++) {
      int s = species_h[i];
      if (s < 0 || s >= hm.S) { set_error("atom " + std::to_string(i) + " has species index " + std::to_string(s) + " outside the model's species"); return 1; }
      boff[s + 1]++;
    }
    for (int i = nlocal; i < ntotal; i++)
      if (species_h[i] < 0 || species_h[i] >= hm.S) { set_error("ghost atom has a species index outside the model's species"); return 1; }
    for (int s = 0; s < hm.S; s++) boff[s + 1] += boff[s];
    {
      uint32_t present = 0;
      for (int i = 0; i < ntotal; i++) present |= 1u << species_h[i];
      if (dense_aev) present = (1u << hm.S) - 1u;
      if (present == 0) present = 1u;
      if ((present & ~active_mask) != 0)
        if (int rc = set_active(active_mask | present)) return rc;
      h_cspecies.resize(ntotal);
      for (int i = 0; i < ntotal; i++) h_cspecies[i] = cmap[species_h[i]];
    }
    h_order.resize(nlocal);
    {
      std::vector<int> cur(boff.begin(), boff.end() - 1);
      for (int i = 0; i < nlocal; i++) h_order[cur[species_h[i]]++] = i;
    }
    }
    const bool csr_dev = csr_dev_maxnn >= 0;
    int maxnn = csr_dev ? csr_dev_maxnn : 0;
    if (!csr_dev)
      for (int i = 0; i < nlocal; i++) maxnn = std::max(maxnn, cnt_h[i]);
    cap = std::max(32, (maxnn + 31) / 32 * 32);
    capA = std::min(cap, 96);
    ANIB_RESERVE(d_species, (size_t)ntotal * sizeof(int));
    ANIB_RESERVE(d_beg, (size_t)std::max(nlocal, 1) * sizeof(int64_t));
    ANIB_RESERVE(d_cnt, (size_t)std::max(nlocal, 1) * sizeof(int));
    ANIB_RESERVE(d_order, (size_t)std::max(nlocal, 1) * sizeof(int));
    if (!bucketed_on_device)
      ANIB_CUDA_OK(cudaMemcpyAsync(d_species.p, h_cspecies.data(), (size_t)ntotal * sizeof(int), cudaMemcpyHostToDevice, st));   // compact indices
    const cudaMemcpyKind ck = csr_dev ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice;
    ANIB_CUDA_OK(cudaMemcpyAsync(d_beg.p, beg_h, (size_t)nlocal * sizeof(int64_t), ck, st));
    ANIB_CUDA_OK(cudaMemcpyAsync(d_cnt.p, cnt_h, (size_t)nlocal * sizeof(int), ck, st));
    if (!bucketed_on_device)
      ANIB_CUDA_OK(cudaMemcpyAsync(d_order.p, h_order.data(), (size_t)nlocal * sizeof(int), cudaMemcpyHostToDevice, st));
    if (jlist_on_device && csr_dev) {
      jlist_ext = jlist;                          // the list the device build just wrote stays where it is (no 600 MB copy)
    } else {
      jlist_ext = nullptr;
      ANIB_RESERVE(d_jlist, (size_t)std::max<int64_t>(npairs, 1) * sizeof(int));
      ANIB_CUDA_OK(cudaMemcpyAsync(d_jlist.p, jlist, (size_t)npairs * sizeof(int),
                                   jlist_on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, st));
    }
    ANIB_CUDA_OK(cudaStreamSynchronize(st));   // host staging vectors may be reused by the caller
    // per-step buffers
    ANIB_RESERVE(d_x4, (size_t)std::max(ntotal, 1) * 4 * sizeof(T));
    ANIB_RESERVE(d_xraw, (size_t)std::max(ntotal, 1) * 3 * sizeof(double));
    ANIB_RESERVE(d_force, (size_t)std::max(ntotal, 1) * 3 * sizeof(T));
    ANIB_RESERVE(d_aev, (size_t)std::max(nlocal, 1) * ld_aev * sizeof(T));
    ANIB_RESERVE(d_daev, (size_t)std::max(nlocal, 1) * ld_aev * sizeof(T));
    ANIB_RESERVE(d_eatom, (size_t)std::max(nlocal, 1) * sizeof(double));
    ANIB_RESERVE(d_eatom_atom, (size_t)std::max(nlocal, 1) * sizeof(double));
    return 0;
  }

  // ---------------------------------------------------------------------------------------------- device neighbour build
  DevBuf nb_x, nb_xs, nb_spec, nb_bnd, nb_key, d_rows, d_sel, nb_cell, nb_atom, nb_cell_s, nb_atom_s, nb_cbeg, nb_cend, nb_cnt, nb_beg, nb_jlist, nb_tmp;
  int build_neighbors(int nlocal_, int ntotal_, const int* species_h, const double* x, bool x_on_device, double cutoff,
                      const double* lo, const double* hi, int64_t* npairs_out) override {
    cudaStream_t st = c->stream;
    if (!(cutoff > 0.0)) { set_error("ani_build_neighbors: cutoff must be positive"); return 1; }
    NbGrid G;
    G.inv_cell = 1.0 / cutoff;            // cubic cells of exactly one cutoff: 27 cells cover every neighbour; atoms outside
    for (int d = 0; d < 3; d++) {         // [lo, hi) are clamped into the border cells (a monotone map: |delta cell| <= 1 holds)
      if (!(hi[d] > lo[d])) { set_error("ani_build_neighbors: empty bounding box"); return 1; }
      G.lo[d] = lo[d];
      G.n[d] = std::max(1, (int)std::ceil((hi[d] - lo[d]) * G.inv_cell));
    }
    const int64_t ncell = (int64_t)G.n[0] * G.n[1] * G.n[2];
    if (ncell > (int64_t)1 << 30) { set_error("ani_build_neighbors: too many cells"); return 1; }
    const int nt = std::max(ntotal_, 1);
    const double* xd = x;
    if (!x_on_device) {
      // a registration of this array left over from a smaller / moved allocation must go before the copy below touches it
      if (pin_x.p != nullptr && (pin_x.p != (const void*)x || (size_t)ntotal_ * 3 * sizeof(double) > pin_x.bytes)) pin_x.release();
      ANIB_RESERVE(nb_x, (size_t)nt * 3 * sizeof(double));
      ANIB_CUDA_OK(cudaMemcpyAsync(nb_x.p, x, (size_t)ntotal_ * 3 * sizeof(double), cudaMemcpyHostToDevice, st));
      xd = nb_x.as<double>();
    }
    ANIB_RESERVE(nb_cell, (size_t)nt * 4); ANIB_RESERVE(nb_atom, (size_t)nt * 4);
    ANIB_RESERVE(nb_cell_s, (size_t)nt * 4); ANIB_RESERVE(nb_atom_s, (size_t)nt * 4);
    ANIB_RESERVE(nb_cbeg, (size_t)ncell * 4); ANIB_RESERVE(nb_cend, (size_t)ncell * 4);
    ANIB_RESERVE(nb_cnt, (size_t)std::max(nlocal_, 1) * 4); ANIB_RESERVE(nb_beg, (size_t)std::max(nlocal_, 1) * 8);
    int64_t P = 0;
    int maxnn = 0;
    bool bucketed = false;
    static const bool nbprof = f16::env_int("ANI_B200_NB_PROF", 0) != 0;
    auto tnow = [] { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    double tp[6] = {0, 0, 0, 0, 0, 0};
    if (nbprof) { cudaStreamSynchronize(st); tp[0] = tnow(); }
    if (ntotal_ > 0 && nlocal_ > 0) {
      // species of all atoms to the device; histogram / present set / range check there (read back with the two list scalars)
      ANIB_RESERVE(nb_spec, (size_t)nt * 4);
      ANIB_RESERVE(d_species, (size_t)nt * sizeof(int));
      ANIB_RESERVE(d_order, (size_t)std::max(nlocal_, 1) * sizeof(int));
      ANIB_RESERVE(nb_red, 128);
      ANIB_RESERVE(nb_bnd, (size_t)std::max(nlocal_, 1) * 4);
      ANIB_RESERVE(nb_key, (size_t)std::max(nlocal_, 1) * 4);
      ANIB_RESERVE(d_rows, (size_t)std::max(nlocal_, 1) * 4);
      ANIB_CUDA_OK(cudaMemcpyAsync(nb_spec.p, species_h, (size_t)ntotal_ * 4, cudaMemcpyHostToDevice, st));
      ANIB_CUDA_OK(cudaMemsetAsync(nb_red.p, 0, 128, st));
      nb_cell_id_kernel<<<(ntotal_ + 255) / 256, 256, 0, st>>>(xd, ntotal_, G, nb_cell.as<int>(), nb_atom.as<int>());
      size_t tmp_bytes = 0, scan_bytes = 0, max_bytes = 0;
      int bits = 1;
      while (((int64_t)1 << bits) < ncell) bits++;
      cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, nb_cell.as<int>(), nb_cell_s.as<int>(), nb_atom.as<int>(), nb_atom_s.as<int>(),
                                      ntotal_, 0, bits, st);
      auto cnt64 = thrust::make_transform_iterator((const int*)nb_cnt.as<int>(), NbToI64());
      cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, cnt64, nb_beg.as<int64_t>(), nlocal_, st);
      cub::DeviceReduce::Max(nullptr, max_bytes, nb_cnt.as<int>(), (int*)nb_red.p, nlocal_, st);
      size_t order_bytes = 0, sum_bytes = 0;
      cub::DeviceReduce::Sum(nullptr, sum_bytes, cnt64, (int64_t*)nullptr, nlocal_, st);
      max_bytes = std::max(max_bytes, sum_bytes);
      int sbits = 1;
      while ((1 << sbits) < hm.S) sbits++;
      sbits++;                                              // the key is 2 * species + boundary flag
      cub::DeviceRadixSort::SortPairs(nullptr, order_bytes, nb_key.as<int>(), nb_cell.as<int>(), nb_atom.as<int>(), d_order.as<int>(),
                                      nlocal_, 0, sbits, st);
      tmp_bytes = std::max(std::max(tmp_bytes, order_bytes), std::max(scan_bytes, max_bytes));
      ANIB_RESERVE(nb_tmp, tmp_bytes + 16);
      ANIB_CUDA_OK(cub::DeviceRadixSort::SortPairs(nb_tmp.p, tmp_bytes, nb_cell.as<int>(), nb_cell_s.as<int>(), nb_atom.as<int>(),
                                                   nb_atom_s.as<int>(), ntotal_, 0, bits, st));
      ANIB_CUDA_OK(cudaMemsetAsync(nb_cbeg.p, 0, (size_t)ncell * 4, st));
      ANIB_CUDA_OK(cudaMemsetAsync(nb_cend.p, 0, (size_t)ncell * 4, st));
      nb_cell_range_kernel<<<(ntotal_ + 255) / 256, 256, 0, st>>>(nb_cell_s.as<int>(), ntotal_, nb_cbeg.as<int>(), nb_cend.as<int>());
      ANIB_RESERVE(nb_xs, (size_t)nt * (3 * sizeof(double) + sizeof(float4)));
      float4* const pk = nb_xs.as<float4>();                // (x, y, z relative to the grid origin, atom index) in sorted order
      double* const sx = (double*)(pk + nt);
      double* const sy = sx + nt;
      double* const sz = sy + nt;
      nb_gather_sorted_kernel<<<(ntotal_ + 255) / 256, 256, 0, st>>>(xd, nb_atom_s.as<int>(), ntotal_, G, sx, sy, sz, pk);
      // fp32 error bound of a squared distance near the cutoff: each relative coordinate is off by at most half an ulp of the
      // grid extent (plus the clamped atoms just outside it: one more cell), a difference by twice that, d^2 by 2 |d| sqrt(3)
      // times that; an order of magnitude of margin on top
      double extent = 0.0;
      for (int d = 0; d < 3; d++) extent = std::max(extent, hi[d] - lo[d] + 2.0 * cutoff);
      const float eps = (float)(64.0 * cutoff * extent * 5.960464477539063e-08 + 1e-5 * cutoff * cutoff);
      const unsigned search_grid = (unsigned)(((size_t)ntotal_ * 32 + 255) / 256);        // one warp per atom, local or ghost
      // One pass when the longest row of the previous build is known: rows of that length plus a margin, counted and written
      // together.  Otherwise (first build, or a row outgrew the stride) count, scan, fill.
      int stride = nb_stride;
      if (stride > 0 && (size_t)nlocal_ * stride * 4 > ((size_t)8 << 30)) stride = 0;     // keep the padded list under 8 GB
      if (stride > 0) {
        ANIB_RESERVE(nb_jlist, (size_t)nlocal_ * stride * 4);
        nb_strided_beg_kernel<<<(nlocal_ + 255) / 256, 256, 0, st>>>(nb_beg.as<int64_t>(), nlocal_, stride);
        nb_search_warp_kernel<2><<<search_grid, 256, 0, st>>>(sx, sy, sz, pk, ntotal_, nlocal_, G, nb_cbeg.as<int>(), nb_cend.as<int>(),
                                                              cutoff * cutoff, eps, nb_cnt.as<int>(), nb_bnd.as<int>(), nullptr, stride,
                                                              nb_jlist.as<int>());
        ANIB_CUDA_OK(cub::DeviceReduce::Sum(nb_tmp.p, tmp_bytes, cnt64, (int64_t*)((char*)nb_red.p + 8), nlocal_, st));
      } else {
        nb_search_warp_kernel<0><<<search_grid, 256, 0, st>>>(sx, sy, sz, pk, ntotal_, nlocal_, G, nb_cbeg.as<int>(), nb_cend.as<int>(),
                                                              cutoff * cutoff, eps, nb_cnt.as<int>(), nb_bnd.as<int>(), nullptr, 0, nullptr);
        // CSR offsets and the total on the device
        ANIB_CUDA_OK(cub::DeviceScan::ExclusiveSum(nb_tmp.p, tmp_bytes, cnt64, nb_beg.as<int64_t>(), nlocal_, st));
        nb_total_kernel<<<1, 1, 0, st>>>(nb_beg.as<int64_t>(), nb_cnt.as<int>(), nlocal_, (int64_t*)((char*)nb_red.p + 8));
      }
      ANIB_CUDA_OK(cub::DeviceReduce::Max(nb_tmp.p, tmp_bytes, nb_cnt.as<int>(), (int*)nb_red.p, nlocal_, st));
      nb_species_stats_kernel<<<std::min(592, (ntotal_ + 255) / 256), 256, 0, st>>>(nb_spec.as<int>(), nb_bnd.as<int>(), nlocal_, ntotal_, hm.S,
                                                                                  nb_key.as<int>(), (int*)((char*)nb_red.p + 16));
      ANIB_RESERVE(h_nb, 128);
      if (nbprof) tp[1] = tnow();
      ANIB_CUDA_OK(cudaMemcpyAsync(h_nb.p, nb_red.p, 128, cudaMemcpyDeviceToHost, st));       // longest row, total, species statistics
      ANIB_CUDA_OK(cudaStreamSynchronize(st));
      if (nbprof) tp[2] = tnow();
      if (stride > 0 && *(const int*)h_nb.p > stride) {      // a row did not fit: this build goes the two-pass way
        stride = 0;
        ANIB_CUDA_OK(cub::DeviceScan::ExclusiveSum(nb_tmp.p, tmp_bytes, cnt64, nb_beg.as<int64_t>(), nlocal_, st));
      }
      maxnn = *(const int*)h_nb.p;
      P = *(const int64_t*)((const char*)h_nb.p + 8);
      {
        const int* stats = (const int*)((const char*)h_nb.p + 16);
        if (stats[17]) { set_error("an atom has a species index outside the model's species"); return 1; }
        boff.assign(hm.S + 1, 0);
        NbRowSplit sp;
        memset(&sp, 0, sizeof(sp));
        sp.S = hm.S;
        int n_int = 0, n_bnd = 0;
        for (int q = 0; q < hm.S; q++) {
          boff[q + 1] = boff[q] + stats[2 * q] + stats[2 * q + 1];
          sp.nint[q] = stats[2 * q]; sp.int_before[q] = n_int; sp.bnd_before[q] = n_bnd;
          n_int += stats[2 * q]; n_bnd += stats[2 * q + 1];
        }
        for (int q = 0; q <= hm.S; q++) sp.boff[q] = boff[q];
        sp.n_int = n_int;
        rows_int = n_int; rows_bnd = n_bnd;
        uint32_t present = (uint32_t)stats[16];
        if (dense_aev) present = (1u << hm.S) - 1u;
        if (present == 0) present = 1u;
        if ((present & ~active_mask) != 0)
          if (int rc = set_active(active_mask | present)) return rc;
        NbSpeciesMap sm;
        for (int q = 0; q < 8; q++) sm.cmap[q] = q < hm.S ? cmap[q] : -1;
        nb_species_compact_kernel<<<(ntotal_ + 255) / 256, 256, 0, st>>>(nb_spec.as<int>(), ntotal_, sm, d_species.as<int>());
        // row order = local atoms bucketed by species, atom order kept inside a bucket: a stable sort of (species, index);
        // nb_atom still holds 0, 1, 2, ... and the unsorted cell ids in nb_cell are no longer needed
        ANIB_CUDA_OK(cub::DeviceRadixSort::SortPairs(nb_tmp.p, tmp_bytes, nb_key.as<int>(), nb_cell.as<int>(), nb_atom.as<int>(),
                                                     d_order.as<int>(), nlocal_, 0, sbits, st));
        nb_row_lists_kernel<<<(nlocal_ + 255) / 256, 256, 0, st>>>(sp, nlocal_, d_rows.as<int>());
        const int hsel[2] = {n_int, n_bnd};
        ANIB_RESERVE(d_sel, 2 * sizeof(int));
        ANIB_CUDA_OK(cudaMemcpyAsync(d_sel.p, hsel, sizeof(hsel), cudaMemcpyHostToDevice, st));   // (staged by the runtime: hsel may go)
        bucketed = true;
      }
      if (stride == 0) {
        ANIB_RESERVE(nb_jlist, (size_t)std::max<int64_t>(P, 1) * 4);
        nb_search_warp_kernel<1><<<search_grid, 256, 0, st>>>(sx, sy, sz, pk, ntotal_, nlocal_, G, nb_cbeg.as<int>(), nb_cend.as<int>(),
                                                              cutoff * cutoff, eps, nullptr, nullptr, nb_beg.as<int64_t>(), 0, nb_jlist.as<int>());
      }
      nb_rows_strided = stride > 0;
      nb_stride = (maxnn + maxnn / 8 + 8 + 7) / 8 * 8;       // rows of the next build
      ANIB_CUDA_OK(cudaGetLastError());
      c->launches = 11;
    }
    if (npairs_out) *npairs_out = P;
    if (nbprof) { tp[3] = tnow(); cudaStreamSynchronize(st); tp[4] = tnow(); }
    const int rc = set_neighbors(nlocal_, ntotal_, species_h, nb_beg.as<int64_t>(), nb_cnt.as<int>(), nb_jlist.as<int>(), P, true, maxnn, bucketed);
    if (nbprof) {
      tp[5] = tnow();
      fprintf(stderr, "[nb prof] enqueue (species H2D, sort, search) %.3f ms | wait for it %.3f | enqueue the rest %.3f | done %.3f | "
                      "set_neighbors %.3f | total %.3f ms\n", tp[1] - tp[0], tp[2] - tp[1], tp[3] - tp[2], tp[4] - tp[3], tp[5] - tp[4], tp[5] - tp[0]);
    }
    return rc;
  }
  DevBuf nb_red;
  PinnedBuf h_nb;
  int nb_stride = 0;              // row stride for the one-pass build (0: unknown yet)
  int rows_int = -1, rows_bnd = 0;   // interior / boundary rows listed in d_rows (device build); -1: no split known
  bool nb_rows_strided = false;   // the installed device-built list has padded rows

  void release_pins() override { pin_x.release(); pin_f.release(); }

  int get_neighbors(int* numneigh, int* jlist) override {
    ANIB_CUDA_OK(cudaStreamSynchronize(c->stream));
    if (numneigh && nlocal > 0) ANIB_CUDA_OK(cudaMemcpy(numneigh, d_cnt.p, (size_t)nlocal * 4, cudaMemcpyDeviceToHost));
    if (jlist && npairs > 0) {
      if (jlist_ext && nb_rows_strided) {                    // padded rows: copy row by row segments through the CSR offsets
        std::vector<int64_t> hb(nlocal);
        std::vector<int> hc(nlocal);
        ANIB_CUDA_OK(cudaMemcpy(hb.data(), d_beg.p, (size_t)nlocal * 8, cudaMemcpyDeviceToHost));
        ANIB_CUDA_OK(cudaMemcpy(hc.data(), d_cnt.p, (size_t)nlocal * 4, cudaMemcpyDeviceToHost));
        const int64_t span = nlocal > 0 ? hb[nlocal - 1] + hc[nlocal - 1] : 0;
        std::vector<int> all((size_t)std::max<int64_t>(span, 1));
        ANIB_CUDA_OK(cudaMemcpy(all.data(), jl(), (size_t)span * 4, cudaMemcpyDeviceToHost));
        int64_t p = 0;
        for (int i = 0; i < nlocal; i++)
          for (int q = 0; q < hc[i]; q++) jlist[p++] = all[(size_t)hb[i] + q];
      } else {
        ANIB_CUDA_OK(cudaMemcpy(jlist, jl(), (size_t)npairs * 4, cudaMemcpyDeviceToHost));
      }
    }
    return 0;
  }

  // ---------------------------------------------------------------------------------------------- MLP (SIMT)
  template <int EPI>
  void gemm(int R, int N, int K, const T* A, int64_t lda, int64_t sAz, const T* B, int64_t ldb, int64_t sBz, T* C,
            int64_t ldc, int64_t sCz, T* C2, const T* bias, int64_t sbz, const T* aux, int64_t ldx, int64_t sXz, T scale,
            int batch) {
    dim3 grid((N + 63) / 64, (R + 63) / 64, batch);
    gemm_nt_kernel<T, EPI><<<grid, 256, 0, c->stream>>>(R, N, K, A, lda, sAz, B, ldb, sBz, C, ldc, sCz, C2, bias, sbz, aux,
                                                        ldx, sXz, scale);
    c->launches++;
  }

  int mlp_simt_species(int s, int row0, int R) {
    SpeciesNet<T>& n = nets[s];
    const int nh = n.nl - 1;   // hidden layers
    // column offsets of each hidden layer inside one row of the [R][sum_l M*w_l] intermediates
    std::vector<int64_t> coff(nh + 1, 0);
    for (int l = 0; l < nh; l++) coff[l + 1] = coff[l] + (int64_t)M * n.dims[l + 1];
    const int64_t rowsz = coff[nh];
    ANIB_RESERVE(d_H, (size_t)chunk_rows * rowsz * sizeof(T));
    ANIB_RESERVE(d_D, (size_t)chunk_rows * rowsz * sizeof(T));
    ANIB_RESERVE(d_G, (size_t)chunk_rows * rowsz * sizeof(T));
    for (int r0 = 0; r0 < R; r0 += chunk_rows) {
      const int Rc = std::min(chunk_rows, R - r0);
      const T* aev = d_aev.as<T>() + (size_t)(row0 + r0) * ld_aev;
      T* daev = d_daev.as<T>() + (size_t)(row0 + r0) * ld_aev;
      // buffers are laid out layer after layer: layer l block = [Rc][M][w_l]
      auto Hl = [&](int l) { return d_H.as<T>() + (size_t)chunk_rows * coff[l]; };
      auto Dl = [&](int l) { return d_D.as<T>() + (size_t)chunk_rows * coff[l]; };
      auto Gl = [&](int l) { return d_G.as<T>() + (size_t)chunk_rows * coff[l]; };
      // forward hidden layers
      for (int l = 0; l < nh; l++) {
        const int ni = n.dims[l], no = n.dims[l + 1];
        const T* A = (l == 0) ? aev : Hl(l - 1);
        const int64_t lda = (l == 0) ? ld_aev : (int64_t)M * ni;
        const int64_t sAz = (l == 0) ? 0 : ni;
        gemm<EPI_BIAS_CELU>(Rc, no, ni, A, lda, sAz, n.W[l].template as<T>(), ni, (int64_t)no * ni, Hl(l), (int64_t)M * no, no,
                            Dl(l), n.b[l].template as<T>(), no, nullptr, 0, 0, T(1), M);
      }
      // output layer + dE/dz of the last hidden layer
      {
        const int h = n.dims[nh];
        mlp_out_kernel<T><<<(Rc + 3) / 4, 128, 0, c->stream>>>(Rc, M, h, Hl(nh - 1), Dl(nh - 1), n.W[nh].template as<T>(),
                                                              n.b[nh].template as<T>(), Gl(nh - 1),
                                                              d_eatom.as<double>() + row0 + r0, hm.sae[s]);
        c->launches++;
      }
      // backward through hidden layers nh-1 .. 1
      for (int l = nh - 1; l >= 1; l--) {
        const int ni = n.dims[l], no = n.dims[l + 1];   // layer l maps ni -> no; gradient goes no -> ni
        gemm<EPI_MUL_AUX>(Rc, ni, no, Gl(l), (int64_t)M * no, no, n.WT[l].template as<T>(), no, (int64_t)ni * no, Gl(l - 1),
                          (int64_t)M * ni, ni, nullptr, nullptr, 0, Dl(l - 1), (int64_t)M * ni, ni, T(1), M);
      }
      // dE/dAEV with the ensemble mean folded into K
      {
        const int h1 = n.dims[1];
        gemm<EPI_SCALE>(Rc, n.dims[0], M * h1, Gl(0), (int64_t)M * h1, 0, n.W1cat.template as<T>(), (int64_t)M * h1, 0, daev,
                        ld_aev, 0, nullptr, nullptr, 0, nullptr, 0, 0, T(1) / T(M), 1);
      }
    }
    return 0;
  }

  int mlp_tc_species(int s, int row0, int R);   // defined for float below
  int mlp_f16_species(int s, int row0, int R);
  int mlp_f16_all();
  bool aev_in_planes = false;

  // ---------------------------------------------------------------------------------------------- one step
  int compute(const void* x, bool x_on_device, bool x_is_float, double* out_energy, double* out_force_host, void* f_dev,
              bool f_is_float, double* out_eatom_host, double* eatom_dev, double* out_virial9,
              HaloPlan* halo = nullptr) override {
    cudaStream_t st = c->stream;
    c->launches = 0;
    const bool tm = c->timing;
    if (tm) cudaEventRecord(c->ev[0], st);
    // ---- (0) coordinates
    const void* xsrc = x;
    if (!x_on_device) {
      const size_t bytes = (size_t)ntotal * 3 * sizeof(double);
      if (pin_x.ensure(x, bytes, c->opt_pin_user != 0)) {
        ANIB_CUDA_OK(cudaMemcpyAsync(d_xraw.p, x, bytes, cudaMemcpyHostToDevice, st));   // DMA straight from the caller's array
      } else {
        ANIB_RESERVE(h_x, bytes);
        memcpy(h_x.p, x, bytes);   // pageable LAMMPS array -> pinned staging
        ANIB_CUDA_OK(cudaMemcpyAsync(d_xraw.p, h_x.p, bytes, cudaMemcpyHostToDevice, st));
      }
      xsrc = d_xraw.p;
      x_is_float = false;
    }
    using V4 = typename Vec4T<T>::type;
    // Ghost exchange of a domain-decomposed step (halo != nullptr: x and f are the caller's DEVICE arrays of doubles).  With
    // the interior / boundary row split of a device-built list the exchange runs on a second stream and overlaps the interior
    // atoms' share of the AEV kernels (their rows read no ghost position and give no ghost force); otherwise it is
    // serialised in front of and behind the step.
    const bool aev2_ok = aev_gen == 2 && ((P.nZ == 4 && P.nA == 8) || (P.nZ == 8 && P.nA == 4) || (P.nZ == 8 && P.nA == 8));
    if (halo && (!x_on_device || x_is_float || !f_dev || f_is_float)) {
      set_error("the ghost exchange needs device arrays of doubles for x and f");
      return 1;
    }
    const int nghost = ntotal - nlocal;
    const bool overlap = halo && c->opt_halo_overlap && rows_int >= 0 && aev2_ok && nlocal > 0 && nghost > 0;
    cudaStream_t cs = nullptr;
    if (overlap) {
      if (!c->comm_stream) {
        ANIB_CUDA_OK(cudaStreamCreateWithFlags(&c->comm_stream, cudaStreamNonBlocking));
        for (int q = 0; q < 5; q++) ANIB_CUDA_OK(cudaEventCreateWithFlags(&c->evh[q], cudaEventDisableTiming));
      }
      cs = c->comm_stream;
      ANIB_CUDA_OK(cudaEventRecord(c->evh[0], st));                 // x of the local atoms is final
      ANIB_CUDA_OK(cudaStreamWaitEvent(cs, c->evh[0], 0));
      if (int rc = halo_forward_on(*halo, (double*)const_cast<void*>(x), cs)) return rc;
      ANIB_CUDA_OK(cudaEventRecord(c->evh[1], cs));                 // ghost positions have arrived
    } else if (halo) {
      if (int rc = halo_forward_on(*halo, (double*)const_cast<void*>(x), st)) return rc;
    }
    auto pack_range = [&](int first, int count) {
      if (count <= 0) return;
      const void* src = x_is_float ? (const void*)((const float*)xsrc + (size_t)first * 3) : (const void*)((const double*)xsrc + (size_t)first * 3);
      pack_x_kernel<T><<<(count + 255) / 256, 256, 0, st>>>(src, x_is_float ? 1 : 0, d_species.as<int>() + first, count, d_x4.as<V4>() + first);
      c->launches++;
    };
    if (overlap) pack_range(0, nlocal);
    else pack_range(0, ntotal);
    ANIB_CUDA_OK(cudaMemsetAsync(d_force.p, 0, (size_t)ntotal * 3 * sizeof(T), st));
    const bool want_virial = out_virial9 != nullptr;
    if (want_virial) ANIB_CUDA_OK(cudaMemsetAsync(d_virial.p, 0, kVirialSlots * 9 * sizeof(double), st));
    // two-pass AEV kernels: first-pass buffers from the survivor counts the previous steps reported (performance only: a row
    // that does not fit is redone by the second pass with list-sized buffers)
    const size_t small_bytes = (4 + kVirialSlots * 9) * sizeof(double);
    ANIB_RESERVE(h_small, small_bytes);
    double* hs = h_small.as<double>();
    if (!stats_pending) memset(hs, 0, small_bytes);
    // flags raised by a step that was not synchronised (device-resident callers) surface here, one call late at most
    if (stats_pending) if (int rc = check_errors()) return rc;
    if (stats_pending) {
      const int* hst = (const int*)(hs + 2);
      const int mn = hst[0], mna = hst[1];
      if (mn > 0) cap_fit = (mn + std::max(4, mn / 8) + 7) / 8 * 8;
      if (mn > 0) capA_fit = (mna + std::max(4, mna / 8) + 7) / 8 * 8;
    }
    const bool two_pass_on = c->opt_aev_two_pass != 0;
    const int cap1 = std::min(cap, cap_fit), capA1 = std::min(capA, std::max(capA_fit, 8));
    const bool two_pass = two_pass_on && aev_gen == 2 && cap_fit > 0 && nlocal > 0 && (cap1 + 16 <= cap || capA1 + 16 <= capA);
    ANIB_RESERVE(d_ovf, 8 * sizeof(int));                 // [max n, max na, deferred rows: forward, backward, backward (interior)]
    ANIB_RESERVE(d_ovf_rows, (size_t)std::max(nlocal, 1) * 3 * sizeof(int));
    ANIB_CUDA_OK(cudaMemsetAsync(d_ovf.p, 0, 8 * sizeof(int), st));
    if (tm) cudaEventRecord(c->ev[1], st);
    // ---- (1) AEV forward
    const bool f16_mode = (sizeof(T) == 4) && c->mlp_mode == ANI_MLP_F16X3;
    const bool tc_mode = (sizeof(T) == 4) && c->mlp_mode != ANI_MLP_FP32_SIMT && !f16_mode;
    aev_in_planes = f16_mode;
    const int npl = 32 / (P.nR <= 16 ? 16 : 32);
    if (nlocal > 0) {
      const int acc_f = P.aev_len + (npl - 1) * P.rad_len;
      AevSmemLayout Lf = aev_smem_layout<T>(cap, capA, acc_f, false);
      const int smem = Lf.bytes * kAevWarps;
      ANIB_CUDA_OK(cudaFuncSetAttribute(aev_forward_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
      void *hi = nullptr, *lo = nullptr;
      int plane_mode = 0;
      int64_t ldp = ld_aev;
      T* rows = d_aev.as<T>();
      if (f16_mode) {
        ANIB_RESERVE(d_aev_hi, (size_t)f16::padded_rows(nlocal) * ld_planes * sizeof(__half));
        ANIB_RESERVE(d_aev_lo, (size_t)f16::padded_rows(nlocal) * ld_planes * sizeof(__half));
        hi = d_aev_hi.p; lo = d_aev_lo.p; plane_mode = 2; ldp = ld_planes;
        rows = nullptr;   // the fp32 rows are not needed: the MLP reads the planes (get_aev rebuilds them from the planes)
      } else if (tc_mode) {
        ANIB_RESERVE(d_aev_hi, (size_t)tc_padded_rows(nlocal, hm.S) * ld_aev * sizeof(float));
        ANIB_RESERVE(d_aev_lo, (size_t)tc_padded_rows(nlocal, hm.S) * ld_aev * sizeof(float));
        hi = d_aev_hi.p; lo = d_aev_lo.p; plane_mode = 1;
      }
      auto launch2 = [&](auto kern, auto kern_strided) -> int {
        const Aev2Layout L2 = aev2_layout<T>(cap, capA, P.aev_len, P.nZ, P.nA, false);
        const Aev2Layout L1 = aev2_layout<T>(cap1, capA1, P.aev_len, P.nZ, P.nA, false);
        const int smem2 = L2.bytes * kAev2Warps;
        ANIB_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem2));
        ANIB_CUDA_OK(cudaFuncSetAttribute(kern_strided, cudaFuncAttributeMaxDynamicSharedMemorySize, smem2));
        const int grid = (nlocal + kAev2Warps - 1) / kAev2Warps;
        // one launch: `direct` = every row, one per warp; otherwise the rows of a list (grid-stride).  small = the buffers that
        // fit recent steps, rows that do not fit being deferred to the list the last launch works off with list-sized buffers
        auto go = [&](bool direct, int g, bool small, Aev2Rows rs) {
          const int cp = small ? cap1 : cap, cpA = small ? capA1 : capA, lb = small ? L1.bytes : L2.bytes;
          if (direct)
            kern<<<g, kAev2Warps * 32, lb * kAev2Warps, st>>>(
                P, c->use_cuaev ? 1 : 0, d_x4.as<V4>(), d_beg.as<int64_t>(), d_cnt.as<int>(), jl(), d_order.as<int>(),
                nlocal, rows, ld_aev, hi, lo, ldp, plane_mode, f16::kScaleAev, cp, cpA, lb, d_err.as<int>(), rs);
          else
            kern_strided<<<g, kAev2Warps * 32, lb * kAev2Warps, st>>>(
                P, c->use_cuaev ? 1 : 0, d_x4.as<V4>(), d_beg.as<int64_t>(), d_cnt.as<int>(), jl(), d_order.as<int>(),
                nlocal, rows, ld_aev, hi, lo, ldp, plane_mode, f16::kScaleAev, cp, cpA, lb, d_err.as<int>(), rs);
          c->launches++;
        };
        Aev2Rows rs;
        rs.stats = d_ovf.as<int>();
        if (two_pass) { rs.ovf_cnt = d_ovf.as<int>() + 2; rs.ovf_rows = d_ovf_rows.as<int>(); }
        if (overlap) {
          // interior rows while the ghost positions travel, then the ghosts' coordinates, then the boundary rows
          Aev2Rows ri = rs;
          ri.sel_cnt = d_sel.as<int>(); ri.sel_rows = d_rows.as<int>();
          if (rows_int > 0) go(false, (rows_int + kAev2Warps - 1) / kAev2Warps, two_pass, ri);
          ANIB_CUDA_OK(cudaStreamWaitEvent(st, c->evh[1], 0));
          pack_range(nlocal, nghost);
          Aev2Rows rb = rs;
          rb.sel_cnt = d_sel.as<int>() + 1; rb.sel_rows = d_rows.as<int>() + rows_int;
          if (rows_bnd > 0) go(false, (rows_bnd + kAev2Warps - 1) / kAev2Warps, two_pass, rb);
        } else {
          go(true, grid, two_pass, rs);
        }
        if (two_pass) {   // the deferred rows
          Aev2Rows r2;
          r2.sel_cnt = d_ovf.as<int>() + 2; r2.sel_rows = d_ovf_rows.as<int>();
          go(false, std::min(grid, 592), false, r2);
        }
        c->launches--;    // (the caller counts the last launch)
        return 0;
      };
      int rc2 = 0;
      if (aev_gen == 2 && P.nZ == 4 && P.nA == 8) rc2 = launch2(aev2_forward_kernel<T, 4, 8>, aev2_forward_kernel<T, 4, 8, true>);
      else if (aev_gen == 2 && P.nZ == 8 && P.nA == 4) rc2 = launch2(aev2_forward_kernel<T, 8, 4>, aev2_forward_kernel<T, 8, 4, true>);
      else if (aev_gen == 2 && P.nZ == 8 && P.nA == 8) rc2 = launch2(aev2_forward_kernel<T, 8, 8>, aev2_forward_kernel<T, 8, 8, true>);
      else
        aev_forward_kernel<T><<<(nlocal + kAevWarps - 1) / kAevWarps, kAevWarps * 32, smem, st>>>(
            P, c->use_cuaev ? 1 : 0, d_x4.as<V4>(), d_beg.as<int64_t>(), d_cnt.as<int>(), jl(),
            d_order.as<int>(), nlocal, rows, ld_aev, hi, lo, ldp, plane_mode, f16::kScaleAev, cap, capA, Lf.bytes,
            d_err.as<int>());
      if (rc2) return rc2;
      c->launches++;
    }
    if (tm) cudaEventRecord(c->ev[2], st);
    // ---- (2) MLP forward + backward per species bucket
    if (f16_mode) {
      if (int rc = mlp_f16_all()) return rc;
    } else
    for (int s = 0; s < hm.S; s++) {
      const int R = boff[s + 1] - boff[s];
      if (R == 0) continue;
      int rc = f16_mode ? mlp_f16_species(s, boff[s], R) : tc_mode ? mlp_tc_species(s, boff[s], R) : mlp_simt_species(s, boff[s], R);
      if (rc) return rc;
    }
    if (tm) cudaEventRecord(c->ev[3], st);
    const bool want_e = out_energy != nullptr;
    // ---- (3b) RepulsionXTB pair term (models with rep_calc): forces / virial into the same arrays, energy into scalars[1]
    // (in front of the AEV backward: with the overlapped exchange the ghost forces leave in the middle of that)
    if (hm.use_repulsion && nlocal > 0) {
      if (want_e) ANIB_CUDA_OK(cudaMemsetAsync(d_scalars.as<double>() + 1, 0, sizeof(double), st));
      repulsion_kernel<T><<<(unsigned)(((int64_t)nlocal * 32 + 255) / 256), 256, 0, st>>>(
          rep, d_x4.as<V4>(), d_beg.as<int64_t>(), d_cnt.as<int>(), jl(), nlocal, d_force.as<T>(),
          want_virial ? d_virial.as<double>() : nullptr, want_e ? d_scalars.as<double>() + 1 : nullptr, (T)kHartree2KcalMol);
      c->launches++;
    }
    // ---- (3) AEV backward: forces on local + ghost atoms, virial
    if (nlocal > 0) {
      const int acc_b = P.S * (P.nR + 1) + P.npairs * (P.ang_sub + 1);
      AevSmemLayout Lb = aev_smem_layout<T>(cap, capA, acc_b, true);
      const int smem = Lb.bytes * kAevWarps;
      auto launch = [&](auto kern) -> int {
        ANIB_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        kern<<<(nlocal + kAevWarps - 1) / kAevWarps, kAevWarps * 32, smem, st>>>(
            P, c->use_cuaev ? 1 : 0, d_x4.as<V4>(), d_beg.as<int64_t>(), d_cnt.as<int>(), jl(),
            d_order.as<int>(), nlocal, d_daev.as<T>(), ld_aev, d_force.as<T>(),
            want_virial ? d_virial.as<double>() : nullptr, (T)kHartree2KcalMol, cap, capA, Lb.bytes);
        c->launches++;
        return 0;
      };
      auto launch2 = [&](auto kern, auto kern_strided) -> int {
        const int acc_b2 = aev2_bwd_acc_elems(P.S, P.nR, P.npairs, P.ang_sub);
        const Aev2Layout L2 = aev2_layout<T>(cap, capA, acc_b2, P.nZ, P.nA, true);
        const Aev2Layout L1 = aev2_layout<T>(cap1, capA1, acc_b2, P.nZ, P.nA, true);
        const int smem2 = L2.bytes * kAev2Warps;
        ANIB_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem2));
        ANIB_CUDA_OK(cudaFuncSetAttribute(kern_strided, cudaFuncAttributeMaxDynamicSharedMemorySize, smem2));
        const int grid = (nlocal + kAev2Warps - 1) / kAev2Warps;
        auto go = [&](bool direct, int g, bool small, Aev2Rows rs) {
          const int cp = small ? cap1 : cap, cpA = small ? capA1 : capA, lb = small ? L1.bytes : L2.bytes;
          rs.range_flag = d_err.as<int>() + 1;
          if (direct)
            kern<<<g, kAev2Warps * 32, lb * kAev2Warps, st>>>(
                P, c->use_cuaev ? 1 : 0, d_x4.as<V4>(), d_beg.as<int64_t>(), d_cnt.as<int>(), jl(), d_order.as<int>(),
                nlocal, d_daev.as<T>(), ld_aev, d_force.as<T>(), want_virial ? d_virial.as<double>() : nullptr,
                (T)kHartree2KcalMol, cp, cpA, lb, rs);
          else
            kern_strided<<<g, kAev2Warps * 32, lb * kAev2Warps, st>>>(
                P, c->use_cuaev ? 1 : 0, d_x4.as<V4>(), d_beg.as<int64_t>(), d_cnt.as<int>(), jl(), d_order.as<int>(),
                nlocal, d_daev.as<T>(), ld_aev, d_force.as<T>(), want_virial ? d_virial.as<double>() : nullptr,
                (T)kHartree2KcalMol, cp, cpA, lb, rs);
          c->launches++;
        };
        // a segment: its rows with the small buffers, then the rows that did not fit (deferred to list `slot`)
        auto segment = [&](bool direct, int g, const int* sel_cnt, const int* sel_rows, int slot) {
          Aev2Rows rs;
          rs.sel_cnt = sel_cnt; rs.sel_rows = sel_rows;
          int* const ocnt = d_ovf.as<int>() + 3 + slot;
          int* const orow = d_ovf_rows.as<int>() + (size_t)(1 + slot) * nlocal;
          if (two_pass) { rs.ovf_cnt = ocnt; rs.ovf_rows = orow; }
          go(direct, g, two_pass, rs);
          if (two_pass) {
            Aev2Rows r2;
            r2.sel_cnt = ocnt; r2.sel_rows = orow;
            go(false, std::min(grid, 592), false, r2);
          }
        };
        if (overlap) {
          // boundary rows first: with them every ghost force is final and can travel while the interior rows are done
          if (rows_bnd > 0) segment(false, (rows_bnd + kAev2Warps - 1) / kAev2Warps, d_sel.as<int>() + 1, d_rows.as<int>() + rows_int, 0);
          axpy_kernel<T, double><<<(unsigned)(((int64_t)nghost * 3 + 255) / 256), 256, 0, st>>>(
              d_force.as<T>() + (size_t)nlocal * 3, (double*)f_dev + (size_t)nlocal * 3, (int64_t)nghost * 3);
          c->launches++;
          ANIB_CUDA_OK(cudaEventRecord(c->evh[2], st));
          ANIB_CUDA_OK(cudaStreamWaitEvent(cs, c->evh[2], 0));
          if (int rc = halo_reverse_exchange_on(*halo, (double*)f_dev, cs)) return rc;
          if (rows_int > 0) segment(false, (rows_int + kAev2Warps - 1) / kAev2Warps, d_sel.as<int>(), d_rows.as<int>(), 1);
        } else {
          segment(true, grid, nullptr, nullptr, 0);
        }
        return 0;
      };
      int rc;
      if (aev_gen == 2 && P.nZ == 4 && P.nA == 8) rc = launch2(aev2_backward_kernel<T, 4, 8>, aev2_backward_kernel<T, 4, 8, true>);
      else if (aev_gen == 2 && P.nZ == 8 && P.nA == 4) rc = launch2(aev2_backward_kernel<T, 8, 4>, aev2_backward_kernel<T, 8, 4, true>);
      else if (aev_gen == 2 && P.nZ == 8 && P.nA == 8) rc = launch2(aev2_backward_kernel<T, 8, 8>, aev2_backward_kernel<T, 8, 8, true>);
      else if (P.nZ == 4) rc = launch(aev_backward_kernel<T, 4>);
      else if (P.nZ == 8) rc = launch(aev_backward_kernel<T, 8>);
      else if (P.nZ == 1) rc = launch(aev_backward_kernel<T, 1>);
      else if (P.nZ == 2) rc = launch(aev_backward_kernel<T, 2>);
      else { set_error("unsupported number of ShfZ values (supported: 1, 2, 4, 8)"); return 1; }
      if (rc) return rc;
    }
    if (tm) cudaEventRecord(c->ev[4], st);
    // ---- (4) reductions and outputs
    if (want_e) {
      if (nlocal > 0) {
        int nb = std::min(1024, (nlocal + 255) / 256);
        reduce_sum_stage1<<<nb, 256, 0, st>>>(d_eatom.as<double>(), nlocal, d_partial.as<double>());
        reduce_sum_stage2<<<1, 256, 0, st>>>(d_partial.as<double>(), nb, d_scalars.as<double>(), kHartree2KcalMol);
        c->launches += 2;
        if (hm.use_repulsion) { add_scaled_scalar_kernel<<<1, 1, 0, st>>>(d_scalars.as<double>(), kHartree2KcalMol); c->launches++; }
      } else {
        ANIB_CUDA_OK(cudaMemsetAsync(d_scalars.p, 0, sizeof(double), st));
      }
    }
    const bool want_eatom = (out_eatom_host != nullptr) || (eatom_dev != nullptr);
    if (want_eatom && nlocal > 0) {
      scatter_eatom_kernel<<<(nlocal + 255) / 256, 256, 0, st>>>(d_eatom.as<double>(), d_order.as<int>(), nlocal,
                                                                kHartree2KcalMol,
                                                                eatom_dev ? eatom_dev : d_eatom_atom.as<double>(),
                                                                eatom_dev ? 1 : 0);
      c->launches++;
    }
    if (f_dev != nullptr && ntotal > 0) {
      const int64_t n3 = (int64_t)(overlap ? nlocal : ntotal) * 3;      // (overlapped exchange: the ghosts' share has left already)
      if (f_is_float) axpy_kernel<T, float><<<(unsigned)((n3 + 255) / 256), 256, 0, st>>>(d_force.as<T>(), (float*)f_dev, n3);
      else axpy_kernel<T, double><<<(unsigned)((n3 + 255) / 256), 256, 0, st>>>(d_force.as<T>(), (double*)f_dev, n3);
      c->launches++;
    }
    if (overlap) {
      // the ghost forces that came back are added once the local forces are in place, and the step's stream waits for that
      ANIB_CUDA_OK(cudaEventRecord(c->evh[3], st));
      ANIB_CUDA_OK(cudaStreamWaitEvent(cs, c->evh[3], 0));
      if (int rc = halo_reverse_unpack_on(*halo, (double*)f_dev, cs)) return rc;
      ANIB_CUDA_OK(cudaEventRecord(c->evh[4], cs));
      ANIB_CUDA_OK(cudaStreamWaitEvent(st, c->evh[4], 0));
    } else if (halo) {
      if (int rc = halo_reverse_exchange_on(*halo, (double*)f_dev, st)) return rc;
      if (int rc = halo_reverse_unpack_on(*halo, (double*)f_dev, st)) return rc;
    }
    // ---- D2H
    bool need_sync = false;
    if (want_e) { ANIB_CUDA_OK(cudaMemcpyAsync(hs, d_scalars.p, sizeof(double), cudaMemcpyDeviceToHost, st)); need_sync = true; }
    if (want_virial) {
      ANIB_CUDA_OK(cudaMemcpyAsync(hs + 4, d_virial.p, kVirialSlots * 9 * sizeof(double), cudaMemcpyDeviceToHost, st));
      need_sync = true;
    }
    ANIB_CUDA_OK(cudaMemcpyAsync(hs + 1, d_err.p, 2 * sizeof(int), cudaMemcpyDeviceToHost, st));
    ANIB_CUDA_OK(cudaMemcpyAsync(hs + 2, d_ovf.p, 4 * sizeof(int), cudaMemcpyDeviceToHost, st));
    stats_pending = true;
    f_direct = false;
    if (out_force_host != nullptr && ntotal > 0 && pin_f.ensure(out_force_host, (size_t)ntotal * 3 * sizeof(double), c->opt_pin_user != 0)) {
      // widen on the device, DMA straight into the caller's array
      const int64_t n3 = (int64_t)ntotal * 3;
      ANIB_RESERVE(d_fout, (size_t)n3 * sizeof(double));
      convert_kernel<T, double><<<(unsigned)((n3 + 255) / 256), 256, 0, st>>>(d_force.as<T>(), d_fout.as<double>(), n3);
      c->launches++;
      ANIB_CUDA_OK(cudaMemcpyAsync(out_force_host, d_fout.p, (size_t)n3 * sizeof(double), cudaMemcpyDeviceToHost, st));
      f_direct = true;
      need_sync = true;
    } else if (out_force_host != nullptr) {
      ANIB_RESERVE(h_f, (size_t)ntotal * 3 * sizeof(T));
      ANIB_CUDA_OK(cudaMemcpyAsync(h_f.p, d_force.p, (size_t)ntotal * 3 * sizeof(T), cudaMemcpyDeviceToHost, st));
      need_sync = true;
    }
    PinnedBuf* eat = nullptr;
    if (out_eatom_host != nullptr && nlocal > 0) {
      ANIB_RESERVE(h_eat, (size_t)nlocal * sizeof(double));
      ANIB_CUDA_OK(cudaMemcpyAsync(h_eat.p, d_eatom_atom.p, (size_t)nlocal * sizeof(double), cudaMemcpyDeviceToHost, st));
      eat = &h_eat;
      need_sync = true;
    }
    if (tm) { cudaEventRecord(c->ev[5], st); need_sync = true; }
    if (need_sync) {
      ANIB_CUDA_OK(cudaStreamSynchronize(st));
      if (int rc = check_errors()) return rc;
    }
    ANIB_CUDA_OK(cudaGetLastError());
    if (want_e) *out_energy = hs[0];
    if (want_virial) {
      double w[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
      for (int sl = 0; sl < kVirialSlots; sl++)
        for (int q = 0; q < 9; q++) w[q] += hs[4 + sl * 9 + q];
      for (int a = 0; a < 3; a++)
        for (int b = 0; b < 3; b++) out_virial9[3 * a + b] = 0.5 * (w[3 * a + b] + w[3 * b + a]);
    }
    if (out_force_host != nullptr && !f_direct) {
      const T* src = h_f.as<T>();
      for (size_t k = 0; k < (size_t)ntotal * 3; k++) out_force_host[k] = (double)src[k];
    }
    if (eat) memcpy(out_eatom_host, eat->p, (size_t)nlocal * sizeof(double));
    if (tm)
      for (int q = 0; q < 5; q++) cudaEventElapsedTime(&c->stage_ms[q], c->ev[q], c->ev[q + 1]);
    return 0;
  }
  PinnedBuf h_eat;

  // Device-side error flags of the last finished step (copied to pinned memory by every step): report once, then clear.
  int check_errors() override {
    if (!h_small.p) return 0;
    int* herr = (int*)(h_small.as<double>() + 1);
    const int e0 = herr[0], e1 = herr[1];
    if (!e0 && !e1) return 0;
    herr[0] = herr[1] = 0;
    cudaMemsetAsync(d_err.p, 0, 2 * sizeof(int), c->stream);
    if (e0) { set_error("an atom has more than " + std::to_string(capA) + " neighbours inside the angular cutoff"); return 2; }
    if (e1 & 1) {      // reported first: gradients that left the fp16 planes as inf also trip the fixed-point check downstream
      set_error("an MLP activation or gradient left the fp16 operand range of the f16x3 tensor-core path (|16 h| or |256 g| >= 65504): "
                "the result of this step is invalid; use ani_ctx_set_mlp_mode(ANI_MLP_TF32X3 or ANI_MLP_FP32_SIMT) for this system");
      return 3;
    }
    set_error("an angular force term left the fixed-point range of the fp32 AEV backward kernel (|term| >= 4096 Hartree/A or NaN: "
              "two atoms are nearly on top of each other); the result of this step is invalid");
    return 3;
  }

  // test hook: rows back in atom order and in the FULL AEV layout (dropped columns are zero)
  int get_aev(double* aev, double* daev) override {
    ANIB_CUDA_OK(cudaStreamSynchronize(c->stream));
    std::vector<T> tmp((size_t)nlocal * ld_aev);
    const int Lf = Pfull.aev_len;
    h_order.resize(nlocal);                   // (the device build leaves the row order on the device only)
    if (nlocal > 0) ANIB_CUDA_OK(cudaMemcpy(h_order.data(), d_order.p, (size_t)nlocal * sizeof(int), cudaMemcpyDeviceToHost));
    for (int pass = 0; pass < 2; pass++) {
      double* out = pass == 0 ? aev : daev;
      if (!out) continue;
      std::fill(out, out + (size_t)nlocal * Lf, 0.0);
      if (pass == 0 && aev_in_planes) {   // f16x3 mode keeps the AEV only as scaled fp16 hi/lo planes
        std::vector<__half> ph((size_t)nlocal * ld_planes), pl((size_t)nlocal * ld_planes);
        ANIB_CUDA_OK(cudaMemcpy(ph.data(), d_aev_hi.p, ph.size() * 2, cudaMemcpyDeviceToHost));
        ANIB_CUDA_OK(cudaMemcpy(pl.data(), d_aev_lo.p, pl.size() * 2, cudaMemcpyDeviceToHost));
        for (int pos = 0; pos < nlocal; pos++) {
          const int i = h_order[pos];
          for (int k = 0; k < P.aev_len; k++)
            out[(size_t)i * Lf + col_full[k]] = ((double)__half2float(ph[(size_t)pos * ld_planes + k]) +
                                                 (double)__half2float(pl[(size_t)pos * ld_planes + k])) / (double)f16::kScaleAev;
        }
        continue;
      }
      ANIB_CUDA_OK(cudaMemcpy(tmp.data(), pass == 0 ? d_aev.p : d_daev.p, tmp.size() * sizeof(T), cudaMemcpyDeviceToHost));
      for (int pos = 0; pos < nlocal; pos++) {
        const int i = h_order[pos];
        for (int k = 0; k < P.aev_len; k++) out[(size_t)i * Lf + col_full[k]] = (double)tmp[(size_t)pos * ld_aev + k];
      }
    }
    return 0;
  }
};

template <> int Engine<double>::upload_tensor_core_planes(const HostModel&) { return 0; }
template <> int Engine<float>::upload_tensor_core_planes(const HostModel& src) {
  int rc = 0;
  for (int s = 0; s < hm.S && !rc; s++) rc = tc_prepare_species(src, M, s, nets[s].tc);
  for (int s = 0; s < hm.S && !rc; s++) rc = f16::prepare_species(src, M, s, nets[s].hw);
  return rc;
}

template <> int Engine<double>::mlp_tc_species(int, int, int) {
  set_error("tensor-core MLP modes are single precision only");
  return 1;
}

template <> int Engine<double>::mlp_f16_species(int, int, int) {
  set_error("tensor-core MLP modes are single precision only");
  return 1;
}

template <> int Engine<float>::mlp_f16_species(int s, int row0, int R) {
  g_dbg_word_fn = f16::fused_dbg_word;
  // default: the whole network of a bucket in one kernel (mlp_f16_fused.cuh); option "mlp_fused" 0 = layer by layer (mlp_f16.cuh)
  if (c->opt_mlp_fused && f16::fused_supported(nets[s].dims)) {
    f16::FusedOptions fo;
    fo.flag = d_err.as<int>() + 1; fo.prof = c->opt_fused_prof; fo.cta_cap = c->opt_gemm_ctas;
    ANIB_CUDA_OK(cudaMemsetAsync(d_daev.as<float>() + (size_t)row0 * ld_aev, 0, (size_t)R * ld_aev * sizeof(float), c->stream));
    return f16::mlp_species_fused(c->stream, nets[s].hw, fwork, fo, M, nets[s].dims, (const __half*)d_aev_hi.p, (const __half*)d_aev_lo.p,
                                  ld_planes, nlocal, row0, R, d_daev.as<float>(), ld_aev, d_eatom.as<double>(), hm.sae[s], &c->launches);
  }
  return f16::mlp_species(c->stream, nets[s].hw, hwork, M, nets[s].dims, (const __half*)d_aev_hi.p, (const __half*)d_aev_lo.p,
                          ld_planes, nlocal, row0, R, d_daev.as<float>(), ld_aev, d_eatom.as<double>(), hm.sae[s], &c->launches, c->opt_gemm_ctas);
}

template <> int Engine<double>::mlp_f16_all() {
  set_error("tensor-core MLP modes are single precision only");
  return 1;
}

template <> int Engine<float>::mlp_f16_all() {
  for (int s = 0; s < hm.S; s++) {
    const int R = boff[s + 1] - boff[s];
    if (R == 0) continue;
    if (!nets[s].hw.ready) { set_error("this network's layer widths are not supported by the f16 tensor-core path (need multiples of 32, <= 256)"); return 1; }
    if (int rc = mlp_f16_species(s, boff[s], R)) return rc;
  }
  return 0;
}

template <> int Engine<float>::mlp_tc_species(int s, int row0, int R) {
  return tc_mlp_species(c->stream, c->mlp_mode == ANI_MLP_TF32X3, nets[s].tc, tcw, M, nets[s].dims, d_aev_hi.as<float>(),
                        d_aev_lo.as<float>(), ld_aev, nlocal, row0, R, d_daev.as<float>(), d_eatom.as<double>(), hm.sae[s],
                        &c->launches);
}

}  // namespace anib

// ================================================================================================== C ABI
extern "C" {

const char* ani_last_error(void) { return g_last_error.c_str(); }
int ani_abi_version(void) { return ANI_B200_ABI_VERSION; }

int ani_model_load(const char* path, ani_model** out) {
  if (!path || !out) { set_error("ani_model_load: null argument"); return 1; }
  std::unique_ptr<ani_model> m(new ani_model());
  std::string err;
  if (!load_aniw(path, m->hm, err)) { set_error(err); return 1; }
  *out = m.release();
  return 0;
}

void ani_model_free(ani_model* m) { delete m; }

int ani_model_info(const ani_model* m, int* num_species, int* aev_length, int* num_models, double* Rcr, double* Rca) {
  if (!m) { set_error("ani_model_info: null model"); return 1; }
  if (num_species) *num_species = m->hm.S;
  if (aev_length) *aev_length = m->hm.aev_len();
  if (num_models) *num_models = m->hm.M;
  if (Rcr) *Rcr = m->hm.Rcr;
  if (Rca) *Rca = m->hm.Rca;
  return 0;
}

int ani_ctx_create(const ani_model* model, int device, int use_num_models, int use_cuaev, int use_fullnbr, int use_single,
                   ani_ctx** out) {
  if (!model || !out) { set_error("ani_ctx_create: null argument"); return 1; }
  if (device < 0) {
    set_error("device 'cpu' is not available: libani_b200 is a CUDA (sm_100a) implementation without a CPU path");
    return 1;
  }
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    cudaGetLastError();
    set_error("no CUDA device available (libani_b200 has no CPU fallback)");
    return 1;
  }
  if (device >= ndev) { set_error("CUDA device ordinal out of range"); return 1; }
  ANIB_CUDA_OK(cudaSetDevice(device));
  int M = use_num_models;
  if (M == -1) M = model->hm.M;   // ani.cpp:75-77
  if (M < 1 || M > model->hm.M) { set_error("use_num_models must be -1 or in 1..num_models"); return 1; }
  std::unique_ptr<ani_ctx> c(new ani_ctx());
  c->device = device; c->use_single = use_single != 0; c->use_cuaev = use_cuaev != 0; c->use_fullnbr = use_fullnbr != 0;
  c->num_models = M; c->hm = &model->hm;
  ANIB_CUDA_OK(cudaStreamCreateWithFlags(&c->own_stream, cudaStreamNonBlocking));
  c->stream = c->own_stream;
  for (int q = 0; q < 6; q++) ANIB_CUDA_OK(cudaEventCreate(&c->ev[q]));
  int rc;
  if (c->use_single) {
    auto* e = new Engine<float>(c.get(), model->hm, M);
    c->eng.reset(e);
    rc = e->init();
  } else {
    auto* e = new Engine<double>(c.get(), model->hm, M);
    c->eng.reset(e);
    rc = e->init();
  }
  if (rc) return rc;
  c->opt_mlp_fused = f16::env_int("ANI_B200_FUSED", 1) != 0;
  c->opt_fused_prof = f16::env_int("ANI_B200_FUSED_PROF", 0) != 0;
  c->opt_gemm_ctas = std::max(0, f16::env_int("ANI_B200_GEMM_CTAS", 0));
  c->opt_aev_two_pass = f16::env_int("ANI_B200_AEV_TWO_PASS", 1) != 0;
  c->opt_pin_user = f16::env_int("ANI_B200_PIN_USER", 0) != 0;
  const char* mode = getenv("ANI_B200_MLP_MODE");
  if (c->use_single) {
    // default: the tensor-core path that keeps fp32 parity, when every network's widths fit it; FFMA otherwise
    c->mlp_mode = ANI_MLP_F16X3;
    for (int s = 0; s < model->hm.S; s++)
      if (!f16::supported_dims(model->hm.dims[s])) c->mlp_mode = ANI_MLP_TF32X3;
    if (c->mlp_mode == ANI_MLP_TF32X3)
      for (int s = 0; s < model->hm.S; s++)
        if (!tc::tc_supported_dims(model->hm.dims[s])) c->mlp_mode = ANI_MLP_FP32_SIMT;
    if (getenv("LAMMPS_ANI_ALLOW_TF32") && atoi(getenv("LAMMPS_ANI_ALLOW_TF32")) != 0) c->mlp_mode = ANI_MLP_TF32;  // ani.cpp:41-43
    if (mode) {
      if (!strcmp(mode, "simt")) c->mlp_mode = ANI_MLP_FP32_SIMT;
      else if (!strcmp(mode, "tf32x3")) c->mlp_mode = ANI_MLP_TF32X3;
      else if (!strcmp(mode, "tf32")) c->mlp_mode = ANI_MLP_TF32;
      else if (!strcmp(mode, "f16x3")) c->mlp_mode = ANI_MLP_F16X3;
    }
  }
  *out = c.release();
  return 0;
}

void ani_ctx_destroy(ani_ctx* c) {
  if (!c) return;
  cudaSetDevice(c->device);
  c->eng.reset();
  for (void* q : c->pinned_host) { cudaHostUnregister(q); cudaGetLastError(); }
  for (int q = 0; q < 5; q++) if (c->evh[q]) cudaEventDestroy(c->evh[q]);
  if (c->comm_stream) cudaStreamDestroy(c->comm_stream);
  if (c->d_word) cudaFree(c->d_word);
  if (c->h_word) cudaFreeHost(c->h_word);
  for (int q = 0; q < 6; q++) if (c->ev[q]) cudaEventDestroy(c->ev[q]);
  if (c->own_stream) cudaStreamDestroy(c->own_stream);
  delete c;
}

int ani_ctx_set_mlp_mode(ani_ctx* c, int mode) {
  if (!c) { set_error("null context"); return 1; }
  if (mode < 0 || mode > 3) { set_error("unknown MLP mode"); return 1; }
  if (!c->use_single && mode != ANI_MLP_FP32_SIMT) { set_error("tensor-core MLP modes are single precision only"); return 1; }
  c->mlp_mode = mode;
  return 0;
}

int ani_ctx_set_stream(ani_ctx* c, void* s) {
  if (!c) { set_error("null context"); return 1; }
  c->stream = (cudaStream_t)s;   // NULL is a real stream too: the legacy default stream (torch's default)
  return 0;
}

int ani_ctx_reset_stream(ani_ctx* c) {
  if (!c) { set_error("null context"); return 1; }
  c->stream = c->own_stream;
  return 0;
}

int ani_ctx_synchronize(ani_ctx* c) {
  if (!c) { set_error("null context"); return 1; }
  ANIB_CUDA_OK(cudaSetDevice(c->device));
  ANIB_CUDA_OK(cudaStreamSynchronize(c->stream));
  return c->eng ? c->eng->check_errors() : 0;
}

int ani_ctx_set_option(ani_ctx* c, const char* name, int value) {
  if (!c || !name) { set_error("ani_ctx_set_option: null argument"); return 1; }
  if (!strcmp(name, "mlp_fused")) c->opt_mlp_fused = value != 0;
  else if (!strcmp(name, "fused_prof")) c->opt_fused_prof = value != 0;
  else if (!strcmp(name, "gemm_ctas")) c->opt_gemm_ctas = value > 0 ? value : 0;
  else if (!strcmp(name, "halo_overlap")) c->opt_halo_overlap = value != 0;
  else if (!strcmp(name, "aev_two_pass")) c->opt_aev_two_pass = value != 0;
  else if (!strcmp(name, "pin_user")) { c->opt_pin_user = value != 0; if (!value && c->eng) c->eng->release_pins(); }
  else { set_error(std::string("ani_ctx_set_option: unknown option '") + name + "'"); return 1; }
  return 0;
}

int ani_ctx_enable_timing(ani_ctx* c, int on) {
  if (!c) { set_error("null context"); return 1; }
  c->timing = on != 0;
  return 0;
}

int ani_last_timing(ani_ctx* c, float stage_ms[5], int* num_launches) {
  if (!c) { set_error("null context"); return 1; }
  if (stage_ms) for (int q = 0; q < 5; q++) stage_ms[q] = c->stage_ms[q];
  if (num_launches) *num_launches = c->launches;
  return 0;
}

// Bring (possibly device-resident) small integer arrays to the host.
static int fetch_ints(std::vector<int>& dst, const int* src, size_t n, int on_device) {
  dst.resize(n);
  if (n == 0) return 0;
  if (on_device) ANIB_CUDA_OK(cudaMemcpy(dst.data(), src, n * sizeof(int), cudaMemcpyDeviceToHost));
  else memcpy(dst.data(), src, n * sizeof(int));
  return 0;
}

int ani_build_neighbors(ani_ctx* c, int nlocal, int ntotal, const int* species, const double* x, int x_on_device, double cutoff,
                        const double lo[3], const double hi[3], int64_t* npairs_out) {
  if (!c || !species || !lo || !hi || (ntotal > 0 && !x)) { set_error("ani_build_neighbors: null argument"); return 1; }
  if (!c->use_fullnbr) { set_error("context was created for a half neighbour list"); return 1; }
  if (nlocal < 0 || ntotal < nlocal) { set_error("ani_build_neighbors: inconsistent sizes"); return 1; }
  ANIB_CUDA_OK(cudaSetDevice(c->device));
  c->h_species.assign(species, species + ntotal);
  return c->eng->build_neighbors(nlocal, ntotal, c->h_species.data(), x, x_on_device != 0, cutoff, lo, hi, npairs_out);
}

int ani_ctx_release_host_buffers(ani_ctx* c) {
  if (!c) { set_error("null context"); return 1; }
  ANIB_CUDA_OK(cudaSetDevice(c->device));
  c->eng->release_pins();
  for (void* q : c->pinned_host) { cudaHostUnregister(q); cudaGetLastError(); }
  c->pinned_host.clear();
  return 0;
}

int ani_ctx_pin_host(ani_ctx* c, void* host, size_t bytes) {
  if (!c) { set_error("null context"); return 1; }
  if (!host || !bytes) return 0;
  ANIB_CUDA_OK(cudaSetDevice(c->device));
  const cudaError_t rc = cudaHostRegister(host, bytes, cudaHostRegisterDefault);
  if (rc == cudaSuccess) { c->pinned_host.push_back(host); return 0; }
  cudaGetLastError();
  if (rc == cudaErrorHostMemoryAlreadyRegistered) return 0;       // the caller (or an earlier call) pinned it: nothing to do
  set_error(std::string("ani_ctx_pin_host: cudaHostRegister failed: ") + cudaGetErrorString(rc));
  return 1;
}

int ani_debug_get_neighbors(ani_ctx* c, int* numneigh, int* jlist) {
  if (!c) { set_error("null context"); return 1; }
  ANIB_CUDA_OK(cudaSetDevice(c->device));
  return c->eng->get_neighbors(numneigh, jlist);
}

int ani_set_neighbors_full(ani_ctx* c, int nlocal, int ntotal, const int* species, int inum, const int* ilist,
                           const int* numneigh, const int* jlist, int64_t npairs, int on_device) {
  if (!c || !species || (inum > 0 && (!ilist || !numneigh)) || (npairs > 0 && !jlist)) { set_error("ani_set_neighbors_full: null argument"); return 1; }
  if (!c->use_fullnbr) { set_error("context was created for a half neighbour list"); return 1; }
  if (nlocal < 0 || ntotal < nlocal || inum < 0 || inum > nlocal) { set_error("ani_set_neighbors_full: inconsistent sizes"); return 1; }
  ANIB_CUDA_OK(cudaSetDevice(c->device));
  if (fetch_ints(c->h_species, species, ntotal, on_device)) return 1;
  std::vector<int> il, nn;
  if (fetch_ints(il, ilist, inum, on_device)) return 1;
  if (fetch_ints(nn, numneigh, nlocal, on_device)) return 1;
  c->h_beg.assign(std::max(nlocal, 1), 0);
  c->h_cnt.assign(std::max(nlocal, 1), 0);
  int64_t run = 0;
  for (int ii = 0; ii < inum; ii++) {
    const int i = il[ii];
    if (i < 0 || i >= nlocal) { set_error("ilist entry outside the local atoms"); return 1; }
    c->h_beg[i] = run;
    c->h_cnt[i] = nn[i];
    run += nn[i];
  }
  if (run != npairs) { set_error("npairs does not match the sum of numneigh over ilist"); return 1; }
  return c->eng->set_neighbors(nlocal, ntotal, c->h_species.data(), c->h_beg.data(), c->h_cnt.data(), jlist, npairs,
                               on_device != 0);
}

int ani_device_count(int* out) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); n = 0; }
  if (out) *out = n;
  if (n == 0) { set_error("no CUDA device available (libani_b200 has no CPU fallback)"); return 1; }
  return 0;
}

int ani_set_neighbors_table(ani_ctx* c, int nlocal, int ntotal, const int* types, int type_offset, int inum, const int* ilist,
                            const int* numneigh, const int* table, int64_t row_stride, int64_t col_stride, int neighmask,
                            int on_device) {
  if (!c || !types || (inum > 0 && (!ilist || !numneigh || !table))) { set_error("ani_set_neighbors_table: null argument"); return 1; }
  if (!c->use_fullnbr) { set_error("context was created for a half neighbour list"); return 1; }
  if (nlocal < 0 || ntotal < nlocal || inum < 0 || inum > nlocal) { set_error("ani_set_neighbors_table: inconsistent sizes"); return 1; }
  ANIB_CUDA_OK(cudaSetDevice(c->device));
  if (fetch_ints(c->h_species, types, ntotal, on_device)) return 1;
  for (int i = 0; i < ntotal; i++) c->h_species[i] -= type_offset;
  std::vector<int> il, nn;
  if (fetch_ints(il, ilist, inum, on_device)) return 1;
  if (fetch_ints(nn, numneigh, nlocal, on_device)) return 1;
  c->h_beg.assign(std::max(nlocal, 1), 0);
  c->h_cnt.assign(std::max(nlocal, 1), 0);
  int64_t run = 0;
  int maxn = 0;
  for (int ii = 0; ii < inum; ii++) {
    const int i = il[ii];
    if (i < 0 || i >= nlocal) { set_error("ilist entry outside the local atoms"); return 1; }
    c->h_beg[i] = run;
    c->h_cnt[i] = nn[i];
    run += nn[i];
    maxn = std::max(maxn, nn[i]);
  }
  // de-pad on the device: table (device, or uploaded once) -> flat list in the context's scratch
  int *d_tab = nullptr, *d_flat = nullptr, *d_cnt = nullptr;
  int64_t* d_beg = nullptr;
  const int* tab = table;
  const size_t tab_elems = (size_t)((int64_t)(nlocal > 0 ? nlocal - 1 : 0) * row_stride + (int64_t)(maxn > 0 ? maxn - 1 : 0) * col_stride + 1);
  int rc = 0;
  auto cleanup = [&]() { cudaFree(d_tab); cudaFree(d_flat); cudaFree(d_cnt); cudaFree(d_beg); };
#define TAB_OK(expr) do { if ((expr) != cudaSuccess) { set_error(std::string(#expr) + " failed: " + cudaGetErrorString(cudaGetLastError())); cleanup(); return 1; } } while (0)
  if (!on_device && run > 0) {
    TAB_OK(cudaMalloc(&d_tab, tab_elems * sizeof(int)));
    TAB_OK(cudaMemcpy(d_tab, table, tab_elems * sizeof(int), cudaMemcpyHostToDevice));
    tab = d_tab;
  }
  TAB_OK(cudaMalloc(&d_flat, std::max<int64_t>(run, 1) * sizeof(int)));
  TAB_OK(cudaMalloc(&d_cnt, std::max(nlocal, 1) * sizeof(int)));
  TAB_OK(cudaMalloc(&d_beg, std::max(nlocal, 1) * sizeof(int64_t)));
  TAB_OK(cudaMemcpy(d_cnt, c->h_cnt.data(), std::max(nlocal, 1) * sizeof(int), cudaMemcpyHostToDevice));
  TAB_OK(cudaMemcpy(d_beg, c->h_beg.data(), std::max(nlocal, 1) * sizeof(int64_t), cudaMemcpyHostToDevice));
  if (run > 0) {
    depad_neighbors_kernel<<<(unsigned)(((int64_t)nlocal * 32 + 255) / 256), 256, 0, c->stream>>>(tab, row_stride, col_stride, d_beg,
                                                                                                 d_cnt, nlocal, neighmask, d_flat);
    TAB_OK(cudaStreamSynchronize(c->stream));
  }
#undef TAB_OK
  rc = c->eng->set_neighbors(nlocal, ntotal, c->h_species.data(), c->h_beg.data(), c->h_cnt.data(), d_flat, run, true);
  cleanup();
  return rc;
}

int ani_set_neighbors_half(ani_ctx* c, int nlocal, int ntotal, const int* species, int64_t npairs,
                           const int64_t* atom_index12, int on_device) {
  if (!c || !species || (npairs > 0 && !atom_index12)) { set_error("ani_set_neighbors_half: null argument"); return 1; }
  if (c->use_fullnbr) { set_error("context was created for a full neighbour list"); return 1; }
  if (nlocal < 0 || ntotal < nlocal) { set_error("ani_set_neighbors_half: inconsistent sizes"); return 1; }
  ANIB_CUDA_OK(cudaSetDevice(c->device));
  if (fetch_ints(c->h_species, species, ntotal, on_device)) return 1;
  std::vector<int64_t> idx((size_t)2 * npairs);
  if (npairs > 0) {
    if (on_device) ANIB_CUDA_OK(cudaMemcpy(idx.data(), atom_index12, idx.size() * sizeof(int64_t), cudaMemcpyDeviceToHost));
    else memcpy(idx.data(), atom_index12, idx.size() * sizeof(int64_t));
  }
  // each stored pair (i,j) is a neighbour of i and, when j is local, of j: per-centre CSR over local atoms
  const int64_t* I = idx.data();
  const int64_t* J = idx.data() + npairs;
  c->h_cnt.assign(std::max(nlocal, 1), 0);
  for (int64_t p = 0; p < npairs; p++) {
    if (I[p] < 0 || I[p] >= ntotal || J[p] < 0 || J[p] >= ntotal) { set_error("atom_index12 entry out of range"); return 1; }
    if (I[p] < nlocal) c->h_cnt[I[p]]++;
    if (J[p] < nlocal) c->h_cnt[J[p]]++;
  }
  c->h_beg.assign(std::max(nlocal, 1), 0);
  int64_t run = 0;
  for (int i = 0; i < nlocal; i++) { c->h_beg[i] = run; run += c->h_cnt[i]; }
  c->h_jl.resize(std::max<int64_t>(run, 1));
  std::vector<int64_t> cur(c->h_beg.begin(), c->h_beg.end());
  for (int64_t p = 0; p < npairs; p++) {
    if (I[p] < nlocal) c->h_jl[cur[I[p]]++] = (int)J[p];
    if (J[p] < nlocal) c->h_jl[cur[J[p]]++] = (int)I[p];
  }
  return c->eng->set_neighbors(nlocal, ntotal, c->h_species.data(), c->h_beg.data(), c->h_cnt.data(), c->h_jl.data(), run,
                               false);
}

int ani_compute(ani_ctx* c, const double* x, double* out_energy, double* out_force, double* out_eatom, double* out_virial9) {
  if (!c || !x) { set_error("ani_compute: null argument"); return 1; }
  ANIB_CUDA_OK(cudaSetDevice(c->device));
  return c->eng->compute(x, false, false, out_energy, out_force, nullptr, false, out_eatom, nullptr, out_virial9);
}

int ani_compute_device(ani_ctx* c, const void* x_dev, int x_is_float, void* f_dev, int f_is_float, double* eatom_dev,
                       double* out_energy, double* out_virial9) {
  if (!c || !x_dev) { set_error("ani_compute_device: null argument"); return 1; }
  ANIB_CUDA_OK(cudaSetDevice(c->device));
  return c->eng->compute(x_dev, true, x_is_float != 0, out_energy, nullptr, f_dev, f_is_float != 0, nullptr, eatom_dev,
                         out_virial9);
}

int ani_debug_get_aev(ani_ctx* c, double* aev, double* daev) {
  if (!c) { set_error("null context"); return 1; }
  ANIB_CUDA_OK(cudaSetDevice(c->device));
  return c->eng->get_aev(aev, daev);
}

// test hook for the f16x3 kernel: C = A . B^T through EPI_D (fp32 rows out)
static int debug_f16_gemm(int device, int R, int N, int K, const float* A, const float* B, float* C) {
  using namespace anib::f16;
  if (K % 16 != 0 || N % 4 != 0) { set_error("ani_debug_tc_gemm(f16x3): K must be a multiple of 16 and N of 4"); return 1; }
  ANIB_CUDA_OK(cudaSetDevice(device));
  std::vector<float> av(A, A + (size_t)R * K), bv(B, B + (size_t)N * K);
  HalfPlanes a, b;
  const float sa = pick_scale(av) / 64.f, sb = pick_scale(bv);
  if (upload_half_planes(av, (size_t)K, sa, a) || upload_half_planes(bv, (size_t)K, sb, b)) return 1;
  anib::DevBufTc c_d;
  if (!c_d.reserve((size_t)R * N * 4)) { set_error("out of memory"); return 1; }
  int num_sms = 0, max_smem = 0;
  cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, device);
  cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, device);
  Launcher L{nullptr, num_sms, nullptr, max_smem};
  Args g{};
  g.BN = (N <= 256 && N % 32 == 0) ? N : 256;
  g.R = R; g.N = N; g.num_n_tiles = (N + g.BN - 1) / g.BN; g.num_row_tiles = (R + BM - 1) / BM; g.K = K; g.batch = 1;
  g.acc_scale = 1.0f / (sa * sb);
  const Operand none{nullptr, nullptr, 0, 0, 0};
  Operand Ao{a.h(), a.l(), (uint64_t)R, (uint64_t)K, (uint64_t)pad64(K)}, Bo{b.h(), b.l(), (uint64_t)N, (uint64_t)K, (uint64_t)pad64(K)};
  if (launch_gemm<EPI_D>(L, Ao, Bo, none, c_d.f(), (uint64_t)R, (uint64_t)N, (uint64_t)N, none, g)) return 1;
  ANIB_CUDA_OK(cudaDeviceSynchronize());
  ANIB_CUDA_OK(cudaMemcpy(C, c_d.p, (size_t)R * N * 4, cudaMemcpyDeviceToHost));
  return 0;
}

int ani_debug_tc_gemm(int device, int R, int N, int K, const float* A, const float* B, float* C, int three_pass) {
  using namespace anib::tc;
  if (three_pass == 2) {
    if (!A || !B || !C || R < 1 || N < 16 || K < 16) { set_error("ani_debug_tc_gemm: bad arguments"); return 1; }
    return debug_f16_gemm(device, R, N, K, A, B, C);
  }
  if (!A || !B || !C || R < 1 || N < 16 || K < 8 || K % 8 != 0 || pick_bn(N) == 0) { set_error("ani_debug_tc_gemm: bad arguments"); return 1; }
  ANIB_CUDA_OK(cudaSetDevice(device));
  DevBufTc a_hi, a_lo, b_hi, b_lo, c_d;
  std::vector<float> av(A, A + (size_t)R * K), bv(B, B + (size_t)N * K);
  if (upload_planes(av, a_hi, a_lo) || upload_planes(bv, b_hi, b_lo)) return 1;
  if (!c_d.reserve((size_t)R * N * 4)) { set_error("out of memory"); return 1; }
  int num_sms = 0, max_smem = 0;
  cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, device);
  cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, device);
  Launcher L{nullptr, num_sms, nullptr, max_smem};
  GemmArgs g{};
  g.BN = pick_bn(N); g.R = R; g.N = N; g.num_n_tiles = N / g.BN; g.num_row_tiles = (R + BM - 1) / BM; g.K = K; g.batch = 1;
  g.three_pass = three_pass ? 1 : 0; g.out_hi = c_d.f(); g.ld_out = N; g.scale = 1.0f;
  if (launch_gemm<EPI_D>(L, a_hi.f(), a_lo.f(), (uint64_t)R, (uint64_t)K, (uint64_t)K, b_hi.f(), b_lo.f(), (uint64_t)N, (uint64_t)K,
                         (uint64_t)K, g))
    return 1;
  ANIB_CUDA_OK(cudaDeviceSynchronize());
  ANIB_CUDA_OK(cudaMemcpy(C, c_d.p, (size_t)R * N * 4, cudaMemcpyDeviceToHost));
  return 0;
}

// ---- plain device-memory helpers for callers that do not link the CUDA runtime themselves (the LAMMPS plugin)
int ani_dev_alloc(ani_ctx* c, size_t bytes, void** out) {
  if (!c || !out) { set_error("ani_dev_alloc: null argument"); return 1; }
  ANIB_CUDA_OK(cudaSetDevice(c->device));
  *out = nullptr;
  ANIB_CUDA_OK(cudaMalloc(out, std::max<size_t>(bytes, 16)));
  return 0;
}
int ani_dev_free(ani_ctx* c, void* p) {
  if (!c) { set_error("ani_dev_free: null context"); return 1; }
  ANIB_CUDA_OK(cudaSetDevice(c->device));
  if (p) { ANIB_CUDA_OK(cudaStreamSynchronize(c->stream)); ANIB_CUDA_OK(cudaFree(p)); }
  return 0;
}
int ani_dev_upload(ani_ctx* c, void* dst_dev, const void* src_host, size_t bytes) {      // ordered on the context's stream, returns when done
  if (!c || (bytes && (!dst_dev || !src_host))) { set_error("ani_dev_upload: null argument"); return 1; }
  if (!bytes) return 0;
  ANIB_CUDA_OK(cudaSetDevice(c->device));
  ANIB_CUDA_OK(cudaMemcpyAsync(dst_dev, src_host, bytes, cudaMemcpyHostToDevice, c->stream));
  ANIB_CUDA_OK(cudaStreamSynchronize(c->stream));
  return 0;
}
int ani_dev_download(ani_ctx* c, void* dst_host, const void* src_dev, size_t bytes) {
  if (!c || (bytes && (!dst_host || !src_dev))) { set_error("ani_dev_download: null argument"); return 1; }
  if (!bytes) return 0;
  ANIB_CUDA_OK(cudaSetDevice(c->device));
  ANIB_CUDA_OK(cudaMemcpyAsync(dst_host, src_dev, bytes, cudaMemcpyDeviceToHost, c->stream));
  ANIB_CUDA_OK(cudaStreamSynchronize(c->stream));
  return c->eng ? c->eng->check_errors() : 0;
}
int ani_dev_zero(ani_ctx* c, void* dst_dev, size_t bytes) {
  if (!c || (bytes && !dst_dev)) { set_error("ani_dev_zero: null argument"); return 1; }
  if (!bytes) return 0;
  ANIB_CUDA_OK(cudaSetDevice(c->device));
  ANIB_CUDA_OK(cudaMemsetAsync(dst_dev, 0, bytes, c->stream));
  return 0;
}

// ---- device-resident time integration (md.cuh)
int ani_md_nve_initial(ani_ctx* c, double* x_dev, double* v_dev, const double* f_dev, const double* mass_dev, int n, double dtv,
                       double dtf) {
  if (!c || (n > 0 && (!x_dev || !v_dev || !f_dev || !mass_dev))) { set_error("ani_md_nve_initial: null argument"); return 1; }
  if (n <= 0) return 0;
  ANIB_CUDA_OK(cudaSetDevice(c->device));
  md_nve_initial_kernel<<<(n + 255) / 256, 256, 0, c->stream>>>(x_dev, v_dev, f_dev, mass_dev, n, dtv, dtf);
  ANIB_CUDA_OK(cudaGetLastError());
  return 0;
}

int ani_md_nve_final(ani_ctx* c, double* v_dev, const double* f_dev, const double* mass_dev, int n, double dtf) {
  if (!c || (n > 0 && (!v_dev || !f_dev || !mass_dev))) { set_error("ani_md_nve_final: null argument"); return 1; }
  if (n <= 0) return 0;
  ANIB_CUDA_OK(cudaSetDevice(c->device));
  md_nve_final_kernel<<<(n + 255) / 256, 256, 0, c->stream>>>(v_dev, f_dev, mass_dev, n, dtf);
  ANIB_CUDA_OK(cudaGetLastError());
  return 0;
}

int ani_md_langevin(ani_ctx* c, const double* v_dev, double* f_dev, const double* mass_dev, const int64_t* tag_dev, int n, double g1,
                    double g2, uint64_t seed, int64_t step) {
  if (!c || (n > 0 && (!v_dev || !f_dev || !mass_dev))) { set_error("ani_md_langevin: null argument"); return 1; }
  if (n <= 0) return 0;
  ANIB_CUDA_OK(cudaSetDevice(c->device));
  md_langevin_kernel<<<(n + 255) / 256, 256, 0, c->stream>>>(v_dev, f_dev, mass_dev, (const long long*)tag_dev, n, g1, g2,
                                                           (unsigned long long)seed, (unsigned long long)step);
  ANIB_CUDA_OK(cudaGetLastError());
  return 0;
}

int ani_md_sum_mv2(ani_ctx* c, const double* v_dev, const double* mass_dev, int n, double* out_host) {
  if (!c || !out_host || (n > 0 && (!v_dev || !mass_dev))) { set_error("ani_md_sum_mv2: null argument"); return 1; }
  *out_host = 0.0;
  if (n <= 0) return 0;
  ANIB_CUDA_OK(cudaSetDevice(c->device));
  const int nb = std::min(1024, (n + 255) / 256);
  double* part = nullptr;
  ANIB_CUDA_OK(cudaMallocAsync((void**)&part, (size_t)nb * sizeof(double), c->stream));
  md_mv2_kernel<<<nb, 256, 0, c->stream>>>(v_dev, mass_dev, n, part);
  std::vector<double> h(nb);
  ANIB_CUDA_OK(cudaMemcpyAsync(h.data(), part, (size_t)nb * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  ANIB_CUDA_OK(cudaFreeAsync(part, c->stream));
  ANIB_CUDA_OK(cudaStreamSynchronize(c->stream));
  for (double q : h) *out_host += q;
  return 0;
}

int ani_md_max_disp2(ani_ctx* c, const double* x_dev, const double* x0_dev, int n, double* out_host) {
  if (!c || !out_host || (n > 0 && (!x_dev || !x0_dev))) { set_error("ani_md_max_disp2: null argument"); return 1; }
  *out_host = 0.0;
  if (n <= 0) return 0;
  ANIB_CUDA_OK(cudaSetDevice(c->device));
  if (!c->d_word) {
    ANIB_CUDA_OK(cudaMalloc((void**)&c->d_word, sizeof(unsigned long long)));
    ANIB_CUDA_OK(cudaMallocHost((void**)&c->h_word, sizeof(unsigned long long)));
  }
  ANIB_CUDA_OK(cudaMemsetAsync(c->d_word, 0, sizeof(unsigned long long), c->stream));
  md_max_disp2_kernel<<<std::min(592, (n + 255) / 256), 256, 0, c->stream>>>(x_dev, x0_dev, n, c->d_word);
  ANIB_CUDA_OK(cudaMemcpyAsync(c->h_word, c->d_word, sizeof(unsigned long long), cudaMemcpyDeviceToHost, c->stream));
  ANIB_CUDA_OK(cudaStreamSynchronize(c->stream));
  memcpy(out_host, c->h_word, sizeof(double));
  return c->eng ? c->eng->check_errors() : 0;
}

// Allocation-free variant of ani_md_sum_mv2 (no counterpart in the reference; added for
// callers that cannot allocate: CUDA graph capture, tight serving loops).  No cudaMalloc / cudaFree / host container: the
// partials go to the file-scope array g_mv2_static_scratch (md.cuh), whose DEVICE address is looked up once per device with
// cudaGetSymbolAddress (in host code the array's name is only a host-side stand-in), and the sum lands in out_dev[0] in stream
// order; nothing synchronises.  All calls on a device share the one array: calls that overlap on the GPU (different streams or
// contexts) race.  Nothing needs zeroing: a call overwrites every partial it later reads.
int ani_md_sum_mv2_static(ani_ctx* c, const double* v_dev, const double* mass_dev, int n, double* out_dev) {
  if (!c || !out_dev || (n > 0 && (!v_dev || !mass_dev))) { set_error("ani_md_sum_mv2_static: null argument"); return 1; }
  if (n < 0 || n > kMv2StaticMaxAtoms) {
    set_error("ani_md_sum_mv2_static: n = " + std::to_string(n) + " is outside 0 .. " + std::to_string(kMv2StaticMaxAtoms) +
              " (the limit of the static scratch variant; use ani_md_sum_mv2)");
    return 1;
  }
  ANIB_CUDA_OK(cudaSetDevice(c->device));
  constexpr int kMaxDev = 64;
  static double* scratch_of[kMaxDev] = {};              // device address of the array on each GPU, filled on first use
  static std::mutex mu;
  if (c->device < 0 || c->device >= kMaxDev) { set_error("ani_md_sum_mv2_static: device index out of range"); return 1; }
  double* part = nullptr;
  {
    std::lock_guard<std::mutex> lk(mu);
    if (!scratch_of[c->device]) {
      void* p = nullptr;
      ANIB_CUDA_OK(cudaGetSymbolAddress(&p, g_mv2_static_scratch));     // with c->device current: the address is per device
      scratch_of[c->device] = (double*)p;
    }
    part = scratch_of[c->device];
  }
  const int nb = std::max(1, std::min(kMv2StaticBlocks, (n + 255) / 256));
  md_mv2_kernel<<<nb, 256, 0, c->stream>>>(v_dev, mass_dev, n, part);     // n == 0: one block writes partial[0] = 0
  md_mv2_final_kernel<<<1, 256, 0, c->stream>>>(part, nb, out_dev);
  ANIB_CUDA_OK(cudaGetLastError());
  return 0;
}

int ani_md_ghost_x(ani_ctx* c, double* x_dev, const int* owner_dev, const double* shift_dev, int nlocal, int nghost) {
  if (!c || (nghost > 0 && (!x_dev || !owner_dev || !shift_dev))) { set_error("ani_md_ghost_x: null argument"); return 1; }
  if (nghost <= 0) return 0;
  ANIB_CUDA_OK(cudaSetDevice(c->device));
  md_ghost_x_kernel<<<(unsigned)(((int64_t)nghost * 3 + 255) / 256), 256, 0, c->stream>>>(x_dev, owner_dev, shift_dev, nlocal, nghost);
  ANIB_CUDA_OK(cudaGetLastError());
  return 0;
}

int ani_md_fold_ghost_f(ani_ctx* c, double* f_dev, const int* owner_dev, int nlocal, int nghost) {
  if (!c || (nghost > 0 && (!f_dev || !owner_dev))) { set_error("ani_md_fold_ghost_f: null argument"); return 1; }
  if (nghost <= 0) return 0;
  ANIB_CUDA_OK(cudaSetDevice(c->device));
  md_fold_ghost_kernel<<<(unsigned)(((int64_t)nghost * 3 + 255) / 256), 256, 0, c->stream>>>(f_dev, owner_dev, nlocal, nghost);
  ANIB_CUDA_OK(cudaGetLastError());
  return 0;
}

// ---- ghost exchange inside the library (halo.cuh)
int ani_halo_unique_id(char out[128]) {
  if (!out) { set_error("ani_halo_unique_id: null argument"); return 1; }
  NcclApi& n = nccl_api();
  if (!n.ok) { set_error("libnccl.so.2 could not be loaded"); return 1; }
  ncclUniqueId id;
  ANIB_NCCL_OK(n.GetUniqueId(&id));
  memcpy(out, id.internal, 128);
  return 0;
}

int ani_halo_plan_create(ani_ctx* c, const char unique_id[128], int nranks, int rank, int nlocal, const int64_t* send_index,
                         const double* send_shift, const int* send_counts, const int* recv_counts, ani_halo_plan** out) {
  if (!c || !out || !send_counts || !recv_counts || nranks < 1 || rank < 0 || rank >= nranks) { set_error("ani_halo_plan_create: bad argument"); return 1; }
  ANIB_CUDA_OK(cudaSetDevice(c->device));
  std::unique_ptr<ani_halo_plan> hp(new ani_halo_plan());
  HaloPlan& p = hp->p;
  hp->ctx = c;
  p.device = c->device; p.nranks = nranks; p.rank = rank; p.nlocal = nlocal; p.stream = c->stream;
  p.send_off.assign(nranks + 1, 0); p.recv_off.assign(nranks + 1, 0);
  for (int r = 0; r < nranks; r++) {
    if (send_counts[r] < 0 || recv_counts[r] < 0) { set_error("ani_halo_plan_create: negative count"); return 1; }
    p.send_off[r + 1] = p.send_off[r] + send_counts[r];
    p.recv_off[r + 1] = p.recv_off[r] + recv_counts[r];
  }
  p.nsend = p.send_off[nranks]; p.nghost = p.recv_off[nranks];
  if (p.nsend > 0 && (!send_index || !send_shift)) { set_error("ani_halo_plan_create: null send arrays"); return 1; }
  if (send_counts[rank] != recv_counts[rank]) { set_error("ani_halo_plan_create: self send / receive counts differ"); return 1; }
  const size_t ns = (size_t)std::max<int64_t>(p.nsend, 1);
  ANIB_CUDA_OK(cudaMalloc((void**)&p.d_send_index, ns * sizeof(int64_t)));
  ANIB_CUDA_OK(cudaMalloc((void**)&p.d_send_shift, ns * 3 * sizeof(double)));
  ANIB_CUDA_OK(cudaMalloc((void**)&p.d_send_buf, ns * 3 * sizeof(double)));
  if (p.nsend > 0) {
    ANIB_CUDA_OK(cudaMemcpy(p.d_send_index, send_index, (size_t)p.nsend * sizeof(int64_t), cudaMemcpyHostToDevice));
    ANIB_CUDA_OK(cudaMemcpy(p.d_send_shift, send_shift, (size_t)p.nsend * 3 * sizeof(double), cudaMemcpyHostToDevice));
  }
  if (nranks > 1) {
    NcclApi& n = nccl_api();
    if (!n.ok) { set_error("libnccl.so.2 could not be loaded"); return 1; }
    if (!unique_id) { set_error("ani_halo_plan_create: a ncclUniqueId is needed for more than one rank"); return 1; }
    ncclUniqueId id;
    memcpy(id.internal, unique_id, 128);
    ANIB_NCCL_OK(n.CommInitRank(&p.comm, nranks, id, rank));
  }
  *out = hp.release();
  return 0;
}

void ani_halo_plan_destroy(ani_halo_plan* hp) {
  if (!hp) return;
  HaloPlan& p = hp->p;
  cudaSetDevice(p.device);
  cudaStreamSynchronize(hp->ctx->stream);
  if (p.comm && nccl_api().ok) nccl_api().CommDestroy(p.comm);
  cudaFree(p.d_send_index); cudaFree(p.d_send_shift); cudaFree(p.d_send_buf);
  delete hp;
}

// ghost positions: x[nlocal + ...] <- owners' x (+ periodic shift), blocks ordered by source rank
int ani_halo_forward(ani_halo_plan* hp, double* x_dev) {
  if (!hp || !x_dev) { set_error("ani_halo_forward: null argument"); return 1; }
  ANIB_CUDA_OK(cudaSetDevice(hp->p.device));
  return halo_forward_on(hp->p, x_dev, hp->ctx->stream);
}

// ghost forces: f[owner] += f[ghost]   (newton off: the pair style folds its own ghost forces, pair_ani.cpp:197-201)
int ani_halo_reverse(ani_halo_plan* hp, double* f_dev) {
  if (!hp || !f_dev) { set_error("ani_halo_reverse: null argument"); return 1; }
  ANIB_CUDA_OK(cudaSetDevice(hp->p.device));
  if (int rc = halo_reverse_exchange_on(hp->p, f_dev, hp->ctx->stream)) return rc;
  return halo_reverse_unpack_on(hp->p, f_dev, hp->ctx->stream);
}

// One domain-decomposed step: ghost positions from their owners, energy / forces, ghost forces back to their owners --
// ani_halo_forward + ani_compute_device + ani_halo_reverse in one call, which lets the library run the exchange on a second
// stream under the interior atoms' share of the AEV kernels (option "halo_overlap", default 1; needs a device-built list).
int ani_compute_device_halo(ani_ctx* c, ani_halo_plan* hp, double* x_dev, double* f_dev, double* eatom_dev, double* out_energy,
                            double* out_virial9) {
  if (!c || !hp || !x_dev || !f_dev) { set_error("ani_compute_device_halo: null argument"); return 1; }
  if (hp->ctx != c) { set_error("ani_compute_device_halo: the plan belongs to another context"); return 1; }
  ANIB_CUDA_OK(cudaSetDevice(c->device));
  return c->eng->compute(x_dev, true, false, out_energy, nullptr, f_dev, false, nullptr, eatom_dev, out_virial9, &hp->p);
}

int ani_halo_pack_x(ani_ctx* c, const void* x_dev, int is_float, const int64_t* send_index_dev, const void* shift_dev,
                    int64_t n, void* out_dev) {
  if (!c) { set_error("null context"); return 1; }
  if (n <= 0) return 0;
  ANIB_CUDA_OK(cudaSetDevice(c->device));
  unsigned g = (unsigned)((n * 3 + 255) / 256);
  if (is_float) halo_pack_kernel<float><<<g, 256, 0, c->stream>>>((const float*)x_dev, send_index_dev, (const float*)shift_dev, n, (float*)out_dev);
  else halo_pack_kernel<double><<<g, 256, 0, c->stream>>>((const double*)x_dev, send_index_dev, (const double*)shift_dev, n, (double*)out_dev);
  ANIB_CUDA_OK(cudaGetLastError());
  return 0;
}

int ani_halo_unpack_add_f(ani_ctx* c, void* f_dev, int is_float, const int64_t* index_dev, const void* buf_dev, int64_t n) {
  if (!c) { set_error("null context"); return 1; }
  if (n <= 0) return 0;
  ANIB_CUDA_OK(cudaSetDevice(c->device));
  unsigned g = (unsigned)((n * 3 + 255) / 256);
  if (is_float) halo_unpack_add_kernel<float><<<g, 256, 0, c->stream>>>((float*)f_dev, index_dev, (const float*)buf_dev, n);
  else halo_unpack_add_kernel<double><<<g, 256, 0, c->stream>>>((double*)f_dev, index_dev, (const double*)buf_dev, n);
  ANIB_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // extern "C"
