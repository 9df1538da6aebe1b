/* ani_b200.h -- C ABI of libani_b200.so: B200 (sm_100a) ANI energy / force evaluation.
 *
 * This is the drop-in boundary for the hot path of roitberg-group/lammps-ani.  Every entry point replaces
 * one piece of the reference's libtorch runner `class ANI` (src/ani_csrc/ani.h:11-84, ani.cpp) as it is
 * called from the LAMMPS pair styles (src/pair_ani.cpp:66-233, src/pair_ani_kokkos.cpp:67-214).  Plain
 * pointers and sizes only: no torch, Kokkos or LAMMPS types.  All functions return 0 on success and a
 * non-zero code on failure; ani_last_error() then describes it (thread-local).  No exceptions cross the ABI.
 * There is NO CPU fallback: ani_ctx_create fails when no CUDA device is usable.
 *
 * Units: inputs in Angstrom, outputs in kcal/mol and kcal/mol/Angstrom (model values x 627.5094738898777,
 * reference ani.h:9, ani.cpp:161-162,246-247).
 */
#ifndef ANI_B200_H
#define ANI_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ANI_B200_ABI_VERSION 1
#define ANI_HARTREE2KCALMOL 627.5094738898777

typedef struct ani_model ani_model; /* host-side parameters (one `.aniw` file)                    */
typedef struct ani_ctx ani_ctx;     /* one device context: weights in HBM, work buffers, a stream */

/* MLP arithmetic of the single-precision mode (ani_ctx_set_mlp_mode).  The default reproduces fp32. */
enum ani_mlp_mode {
  ANI_MLP_FP32_SIMT = 0, /* FFMA GEMMs, fp32 accumulate: the strict reference-precision path (and fp64 mode)   */
  ANI_MLP_TF32X3 = 1,    /* tcgen05.mma kind::tf32, 3-pass hi/lo split, fp32 accumulate in TMEM (~fp32 accuracy) */
  ANI_MLP_TF32 = 2,      /* single-pass TF32 on tcgen05 -- the reference's LAMMPS_ANI_ALLOW_TF32=1 (ani.cpp:41-43) */
  ANI_MLP_F16X3 = 3,     /* tcgen05.mma kind::f16 on scaled fp16 hi/lo planes, 3-pass, fp32 accumulate (~fp32 accuracy,
                            twice the TF32x3 rate): the default when the network widths allow it                       */
};

/* ---- model parameters ------------------------------------------------------------------------------
 * Replaces torch::jit::load + module_to_dtype (ani.cpp:13-33,46-61).  `path` is an `.aniw` blob written
 * once by `python -m lammps_ani_b200.extract` from ani2x.pt / ani1x_nr.pt / a NeuroChem .info file.  If
 * `path` ends in ".pt" the loader looks for "<path>.aniw" next to it and fails with instructions if absent. */
int ani_model_load(const char* path, ani_model** out);
void ani_model_free(ani_model* model);
/* attrs `num_models`, AEV geometry (lammps_ani.py:104-105).  Any out pointer may be NULL. */
int ani_model_info(const ani_model* model, int* num_species, int* aev_length, int* num_models, double* Rcr,
                   double* Rca);

/* ---- context ----------------------------------------------------------------------------------------
 * Replaces ANI::ANI(model_file, local_rank, use_num_models, use_cuaev, use_fullnbr, use_single)
 * (ani.h:32-37, ani.cpp:35-97): `device` = CUDA ordinal (the reference's local_rank; -1 = "cpu" is
 * REJECTED here), use_num_models = -1 for all members (select_models, lammps_ani.py:332-343),
 * use_cuaev: 1 = radial terms re-filtered at Rcr (cuaev behaviour), 0 = `pyaev` behaviour: every pair of
 * the supplied list contributes a radial term; use_single: 1 fp32, 0 fp64 arithmetic. */
int ani_ctx_create(const ani_model* model, int device, int use_num_models, int use_cuaev, int use_fullnbr,
                   int use_single, ani_ctx** out);
void ani_ctx_destroy(ani_ctx* ctx);
int ani_ctx_set_mlp_mode(ani_ctx* ctx, int mode); /* enum ani_mlp_mode; single precision only */
/* Run on the caller's CUDA stream (cudaStream_t as void*) instead of the context's own, so device-resident
 * callers (ani/kk, the multi-GPU halo exchange) keep stream order with their own kernels / NCCL.  The handle is
 * used verbatim: NULL means the CUDA legacy default stream (what Kokkos and torch use by default), not "unset";
 * ani_ctx_reset_stream goes back to the context's own non-blocking stream. */
int ani_ctx_set_stream(ani_ctx* ctx, void* cuda_stream);
int ani_ctx_reset_stream(ani_ctx* ctx);
/* Waits for the context's stream AND reports (once) any device-side error flag a step that was not synchronised raised:
 * 2 = more angular neighbours than the kernels hold, 3 = a value left the numeric range of a kernel (an MLP activation or gradient
 * outside the fp16 operand range of the f16x3 path, or an angular force term outside the fixed-point range of the fp32 AEV backward
 * kernel): ani_last_error() says which; the step's result is invalid.  */
int ani_ctx_synchronize(ani_ctx* ctx);
/* Per-context switches (the reference has only LAMMPS_ANI_ALLOW_TF32 / LAMMPS_ANI_PROFILING, ani.cpp:41, pair_ani.cpp:49).
 * "mlp_fused" 1 (default): the f16x3 MLP of a species bucket runs as one fused kernel, 0: layer by layer;
 * "pin_user" 1 (default 0): page-lock caller x / out_force arrays seen at a stable address and DMA from / to them
 * directly -- the caller then must call ani_ctx_release_host_buffers before freeing or reallocating them (pair style, bench);
 * "aev_two_pass" 1 (default) / 0; "gemm_ctas" n > 0: cap the persistent GEMM grids at n CTAs; "fused_prof" 1: print the
 * fused kernel's cycle counters to stderr; "halo_overlap" 1 (default) / 0: ani_compute_device_halo may run the ghost exchange on
 * a second stream under the interior atoms' AEV kernels.  The environment (ANI_B200_FUSED, ANI_B200_AEV_TWO_PASS, ANI_B200_GEMM_CTAS,
 * ANI_B200_FUSED_PROF) only sets the defaults a context starts with. */
int ani_ctx_set_option(ani_ctx* ctx, const char* name, int value);
/* Page-lock a caller array (cudaHostRegister) so that ani_dev_upload / ani_dev_download / ani_compute on it are direct DMA
 * instead of staged copies (LAMMPS's atom->x, v, f are pageable: 6 GB/s staged, > 25 GB/s pinned).  The registration is dropped
 * by ani_ctx_release_host_buffers, which the caller MUST call before the array is freed or reallocated (the pair style does when
 * atom->nmax changes), and by ani_ctx_destroy.  An array that is already registered is accepted as is. */
int ani_ctx_pin_host(ani_ctx* ctx, void* host, size_t bytes);

/* ---- neighbour data: call when the list was rebuilt (neighbor->ago == 0, pair_ani.cpp:107-151) ---------
 * species[ntotal] = LAMMPS type-1 for local AND ghost atoms (pair_ani.cpp:109-111).
 * Full list (ani.cpp:213-229): ilist[inum] local atom indices, numneigh[] INDEXED BY ATOM like LAMMPS'
 * list->numneigh, jlist = the flattened neighbours in ilist order with the special bits already masked
 * (j &= NEIGHMASK, pair_ani.cpp:138-139), npairs = sum numneigh[ilist[ii]].
 * Half list (ani.cpp:122-137): atom_index12 = int64[2][npairs], first row i, second row j.
 * on_device != 0: all pointers are device pointers (ani/kk, pair_ani_kokkos.cpp:143-169).               */
int ani_set_neighbors_full(ani_ctx* ctx, int nlocal, int ntotal, const int* species, int inum, const int* ilist,
                           const int* numneigh, const int* jlist, int64_t npairs, int on_device);
int ani_set_neighbors_half(ani_ctx* ctx, int nlocal, int ntotal, const int* species, int64_t npairs,
                           const int64_t* atom_index12, int on_device);

/* Padded neighbour table as LAMMPS' KOKKOS package keeps it (pair_ani_kokkos.cpp:114-169): table entry (i, k) lives at
 * table[i*row_stride + k*col_stride] (LayoutLeft: row_stride 1, col_stride = extent(0); LayoutRight: row_stride =
 * max_neighs, col_stride 1).  types[ntotal] are LAMMPS types; species = types[i] - type_offset (1 for LAMMPS).  The
 * table is de-padded ON THE DEVICE into the same flat list ani_set_neighbors_full takes (entries & neighmask), once per
 * rebuild instead of the reference's per-step masked_select.  on_device: all pointers are device pointers.            */
int ani_set_neighbors_table(ani_ctx* ctx, int nlocal, int ntotal, const int* types, int type_offset, int inum,
                            const int* ilist, const int* numneigh, const int* table, int64_t row_stride,
                            int64_t col_stride, int neighmask, int on_device);

/* Build the full neighbour list ON THE DEVICE and install it: replaces LAMMPS's `neighbor->build`
 * (npair_full_bin.cpp) + the host flatten of pair_ani.cpp:107-151 + the H2D copy of the list (ani.cpp:216-218).
 * x: [ntotal][3] doubles, local atoms first, on the host or (x_on_device != 0) on the device; species: host, [ntotal],
 * 0-based model species.  Every local atom i gets all j != i (local or ghost) with |x_i - x_j| <= cutoff (fp64).
 * lo / hi: a box that contains every atom (LAMMPS: sublo - cutghost, subhi + cutghost); atoms outside are clamped into the
 * border cells, which keeps the result correct and only costs time.  npairs_out (optional): stored pairs. */
int ani_build_neighbors(ani_ctx* ctx, int nlocal, int ntotal, const int* species, const double* x, int x_on_device,
                        double cutoff, const double lo[3], const double hi[3], int64_t* npairs_out);
/* Drop the page-locks the library took on caller buffers (with option "pin_user" 1, ani_compute page-locks an x / out_force
 * array it sees at a stable address).  Call it BEFORE freeing or reallocating such a buffer (the pair style does when atom->nmax or its out_force vector
 * grows); harmless otherwise. */
int ani_ctx_release_host_buffers(ani_ctx* ctx);
/* Copy the installed full list to host memory: numneigh [nlocal], jlist [npairs] in atom order (tests). */
int ani_debug_get_neighbors(ani_ctx* ctx, int* out_numneigh, int* out_jlist);

/* Number of usable CUDA devices (rank -> GPU mapping of the pair style, pair_ani.cpp:255-283). */
int ani_device_count(int* out);

/* ---- one timestep, host arrays (plain `ani` style; ANI::compute host overloads ani.cpp:100-265) ---------
 * x[ntotal][3] doubles.  out_energy: sum over LOCAL atoms.  out_force[ntotal][3] is OVERWRITTEN with the
 * force on local and ghost atoms (the caller folds ghosts back: pair_ani.cpp:197-210).  out_eatom[nlocal]
 * may be NULL (eflag_atom off).  out_virial9 (3x3 row major) may be NULL (vflag off); unlike cuaev, which
 * returns zeros (lammps_ani.py:151-153,208-211), it is always the true -sym(sum dE/dd (x) d).            */
int ani_compute(ani_ctx* ctx, const double* x, double* out_energy, double* out_force, double* out_eatom,
                double* out_virial9);

/* ---- one timestep, device arrays (`ani/kk`; ANI::compute tensor overload ani.cpp:268-316) -----------------
 * x_dev: [ntotal][3] of double (x_is_float=0) or float (1).  f_dev: [ntotal][3] double/float, ACCUMULATED in
 * place (f += force, pair_ani_kokkos.cpp:190-191).  eatom_dev: double[nlocal] accumulated, or NULL.
 * out_energy / out_virial9 are HOST pointers (may be NULL to skip the read-back and its sync).           */
int ani_compute_device(ani_ctx* ctx, const void* x_dev, int x_is_float, void* f_dev, int f_is_float,
                       double* eatom_dev, double* out_energy, double* out_virial9);

/* ---- halo helpers for the domain-decomposed driver (CommBrick forward_comm / reverse_comm(Pair*),
 * comm_brick.cpp:538-593,1023-1053; pair_ani.cpp:461-484) -- device pack / unpack, T = float or double.   */
int ani_halo_pack_x(ani_ctx* ctx, const void* x_dev, int is_float, const int64_t* send_index_dev, const void* shift_dev,
                    int64_t n, void* out_dev);
int ani_halo_unpack_add_f(ani_ctx* ctx, void* f_dev, int is_float, const int64_t* index_dev, const void* buf_dev,
                          int64_t n);

/* ---- device-resident time integration: what `modify->initial_integrate / post_force / final_integrate`
 * do on the host around the pair compute in the reference's run loop (external/lammps/src/verlet.cpp:253-256,347-349), so
 * that x, v, f stay on the GPU between neighbour rebuilds.  All arrays are DEVICE pointers, [n][3] doubles, mass_dev[n] per
 * atom; the kernels run on the context's stream.  Unit factors are the caller's (LAMMPS: dtv = dt, dtf = 0.5 dt ftm2v).
 *   ani_md_nve_initial : v += dtf f / m ; x += dtv v              (fix_nve.cpp FixNVE::initial_integrate)
 *   ani_md_nve_final   : v += dtf f / m                           (FixNVE::final_integrate)
 *   ani_md_langevin    : f += -m g1 v + sqrt(m) g2 (u - 0.5), u uniform on [0,1)   (fix_langevin.cpp post_force, the default
 *                        uniform-noise variant; g1 = 1/t_period/ftm2v, g2 = sqrt(24 boltz/t_period/dt/mvv2e)/ftm2v*sqrt(T));
 *                        u is a counter-based hash of (seed, step, tag[i] or i, component): independent of the decomposition
 *   ani_md_sum_mv2     : sum_i m_i |v_i|^2 to the host (kinetic energy / temperature on thermo steps; synchronises)
 *   ani_md_ghost_x     : x[nlocal+g] = x[owner[g]] + shift[g]    (forward_comm of a single periodic domain, comm_brick.cpp:549-593)
 *   ani_md_fold_ghost_f: f[owner[g]] += f[nlocal+g]               (reverse_comm(Pair*), pair_ani.cpp:197-201,461-484)          */
int ani_md_nve_initial(ani_ctx* ctx, double* x_dev, double* v_dev, const double* f_dev, const double* mass_dev, int n,
                       double dtv, double dtf);
int ani_md_nve_final(ani_ctx* ctx, double* v_dev, const double* f_dev, const double* mass_dev, int n, double dtf);
int ani_md_langevin(ani_ctx* ctx, const double* v_dev, double* f_dev, const double* mass_dev, const int64_t* tag_dev, int n,
                    double g1, double g2, uint64_t seed, int64_t step);
int ani_md_sum_mv2(ani_ctx* ctx, const double* v_dev, const double* mass_dev, int n, double* out_host);
/* max_i |x_i - x0_i|^2 over n atoms, to the host (synchronises): the displacement test of Neighbor::check_distance
 * (external/lammps/src/neighbor.cpp: rebuild when an atom has moved more than half the skin since the last build) without
 * copying the positions to the host.  fix ani/md asks it on the steps where neighbor->decide() would look at atom->x and
 * only brings x, v back when the answer says LAMMPS will rebuild. */
int ani_md_max_disp2(ani_ctx* ctx, const double* x_dev, const double* x0_dev, int n, double* out_host);
/* Allocation-free variant of ani_md_sum_mv2.  It has no counterpart in the reference and was added
 * for callers that cannot allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc): it calls no cudaMalloc / cudaFree / allocating routine and does not synchronise.  The per-block partial sums go
 * to a file-scope static device array (1024 doubles = 8 KB, held by every GPU that loads the library until the process exits);
 * the sum is written to out_dev[0] (DEVICE pointer, one double) in the order of the context's stream.  n must be within
 * 0 .. 16 777 216 atoms (ANI_MD_SUM_MV2_STATIC_MAX); beyond that the call fails with status 1 and a message.  All calls on a
 * device share the one scratch array: two calls that overlap on the GPU (different streams / contexts) race -- serialise them. */
#define ANI_MD_SUM_MV2_STATIC_MAX 16777216
int ani_md_sum_mv2_static(ani_ctx* ctx, const double* v_dev, const double* mass_dev, int n, double* out_dev);
int ani_md_ghost_x(ani_ctx* ctx, double* x_dev, const int* owner_dev, const double* shift_dev, int nlocal, int nghost);
int ani_md_fold_ghost_f(ani_ctx* ctx, double* f_dev, const int* owner_dev, int nlocal, int nghost);

/* ---- ghost exchange of a domain-decomposed run, inside the library: CommBrick::forward_comm and
 * reverse_comm(Pair*) (comm_brick.cpp:538-593,1023-1053; KOKKOS/comm_kokkos.cpp:187-360) + pair_ani.cpp:461-484 as one pack
 * kernel, one ncclSend / ncclRecv group over NVLink straight into / out of the ghost slab, one fold kernel -- all on the
 * context's stream, no host synchronisation.
 *   ani_halo_unique_id   : rank 0 creates the ncclUniqueId (128 bytes); the caller distributes it (MPI_Bcast / any control plane).
 *   ani_halo_plan_create : collective over the nranks callers.  send_index[nsend] local atom index and send_shift[nsend][3]
 *                          displacement added on the way out, concatenated in DESTINATION-rank order (send_counts[nranks]); the
 *                          ghosts of x / f, rows nlocal.., are ordered by SOURCE rank (recv_counts[nranks]).  A rank's send to
 *                          itself (periodic images) is a device copy.  nranks == 1 needs no id and no NCCL.
 *   ani_halo_forward     : ghost rows of x_dev <- owners' positions (+ shift).      [ntotal][3] doubles, device
 *   ani_halo_reverse     : f_dev[owner] += ghost rows of f_dev.                                                                 */
typedef struct ani_halo_plan ani_halo_plan;
int ani_halo_unique_id(char out[128]);
int ani_halo_plan_create(ani_ctx* ctx, const char unique_id[128], int nranks, int rank, int nlocal, const int64_t* send_index,
                         const double* send_shift, const int* send_counts, const int* recv_counts, ani_halo_plan** out);
void ani_halo_plan_destroy(ani_halo_plan* plan);
int ani_halo_forward(ani_halo_plan* plan, double* x_dev);
int ani_halo_reverse(ani_halo_plan* plan, double* f_dev);
/* One domain-decomposed step in one call: ani_halo_forward(x), ani_compute_device(x, f), ani_halo_reverse(f), with x / f the
 * caller's device arrays of doubles [ntotal][3] (f cleared by the caller: Verlet::force_clear).  What the single call adds:
 * when the neighbour list was built on the device (ani_build_neighbors records which local atoms have a ghost neighbour), the
 * exchange runs on a second stream under the INTERIOR atoms' share of the AEV kernels -- their rows read no ghost position and
 * produce no ghost force -- instead of in front of and behind the step (comm_brick.cpp:538-593,1023-1053 are serial in the
 * reference; KOKKOS overlaps nothing either).  ani_ctx_set_option(ctx, "halo_overlap", 0) serialises it again; a host-installed
 * list always runs serialised.  The context's stream sees the step complete, ghost forces folded, when the call returns. */
int ani_compute_device_halo(ani_ctx* ctx, ani_halo_plan* plan, double* x_dev, double* f_dev, double* eatom_dev,
                            double* out_energy, double* out_virial9);

/* Device-memory helpers on the context's device and stream, for callers that do not link the CUDA runtime (the LAMMPS
 * plugin's device-resident mode): alloc / free, blocking upload / download ordered after everything queued on the context's
 * stream (download also reports pending device-side error flags), asynchronous zero fill. */
int ani_dev_alloc(ani_ctx* ctx, size_t bytes, void** out_dev);
int ani_dev_free(ani_ctx* ctx, void* dev);
int ani_dev_upload(ani_ctx* ctx, void* dst_dev, const void* src_host, size_t bytes);
int ani_dev_download(ani_ctx* ctx, void* dst_host, const void* src_dev, size_t bytes);
int ani_dev_zero(ani_ctx* ctx, void* dst_dev, size_t bytes);

/* ---- introspection for tests, profiling and bench ------------------------------------------------------- */
/* Copy the AEV rows / dE/dAEV rows of the local atoms (atom order, [nlocal][aev_length] doubles) of the
 * last compute to host memory. */
int ani_debug_get_aev(ani_ctx* ctx, double* out_aev, double* out_daev);
/* Device time (ms, CUDA events on the context's stream) of the stages of the last compute:
 * [0] pack/H2D, [1] AEV forward, [2] MLP forward+backward, [3] AEV backward / force scatter, [4] reduce/D2H,
 * and the number of kernels launched in that compute. */
int ani_last_timing(ani_ctx* ctx, float stage_ms[5], int* num_launches);
int ani_ctx_enable_timing(ani_ctx* ctx, int on);

/* Test hook for the tensor-core GEMM kernel alone: C[R][N] = A[R][K] . B[N][K]^T (host pointers, row major),
 * three_pass = 1: TF32x3 split, 0: single TF32, 2: the fp16 hi/lo (F16X3) kernel.  N must have a divisor <= 256 that is a multiple of 16; K % 8 == 0. */
int ani_debug_tc_gemm(int device, int R, int N, int K, const float* A, const float* B, float* C, int three_pass);

const char* ani_last_error(void);
int ani_abi_version(void);

#ifdef __cplusplus
}
#endif
#endif /* ANI_B200_H */
