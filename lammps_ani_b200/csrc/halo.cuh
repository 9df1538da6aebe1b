// halo.cuh -- the ghost exchange of a domain-decomposed run inside the library.
//
// Replaces CommBrick::forward_comm / reverse_comm(Pair*) (external/lammps/src/comm_brick.cpp:538-593,1023-1053; device-buffer
// variant KOKKOS/comm_kokkos.cpp:187-360) and the pair style's pack / unpack_reverse_comm (src/pair_ani.cpp:461-484): instead
// of six host-staged MPI swaps, one pack kernel (all directions, periodic shift applied), ONE group of ncclSend / ncclRecv
// over NVLink straight into the ghost slab of the receiver's coordinate array (ghosts are ordered by source rank, so nothing
// is unpacked), and on the way back one group + one atomic fold kernel.  The communicator is created from a caller-supplied
// ncclUniqueId; Python / the pair style only hand over the plan.  NCCL is loaded with dlopen (libnccl.so.2: torch's copy when
// torch is in the process, the system's otherwise), so single-GPU users do not depend on it.
#pragma once
#include <dlfcn.h>
#include <nccl.h>

#include <vector>

#include "common.cuh"
#include "misc.cuh"

namespace anib {

struct NcclApi {
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
  bool ok = false;
};

inline NcclApi& nccl_api() {
  static NcclApi api;
  static bool tried = false;
  if (tried) return api;
  tried = true;
  void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
  if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
  if (!h) return api;
  api.GetUniqueId = (decltype(api.GetUniqueId))dlsym(h, "ncclGetUniqueId");
  api.CommInitRank = (decltype(api.CommInitRank))dlsym(h, "ncclCommInitRank");
  api.CommDestroy = (decltype(api.CommDestroy))dlsym(h, "ncclCommDestroy");
  api.Send = (decltype(api.Send))dlsym(h, "ncclSend");
  api.Recv = (decltype(api.Recv))dlsym(h, "ncclRecv");
  api.GroupStart = (decltype(api.GroupStart))dlsym(h, "ncclGroupStart");
  api.GroupEnd = (decltype(api.GroupEnd))dlsym(h, "ncclGroupEnd");
  api.GetErrorString = (decltype(api.GetErrorString))dlsym(h, "ncclGetErrorString");
  api.ok = api.GetUniqueId && api.CommInitRank && api.CommDestroy && api.Send && api.Recv && api.GroupStart && api.GroupEnd;
  return api;
}

#define ANIB_NCCL_OK(expr)                                                                                             \
  do {                                                                                                                 \
    ncclResult_t _r = (expr);                                                                                          \
    if (_r != ncclSuccess) {                                                                                           \
      ::anib::set_error(std::string(#expr) + " failed: " + (nccl_api().GetErrorString ? nccl_api().GetErrorString(_r) : "?")); \
      return 1;                                                                                                        \
    }                                                                                                                  \
  } while (0)

struct HaloPlan {
  int device = 0, nranks = 1, rank = 0, nlocal = 0;
  cudaStream_t stream = nullptr;
  ncclComm_t comm = nullptr;
  int64_t nsend = 0, nghost = 0;
  std::vector<int64_t> send_off, recv_off;      // [nranks + 1] rows
  int64_t* d_send_index = nullptr;
  double* d_send_shift = nullptr;
  double* d_send_buf = nullptr;                 // [nsend][3]: packed positions out / ghost forces back in
};

}  // namespace anib
