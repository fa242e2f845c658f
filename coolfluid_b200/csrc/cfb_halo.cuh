// cfb200 — multi-GPU: one partition per GPU, state halo exchange over NCCL.
// Included by cfb_capi.cu only (CUDA build).  Replaces DataHandle<State*,GLOBAL>::synchronize()
// (src/Common/MPI/MPICommPattern.hh:1735-1782) for the states array: after every update each rank sends the states
// of its updatable cells that are overlap cells elsewhere and receives its own overlap cells.
//
//   cfb200_halo_begin     pack kernel on the engine's stream -> event -> NCCL send/recv group + unpack kernel on a
//                         dedicated communication stream -> event; returns at once
//   cfb200_compute_rhs    with an exchange in flight: applyBC + gradients of the leading `nIndependent` cells (their
//                         stencils hold no received cell) run while the halo is on the wire, then the engine's
//                         stream waits for the unpack and continues with the rest
//   cfb200_halo_exchange  begin + wait (what a caller that does not overlap uses)
//
// NCCL is bound at run time (dlopen of the libnccl.so.2 already loaded in the process, e.g. PyTorch's, else the
// system one), so that libcfb200.so neither drags in a second NCCL nor fails to load on a box without it.
#pragma once
#include <dlfcn.h>
#include <nccl.h>

namespace cfb {

struct NcclApi {
  void* dl = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;

  static NcclApi& get() {
    static NcclApi api;
    if (api.dl) return api;
    void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);
    if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!h) throw DeviceError(CFB200_ERR_NCCL, std::string("libnccl.so.2 cannot be loaded: ") + dlerror());
    auto sym = [&](const char* n) { void* p = dlsym(h, n); if (!p) throw DeviceError(CFB200_ERR_NCCL, std::string("missing NCCL symbol ") + n); return p; };
    api.GetUniqueId = (decltype(api.GetUniqueId))sym("ncclGetUniqueId");
    api.CommInitRank = (decltype(api.CommInitRank))sym("ncclCommInitRank");
    api.CommDestroy = (decltype(api.CommDestroy))sym("ncclCommDestroy");
    api.GroupStart = (decltype(api.GroupStart))sym("ncclGroupStart");
    api.GroupEnd = (decltype(api.GroupEnd))sym("ncclGroupEnd");
    api.Send = (decltype(api.Send))sym("ncclSend");
    api.Recv = (decltype(api.Recv))sym("ncclRecv");
    api.AllReduce = (decltype(api.AllReduce))sym("ncclAllReduce");
    api.GetErrorString = (decltype(api.GetErrorString))sym("ncclGetErrorString");
    api.dl = h;
    return api;
  }
};

#define CFB_NCCL(call) do { ncclResult_t r_ = (call); if (r_ != ncclSuccess) \
  throw cfb::DeviceError(CFB200_ERR_NCCL, std::string(#call) + ": " + cfb::NcclApi::get().GetErrorString(r_)); } while (0)

// states of the listed cells <-> packed [cell][q] buffer (the layout on the wire)
struct HaloPack {     // idx over nCells*q
  const double* U; int nStates, q; const int* cells; double* buf;
  CFB_HD void operator()(int idx) const { const int i = idx/q, v = idx - i*q; buf[idx] = U[(size_t)v*nStates + cells[i]]; }
};
struct HaloUnpack {
  double* U; int nStates, q; const int* cells; const double* buf;
  CFB_HD void operator()(int idx) const { const int i = idx/q, v = idx - i*q; U[(size_t)v*nStates + cells[i]] = buf[idx]; }
};

struct Halo {
  ncclComm_t comm = nullptr;
  int rank = 0, nRanks = 1, dev = 0;
  std::vector<int32_t> sendPtr, recvPtr;
  int* dSend = nullptr; int* dRecv = nullptr;
  double* sbuf = nullptr; double* rbuf = nullptr; double* red = nullptr;
  cudaStream_t cs = nullptr;
  cudaEvent_t packed = nullptr, done = nullptr;
  bool haveLists = false;

  ~Halo() {
    cudaSetDevice(dev);
    if (cs) cudaStreamSynchronize(cs);      // no NCCL / unpack kernel may still be using the buffers freed below
    if (comm) NcclApi::get().CommDestroy(comm);
    cudaFree(dSend); cudaFree(dRecv); cudaFree(sbuf); cudaFree(rbuf); cudaFree(red);
    if (packed) cudaEventDestroy(packed);
    if (done) cudaEventDestroy(done);
    if (cs) cudaStreamDestroy(cs);
  }
};

} // namespace cfb
