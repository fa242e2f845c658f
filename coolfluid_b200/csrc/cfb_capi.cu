// cfb200 — the product library: CUDA backend of cfb::Engine + the extern "C" surface of include/cfb200.h.
// Compiled for sm_100a only (nvcc -gencode arch=compute_100a,code=sm_100a -fmad=false).  There is no CPU
// execution path in this library: without a usable CUDA device cfb200_create returns CFB200_ERR_NO_DEVICE.
#include <type_traits>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <string>
#include <vector>
#include <mutex>
#include "cfb_engine.hpp"

namespace cfb {

#define CFB_CUDA(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) \
  throw cfb::DeviceError(CFB200_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_)); } while (0)

template <class F, int BS, int MINB>
__global__ void __launch_bounds__(BS, MINB) k_run(const F f, const int n, const int first) {
  const int i = first + blockIdx.x*BS + threadIdx.x;     // items [first, n): the launch window (CudaBackend::window)
  if (i < n) f(i);
}

} // namespace cfb
#include "cfb_staged.cuh"
namespace cfb {
// compiled in cfb_hot.cu (its own translation unit: the register allocation of this kernel moved between 238 and 252
// registers, 28.3 vs 27.4 ms at 220^3, with unrelated edits elsewhere in this file; alone it is stable)
extern template __global__ void k_staged<InnerFaceK<3, true, 0, 0, 1>, 128, 2>(const InnerFaceK<3, true, 0, 0, 1>, const int, const int);

// Two-phase variant for functors that stage ONE block per phase (f.run(i, phase, ...), f.slot(i, phase)): q = 9 blocks.
template <class F, int BS, int MINB>
__global__ void __launch_bounds__(BS, MINB) k_staged2(const F f, const int n, const int first) {
  extern __shared__ double sm[];
  constexpr int STRIDE = BS + 1;
  static_assert(F::BLK >= 32 && F::STAGE >= F::BLK + 1, "staging layout");
  const int t = threadIdx.x, lane = t & 31, warpBase = t & ~31;
  const int i = first + blockIdx.x*BS + t;
  double* out = f.out();
  typename F::State st;            // what phase 1 keeps of phase 0, in registers across the write-out
  for (int phase = 0; phase < 2; ++phase) {
    int s = -1;
    if (i < n) { f.run(i, phase, st, sm + t, STRIDE); s = f.slot(i, phase); }
    sm[F::BLK*STRIDE + t] = __hiloint2double(0, s);
    __syncwarp();
    // batches of WB entries, as in k_staged: the shared-memory reads of a batch are issued before its first global store
    constexpr int WB = (F::BLK % 9 == 0) ? 9 : ((F::BLK % 5 == 0) ? 5 : 1);
    int k = lane, lf = 0;
#pragma unroll 1
    for (int it = 0; it < F::BLK; it += WB) {
      double x[WB]; size_t dst[WB]; bool on[WB];
#pragma unroll
      for (int j = 0; j < WB; ++j) {
        const int slot = __double2loint(sm[F::BLK*STRIDE + warpBase + lf]);
        x[j] = sm[k*STRIDE + warpBase + lf];
        on[j] = slot >= 0;
        dst[j] = (size_t)slot*F::BLK + k;
        k += 32;
        if (k >= F::BLK) { k -= F::BLK; ++lf; }
      }
#pragma unroll
      for (int j = 0; j < WB; ++j) if (on[j]) out[dst[j]] = x[j];
    }
    __syncwarp();
  }
}

// Phased functors (F::begin, F::phase(ph), F::slot(f, ph)): the per-face state lives in registers across the two phases,
// each phase leaves ONE finished block in the thread's staging area, which the warp writes out (whole rows, coalesced)
// before the next phase reuses the area.  Warps stay independent (no block-wide barrier).
template <class F, int BS, int MINB>
__global__ void __launch_bounds__(BS, MINB) k_phased(const F f, const int n, const int first) {
  extern __shared__ double sm[];
  constexpr int STRIDE = BS + 1;
  const int t = threadIdx.x, lane = t & 31, warpBase = t & ~31;
  const int i = first + blockIdx.x*BS + t;
  typename F::State st;
  const bool valid = i < n;
  const bool active = valid && f.begin(i, st, sm + t, STRIDE);
  double* out = f.out();
  constexpr int WB = (F::BLK % 5 == 0) ? 5 : 4;
  static_assert(F::BLK % WB == 0 && F::BLK <= 32 && 2*F::BLK >= 32, "write-out batch");
#pragma unroll 1
  for (int ph = 0; ph < 2; ++ph) {
    if (active) f.phase(ph, st, sm + t, STRIDE);
    sm[F::SLOTROW*STRIDE + t] = __hiloint2double(0, valid ? f.slot(i, ph) : -1);
    __syncwarp();
    // entry e = lane + 32*it of the warp's 32 x BLK finished values: face lf = e / BLK, k = e % BLK, kept incrementally
    int k = lane, lf = 0;
    if (k >= F::BLK) { k -= F::BLK; ++lf; }
#pragma unroll 1
    for (int it = 0; it < F::BLK; it += WB) {
      double x[WB]; size_t dst[WB]; bool on[WB];
#pragma unroll
      for (int j = 0; j < WB; ++j) {
        const int slot = __double2loint(sm[F::SLOTROW*STRIDE + warpBase + lf]);
        x[j] = sm[k*STRIDE + warpBase + lf];
        on[j] = slot >= 0;
        dst[j] = (size_t)slot*F::BLK + k;
        k += 32;
        if (k >= F::BLK) { k -= F::BLK; ++lf; }
        if (k >= F::BLK) { k -= F::BLK; ++lf; }
      }
#pragma unroll
      for (int j = 0; j < WB; ++j) if (on[j]) out[dst[j]] = x[j];
    }
    __syncwarp();
  }
}

// One cell per thread with F::BLK doubles of staging per thread (entry k of thread t at sm[k*(BS+1) + t]); after the cells
// are evaluated the BLOCK writes the BS finished q x q blocks to their slots, whole rows at a time (coalesced).
template <class F, int BS>
__global__ void __launch_bounds__(BS) k_cell_staged(const F f, const int n, const int first) {
  extern __shared__ double sm[];
  constexpr int STRIDE = BS + 1;
  __shared__ int slots[BS];
  const int t = threadIdx.x;
  const int i = first + blockIdx.x*BS + t;
  int s = -1;
  if (i < n) { f.run(i, sm + t, STRIDE); s = f.slot(i); }
  slots[t] = s;
  __syncthreads();
  double* out = f.out();
  int k = t, lc = 0;
  while (k >= F::BLK) { k -= F::BLK; ++lc; }
#pragma unroll 1
  for (int it = 0; it < F::BLK; ++it) {      // entry e = t + BS*it of the block's BS x BLK values: cell lc = e / BLK, k = e % BLK
    const int slot = slots[lc];
    if (slot >= 0) out[(size_t)slot*F::BLK + k] = sm[k*STRIDE + lc];
    k += BS;
    while (k >= F::BLK) { k -= F::BLK; ++lc; }
  }
}

// fixed-shape block sum (run-to-run deterministic): each block reduces 1024 consecutive entries
__global__ void __launch_bounds__(256) k_block_sum(const double* __restrict__ x, double* __restrict__ out, int n) {
  __shared__ double s[256];
  const int base = blockIdx.x*1024;
  double a = 0.;
  for (int k = 0; k < 4; ++k) { const int i = base + k*256 + threadIdx.x; if (i < n) a += x[i]; }
  s[threadIdx.x] = a;
  __syncthreads();
  for (int w = 128; w > 0; w >>= 1) { if (threadIdx.x < w) s[threadIdx.x] += s[threadIdx.x + w]; __syncthreads(); }
  if (threadIdx.x == 0) out[blockIdx.x] = s[0];
}

// fixed-shape block minimum (the minimum is exact whatever the order)
__global__ void __launch_bounds__(256) k_block_min(const double* __restrict__ x, double* __restrict__ out, int n) {
  __shared__ double s[256];
  const int base = blockIdx.x*1024;
  double a = 1.7976931348623157e308;
  for (int k = 0; k < 4; ++k) { const int i = base + k*256 + threadIdx.x; if (i < n) a = (x[i] < a) ? x[i] : a; }
  s[threadIdx.x] = a;
  __syncthreads();
  for (int w = 128; w > 0; w >>= 1) {
    if (threadIdx.x < w) { const double b = s[threadIdx.x + w]; if (b < s[threadIdx.x]) s[threadIdx.x] = b; }
    __syncthreads();
  }
  if (threadIdx.x == 0) out[blockIdx.x] = s[0];
}

// resident 128-thread blocks per SM a staged face functor asks for (F::STAGED_MINB), 2 unless it says otherwise
template <class F, class = void> struct StagedMinB { static constexpr int value = 2; };
template <class F> struct StagedMinB<F, std::void_t<decltype(F::STAGED_MINB)> > { static constexpr int value = F::STAGED_MINB; };

template <class F, class = void> struct HasMinB : std::false_type {};
template <class F> struct HasMinB<F, std::void_t<decltype(F::MINB)> > : std::true_type {};

struct CudaBackend {
  int dev = 0;
  cudaStream_t stream = nullptr;
  cudaStream_t copyStream = nullptr;       // device -> host copies that overlap the kernels still to run (d2hOverlapped)
  cudaEvent_t copyReady = nullptr;
  bool copiesInFlight = false;
  // launch window: the NEXT launch covers the items [first, end) of its index range instead of all of them (Engine uses it
  // to run the face kernel and the residual gather in two parts so that the host copy of the first part's rhs overlaps
  // the second part)
  int winFirst = 0, winEnd = -1;
  void window(int first, int end) { winFirst = first; winEnd = end; }
  int takeWindow(int& n) {
    int w0 = 0;
    if (winEnd >= 0) { w0 = winFirst; if (winEnd < n) n = winEnd; }
    winFirst = 0; winEnd = -1;
    return w0;
  }
  bool timing = false;
  struct Ev { const char* name; cudaEvent_t a, b; };
  std::vector<Ev> evs;
  size_t nEv = 0;
  std::vector<std::string> nameStore;

  explicit CudaBackend(int device) : dev(device) {
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count <= 0)
      throw DeviceError(CFB200_ERR_NO_DEVICE, "no usable CUDA device (cfb200 has no CPU path)");
    if (device < 0 || device >= count) throw DeviceError(CFB200_ERR_INVALID_ARG, "device ordinal out of range");
    CFB_CUDA(cudaSetDevice(dev));
    CFB_CUDA(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
    CFB_CUDA(cudaStreamCreateWithFlags(&copyStream, cudaStreamNonBlocking));
    CFB_CUDA(cudaEventCreateWithFlags(&copyReady, cudaEventDisableTiming));
  }
  ~CudaBackend() {
    cudaSetDevice(dev);
    if (copyStream) { cudaStreamSynchronize(copyStream); cudaStreamDestroy(copyStream); }
    if (copyReady) cudaEventDestroy(copyReady);
    for (auto& e : evs) { cudaEventDestroy(e.a); cudaEventDestroy(e.b); }
    if (stream) cudaStreamDestroy(stream);
  }
  CudaBackend(const CudaBackend&) = delete;
  CudaBackend& operator=(const CudaBackend&) = delete;

  void activate() { CFB_CUDA(cudaSetDevice(dev)); }
  void* alloc(size_t bytes) { void* p = nullptr; CFB_CUDA(cudaMalloc(&p, bytes)); return p; }
  void free(void* p) { cudaFree(p); }
  void h2d(void* dst, const void* src, size_t bytes) {
    CFB_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, stream));
    CFB_CUDA(cudaStreamSynchronize(stream));
  }
  void d2h(void* dst, const void* src, size_t bytes) {
    CFB_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, stream));
    CFB_CUDA(cudaStreamSynchronize(stream));
  }
  // device -> host copy of data that is final at this point of the stream, on the copy stream: it runs while the kernels
  // launched afterwards execute (PCIe and the SMs are independent); joinCopies() before the host reads it
  void d2hOverlapped(void* dst, const void* src, size_t bytes) {
    // every copy waits for what is queued on the engine's stream at this point (a re-recorded event keeps earlier waits)
    CFB_CUDA(cudaEventRecord(copyReady, stream));
    CFB_CUDA(cudaStreamWaitEvent(copyStream, copyReady, 0));
    copiesInFlight = true;
    CFB_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, copyStream));
  }
  void joinCopies() {
    if (!copiesInFlight) return;
    copiesInFlight = false;
    CFB_CUDA(cudaStreamSynchronize(copyStream));
  }
  void d2d(void* dst, const void* src, size_t bytes) { CFB_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToDevice, stream)); }
  void memset(void* p, int v, size_t bytes) { CFB_CUDA(cudaMemsetAsync(p, v, bytes, stream)); }
  void sync() { CFB_CUDA(cudaStreamSynchronize(stream)); }
  void* streamPtr() { return (void*)stream; }

  void enableTiming(bool on) { timing = on; }
  void resetTimings() { nEv = 0; }
  int timings(const char** names, float* ms, int cap) {
    CFB_CUDA(cudaStreamSynchronize(stream));
    int n = 0;
    for (size_t i = 0; i < nEv && n < cap; ++i, ++n) {
      float t = 0.f;
      CFB_CUDA(cudaEventElapsedTime(&t, evs[i].a, evs[i].b));
      names[n] = evs[i].name; ms[n] = t;
    }
    return n;
  }

  template <class F> void launch(const char* name, const F& f, int n, int bs) {
    const int w0 = takeWindow(n), m = n - w0;            // items [w0, n) of the launch window, else all
    if (m <= 0) return;
    Ev* ev = nullptr;
    if (timing) {
      if (nEv == evs.size()) { Ev e; e.name = name; CFB_CUDA(cudaEventCreate(&e.a)); CFB_CUDA(cudaEventCreate(&e.b)); evs.push_back(e); }
      ev = &evs[nEv++]; ev->name = name;
      CFB_CUDA(cudaEventRecord(ev->a, stream));
    }
    // launch shape: a functor that declares `static constexpr int MINB = 4` runs as 128 threads with >= 4 resident
    // blocks per SM (<= 128 registers); otherwise bs = 256 -> 256 threads, bs = 128 -> 128 threads, >= 2 blocks/SM
    if constexpr (HasMinB<F>::value) k_run<F, 128, F::MINB><<<(m + 127)/128, 128, 0, stream>>>(f, n, w0);
    else if (bs == 128) k_run<F, 128, 2><<<(m + 127)/128, 128, 0, stream>>>(f, n, w0);
    else k_run<F, 256, 1><<<(m + 255)/256, 256, 0, stream>>>(f, n, w0);
    CFB_CUDA(cudaGetLastError());
    if (ev) CFB_CUDA(cudaEventRecord(ev->b, stream));
  }

  // 128-thread blocks with F::STAGE doubles of shared-memory staging per thread; minBlocks = resident blocks per SM
  template <class F> void launchStaged(const char* name, const F& f, int n, int minBlocks) {
    const int w0 = takeWindow(n);       // consumed even when nothing is launched
    if (n - w0 <= 0) return;
    Ev* ev = nullptr;
    if (timing) {
      if (nEv == evs.size()) { Ev e; e.name = name; CFB_CUDA(cudaEventCreate(&e.a)); CFB_CUDA(cudaEventCreate(&e.b)); evs.push_back(e); }
      ev = &evs[nEv++]; ev->name = name;
      CFB_CUDA(cudaEventRecord(ev->a, stream));
    }
    auto go = [&](auto kern, int BS) {
      const size_t smem = (size_t)F::STAGE*(BS + 1)*sizeof(double);
      CFB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      // without this the driver sizes the shared-memory carve-out for ONE block and the intended number of resident
      // blocks does not fit.  Ask for just what two resident blocks need (+ 1 KB each of system use): the rest of the
      // 256 KB stays L1 for the gathers of the prologue (measured 31.45 -> 30.56 ms against the maximum carve-out)
      int pct = (int)((StagedMinB<F>::value*(smem + 1024)*100 + 228*1024 - 1)/(228*1024)) + 3;
      if (pct > 100) pct = 100;
      CFB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, pct));
      kern<<<(n - w0 + BS - 1)/BS, BS, smem, stream>>>(f, n, w0);
    };
    // launch shape = resident 128-thread blocks per SM = warps per scheduler (each scheduler owns 16384 registers, so the
    // only register budgets that change occupancy are 255 / 168 / 128 per thread; measured on B200: 2 is fastest)
    // (3 and 4 resident blocks = 168 / 128 registers were measured 17 % and more slower, DESIGN.md §4.2, and are no longer
    // compiled: they tripled the build time of the face kernels)
    (void)minBlocks;
    go(k_staged<F, 128, StagedMinB<F>::value>, 128);    //  2: 8 warps per SM, 255 registers (the FD kernels)
    CFB_CUDA(cudaGetLastError());
    if (ev) CFB_CUDA(cudaEventRecord(ev->b, stream));
  }

  template <class F> void launchCellStaged(const char* name, const F& f, int n) {
    const int w0 = takeWindow(n);       // consumed even when nothing is launched
    if (n - w0 <= 0) return;
    Ev* ev = nullptr;
    if (timing) {
      if (nEv == evs.size()) { Ev e; e.name = name; CFB_CUDA(cudaEventCreate(&e.a)); CFB_CUDA(cudaEventCreate(&e.b)); evs.push_back(e); }
      ev = &evs[nEv++]; ev->name = name;
      CFB_CUDA(cudaEventRecord(ev->a, stream));
    }
    constexpr int BS = 128;
    const size_t smem = (size_t)F::BLK*(BS + 1)*sizeof(double);
    auto kern = k_cell_staged<F, BS>;
    CFB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<(n - w0 + BS - 1)/BS, BS, smem, stream>>>(f, n, w0);
    CFB_CUDA(cudaGetLastError());
    if (ev) CFB_CUDA(cudaEventRecord(ev->b, stream));
  }

  // three resident 128-thread blocks per SM (168 registers), F::STAGE doubles of staging per thread
  template <class F> void launchPhased(const char* name, const F& f, int n) {
    const int w0 = takeWindow(n);       // consumed even when nothing is launched
    if (n - w0 <= 0) return;
    Ev* ev = nullptr;
    if (timing) {
      if (nEv == evs.size()) { Ev e; e.name = name; CFB_CUDA(cudaEventCreate(&e.a)); CFB_CUDA(cudaEventCreate(&e.b)); evs.push_back(e); }
      ev = &evs[nEv++]; ev->name = name;
      CFB_CUDA(cudaEventRecord(ev->a, stream));
    }
    constexpr int BS = 128;
    const size_t smem = (size_t)F::STAGE*(BS + 1)*sizeof(double);
    auto kern = k_phased<F, BS, 3>;
    CFB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    CFB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
    kern<<<(n - w0 + BS - 1)/BS, BS, smem, stream>>>(f, n, w0);
    CFB_CUDA(cudaGetLastError());
    if (ev) CFB_CUDA(cudaEventRecord(ev->b, stream));
  }

  template <class F> void launchStaged2(const char* name, const F& f, int n) {
    const int w0 = takeWindow(n);       // consumed even when nothing is launched
    if (n - w0 <= 0) return;
    Ev* ev = nullptr;
    if (timing) {
      if (nEv == evs.size()) { Ev e; e.name = name; CFB_CUDA(cudaEventCreate(&e.a)); CFB_CUDA(cudaEventCreate(&e.b)); evs.push_back(e); }
      ev = &evs[nEv++]; ev->name = name;
      CFB_CUDA(cudaEventRecord(ev->a, stream));
    }
    constexpr int BS = 128;
    const size_t smem = (size_t)F::STAGE*(BS + 1)*sizeof(double);
    auto kern = k_staged2<F, BS, 2>;
    CFB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int pct = (int)((2*(smem + 1024)*100 + 228*1024 - 1)/(228*1024)) + 3;      // as in launchStaged
    if (pct > 100) pct = 100;
    CFB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, pct));
    kern<<<(n - w0 + BS - 1)/BS, BS, smem, stream>>>(f, n, w0);
    CFB_CUDA(cudaGetLastError());
    if (ev) CFB_CUDA(cudaEventRecord(ev->b, stream));
  }

  // sum of x[0..n) with a fixed reduction shape (run-to-run deterministic); the scratch behind x[n) is preallocated by
  // the engine (Engine::setMesh: nCells + ceil(nCells/1024) + 1024 doubles), so no allocation happens per call
  double sumOrdered(double* x, int n) {
    if (n <= 0) return 0.;
    const int nb = (n + 1023)/1024;
    double* out = x + n;
    std::vector<double> h;
    if (nb > 1024) {
      // two levels: reduce the block sums again
      const int nb2 = (nb + 1023)/1024;
      double* out2 = out + nb;
      k_block_sum<<<nb, 256, 0, stream>>>(x, out, n);
      k_block_sum<<<nb2, 256, 0, stream>>>(out, out2, nb);
      CFB_CUDA(cudaGetLastError());
      h.resize(nb2);
      CFB_CUDA(cudaMemcpyAsync(h.data(), out2, sizeof(double)*nb2, cudaMemcpyDeviceToHost, stream));
    } else {
      k_block_sum<<<nb, 256, 0, stream>>>(x, out, n);
      CFB_CUDA(cudaGetLastError());
      h.resize(nb);
      CFB_CUDA(cudaMemcpyAsync(h.data(), out, sizeof(double)*nb, cudaMemcpyDeviceToHost, stream));
    }
    CFB_CUDA(cudaStreamSynchronize(stream));
    double s = 0.;
    for (double v : h) s += v;
    return s;
  }
  double minOrdered(double* x, int n) {
    if (n <= 0) return 0.;
    const int nb = (n + 1023)/1024;
    double* out = x + n;
    k_block_min<<<nb, 256, 0, stream>>>(x, out, n);
    CFB_CUDA(cudaGetLastError());
    std::vector<double> h(nb);
    CFB_CUDA(cudaMemcpyAsync(h.data(), out, sizeof(double)*nb, cudaMemcpyDeviceToHost, stream));
    CFB_CUDA(cudaStreamSynchronize(stream));
    double m = h[0];
    for (double v : h) m = (v < m) ? v : m;
    return m;
  }
};

} // namespace cfb

typedef cfb::CudaBackend BackendT;
#define CFB_FN(x) cfb200_##x
#include "cfb_capi.inc"
#include "cfb_halo.cuh"
#include "cfb_solver.cuh"

// ---- multi-GPU entry points (include/cfb200.h) -----------------------------------------------------------------
namespace {
cfb::Halo* haloOf(cfb200_handle* h, bool create) {
  if (!h->ext && create) {
    h->ext = new cfb::Halo();
    h->extFree = [](void* p) { delete static_cast<cfb::Halo*>(p); };
  }
  return static_cast<cfb::Halo*>(h->ext);
}
void haloBegin(cfb::Engine<BackendT>& e, cfb::Halo& H) {
  if (!H.comm || !H.haveLists) throw std::runtime_error("comm_init / set_halo have not been called");
  if (e.haloPending) throw std::runtime_error("a halo exchange is already in flight");
  cfb::NcclApi& N = cfb::NcclApi::get();
  const int q = e.q();
  const int nSend = H.sendPtr[H.nRanks], nRecv = H.recvPtr[H.nRanks];
  if (nSend) e.bk.launch("haloPack", cfb::HaloPack{e.d.U, e.d.nStates, q, H.dSend, H.sbuf}, nSend*q, 256);
  CFB_CUDA(cudaEventRecord(H.packed, e.bk.stream));
  CFB_CUDA(cudaStreamWaitEvent(H.cs, H.packed, 0));
  CFB_NCCL(N.GroupStart());
  for (int p = 0; p < H.nRanks; ++p) {
    const int ns = H.sendPtr[p+1] - H.sendPtr[p], nr = H.recvPtr[p+1] - H.recvPtr[p];
    if (ns) CFB_NCCL(N.Send(H.sbuf + (size_t)H.sendPtr[p]*q, (size_t)ns*q, ncclDouble, p, H.comm, H.cs));
    if (nr) CFB_NCCL(N.Recv(H.rbuf + (size_t)H.recvPtr[p]*q, (size_t)nr*q, ncclDouble, p, H.comm, H.cs));
  }
  CFB_NCCL(N.GroupEnd());
  if (nRecv) {
    const int n = nRecv*q;
    cfb::k_run<cfb::HaloUnpack, 256, 1><<<(n + 255)/256, 256, 0, H.cs>>>(cfb::HaloUnpack{e.d.U, e.d.nStates, q, H.dRecv, H.rbuf}, n, 0);
    CFB_CUDA(cudaGetLastError());
  }
  CFB_CUDA(cudaEventRecord(H.done, H.cs));
  e.haloPending = true;
  cudaEvent_t done = H.done;
  cudaStream_t s = e.bk.stream;
  e.haloWait = [done, s]() { CFB_CUDA(cudaStreamWaitEvent(s, done, 0)); };
}
}

extern "C" {

int cfb200_nccl_unique_id(void* id128) {
  if (!id128) return CFB200_ERR_INVALID_ARG;
  try {
    ncclUniqueId id;
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
    if (cfb::NcclApi::get().GetUniqueId(&id) != ncclSuccess) return CFB200_ERR_NCCL;
    std::memcpy(id128, &id, 128);
    return CFB200_OK;
  } catch (const cfb::DeviceError& e) { return e.code; }
}

int cfb200_comm_init(cfb200_handle* h, int rank, int nRanks, const void* id128) {
  if (!id128 || nRanks < 1 || rank < 0 || rank >= nRanks) return CFB200_ERR_INVALID_ARG;
  return guarded(h, [&](cfb::Engine<BackendT>& e) {
    cfb::Halo* H = haloOf(h, true);
    if (H->comm) throw std::runtime_error("communicator already initialised");
    ncclUniqueId id;
    std::memcpy(&id, id128, 128);
    H->rank = rank; H->nRanks = nRanks; H->dev = e.bk.dev;
    CFB_NCCL(cfb::NcclApi::get().CommInitRank(&H->comm, nRanks, id, rank));
    CFB_CUDA(cudaStreamCreateWithFlags(&H->cs, cudaStreamNonBlocking));
    CFB_CUDA(cudaEventCreateWithFlags(&H->packed, cudaEventDisableTiming));
    CFB_CUDA(cudaEventCreateWithFlags(&H->done, cudaEventDisableTiming));
    CFB_CUDA(cudaMalloc((void**)&H->red, sizeof(double)*64));
    return CFB200_OK;
  });
}

int cfb200_set_halo(cfb200_handle* h, const int32_t* sendPtr, const int32_t* sendCells, const int32_t* recvPtr,
                    const int32_t* recvCells, int32_t nIndependent) {
  if (!sendPtr || !recvPtr) return CFB200_ERR_INVALID_ARG;
  return guarded(h, [&](cfb::Engine<BackendT>& e) {
    CFB_NEED_MESH(e);
    cfb::Halo* H = haloOf(h, false);
    if (!H || !H->comm) throw std::runtime_error("comm_init has not been called");
    const int n = H->nRanks, q = e.q();
    H->sendPtr.assign(sendPtr, sendPtr + n + 1); H->recvPtr.assign(recvPtr, recvPtr + n + 1);
    const int ns = H->sendPtr[n], nr = H->recvPtr[n];
    if ((ns && !sendCells) || (nr && !recvCells)) throw std::invalid_argument("set_halo: null cell list");
    for (int i = 0; i < ns; ++i) if (sendCells[i] < 0 || sendCells[i] >= e.d.nCells) throw std::invalid_argument("set_halo: send cell out of range");
    for (int i = 0; i < nr; ++i) if (recvCells[i] < 0 || recvCells[i] >= e.d.nCells) throw std::invalid_argument("set_halo: receive cell out of range");
    if (nIndependent < 0 || nIndependent > e.d.nCells) throw std::invalid_argument("set_halo: nIndependent out of range");
    cudaFree(H->dSend); cudaFree(H->dRecv); cudaFree(H->sbuf); cudaFree(H->rbuf);
    H->dSend = H->dRecv = nullptr; H->sbuf = H->rbuf = nullptr;
    CFB_CUDA(cudaMalloc((void**)&H->dSend, sizeof(int)*std::max(ns, 1)));
    CFB_CUDA(cudaMalloc((void**)&H->dRecv, sizeof(int)*std::max(nr, 1)));
    CFB_CUDA(cudaMalloc((void**)&H->sbuf, sizeof(double)*std::max(ns, 1)*q));
    CFB_CUDA(cudaMalloc((void**)&H->rbuf, sizeof(double)*std::max(nr, 1)*q));
    std::vector<int32_t> sc(ns), rc(nr);                 // the caller's cell IDs -> storage IDs (cfg.renumber)
    for (int i = 0; i < ns; ++i) sc[i] = e.cinv(sendCells[i]);
    for (int i = 0; i < nr; ++i) rc[i] = e.cinv(recvCells[i]);
    if (ns) CFB_CUDA(cudaMemcpy(H->dSend, sc.data(), sizeof(int)*ns, cudaMemcpyHostToDevice));
    if (nr) CFB_CUDA(cudaMemcpy(H->dRecv, rc.data(), sizeof(int)*nr, cudaMemcpyHostToDevice));
    H->haveLists = true;
    e.nIndep = nIndependent;
    return CFB200_OK;
  });
}

int cfb200_set_independent_cells(cfb200_handle* h, const int32_t* cells, int32_t n) {
  if (n < 0 || (n > 0 && !cells)) return CFB200_ERR_INVALID_ARG;
  return guarded(h, [&](cfb::Engine<BackendT>& e) {
    CFB_NEED_MESH(e);
    e.setIndependentCells(cells, n);
    return CFB200_OK;
  });
}

int cfb200_halo_begin(cfb200_handle* h) {
  return guarded(h, [&](cfb::Engine<BackendT>& e) {
    cfb::Halo* H = haloOf(h, false);
    if (!H) throw std::runtime_error("comm_init has not been called");
    haloBegin(e, *H);
    return CFB200_OK;
  });
}

int cfb200_halo_exchange(cfb200_handle* h) {
  return guarded(h, [&](cfb::Engine<BackendT>& e) {
    cfb::Halo* H = haloOf(h, false);
    if (!H) throw std::runtime_error("comm_init has not been called");
    haloBegin(e, *H);
    e.joinHalo();
    e.bk.sync();
    return CFB200_OK;
  });
}

int cfb200_allreduce_sum(cfb200_handle* h, double* v, int n) {
  if (!v || n < 1 || n > 64) return CFB200_ERR_INVALID_ARG;
  return guarded(h, [&](cfb::Engine<BackendT>& e) {
    cfb::Halo* H = haloOf(h, false);
    if (!H || !H->comm) throw std::runtime_error("comm_init has not been called");
    CFB_CUDA(cudaMemcpyAsync(H->red, v, sizeof(double)*n, cudaMemcpyHostToDevice, e.bk.stream));
    CFB_NCCL(cfb::NcclApi::get().AllReduce(H->red, H->red, (size_t)n, ncclDouble, ncclSum, H->comm, e.bk.stream));
    CFB_CUDA(cudaMemcpyAsync(v, H->red, sizeof(double)*n, cudaMemcpyDeviceToHost, e.bk.stream));
    e.bk.sync();
    return CFB200_OK;
  });
}

} // extern "C"

// ---- device-side consumers of the block Jacobian (cfb_solver.cuh) ------------------------------
namespace {
struct SolverWork {      // scratch of cfb200_solve_gmres / cfb200_export_csr_device, owned by the handle's engine
  int* blockRow = nullptr;
  double* V = nullptr; double* w = nullptr; double* z = nullptr; double* x = nullptr; double* b = nullptr;
  double* tmp = nullptr; double* dinv = nullptr;
  int restart = 0; size_t n = 0;
};
const int* blockRows(cfb::Engine<BackendT>& e, SolverWork& W) {
  if (!W.blockRow) {
    std::vector<int32_t> br((size_t)e.nnzb);
    for (int c = 0; c < e.d.nCells; ++c)
      for (int k = e.hRowPtr[c]; k < e.hRowPtr[c+1]; ++k) br[k] = c;
    W.blockRow = e.upload(br);
  }
  return W.blockRow;
}
SolverWork& solverWork(cfb::Engine<BackendT>& e) {
  if (!e.solverExt) { e.solverExt = new SolverWork(); e.solverExtFree = [](void* p) { delete static_cast<SolverWork*>(p); }; }
  return *static_cast<SolverWork*>(e.solverExt);
}
template <class F> void launchN(cfb::Engine<BackendT>& e, const char* name, const F& f, long long n) {
  if (n > 2147483647LL) throw std::invalid_argument("vector too long for one launch");
  e.bk.launch(name, f, (int)n, 256);
}
}

extern "C" {

// scalar CSR written on the device into caller-provided DEVICE arrays (rowOff [nCells*q + 1], col / val [nnzb*q*q]); any
// of them may be NULL.  index64 != 0: 64-bit offsets and columns (needed above 2^31 scalar entries, e.g. 220^3 NS)
int cfb200_export_csr_device(cfb200_handle* h, void* rowOff, void* col, double* val, int index64) {
  return guarded(h, [&](cfb::Engine<BackendT>& e) {
    CFB_NEED_MESH(e);
    if (!e.d.bsrVal) throw std::runtime_error("no Jacobian has been computed");
    const int q = e.q();
    const long long total = e.nnzb*q*q;
    if (!index64 && (total > 2147483647LL || (long long)e.d.nCells*q > 2147483647LL))
      throw std::invalid_argument("export_csr_device: scalar nnz exceeds 32-bit offsets: pass index64 = 1");
    const int* br = blockRows(e, solverWork(e));
    const long long CH = 1LL << 30;
    for (long long b0 = 0; b0 < total; b0 += CH) {
      const int n = (int)std::min(CH, total - b0);
      if (index64) e.bk.launch("bsrToCsr", cfb::BsrToCsr<long long>{e.d.bsrRowPtr, e.d.bsrColInd, br, e.d.bsrVal, q, b0,
                                   (long long*)rowOff, (long long*)col, val}, n, 256);
      else e.bk.launch("bsrToCsr", cfb::BsrToCsr<int>{e.d.bsrRowPtr, e.d.bsrColInd, br, e.d.bsrVal, q, b0,
                                   (int*)rowOff, (int*)col, val}, n, 256);
    }
    if (rowOff) {      // the closing offset
      if (index64) { const long long t = total; CFB_CUDA(cudaMemcpyAsync((long long*)rowOff + (size_t)e.d.nCells*q, &t, 8, cudaMemcpyHostToDevice, e.bk.stream)); }
      else { const int t = (int)total; CFB_CUDA(cudaMemcpyAsync((int*)rowOff + (size_t)e.d.nCells*q, &t, 4, cudaMemcpyHostToDevice, e.bk.stream)); }
    }
    e.bk.sync();
    return CFB200_OK;
  });
}

// J x = rhs on the device: restarted GMRES, right-preconditioned with the inverted diagonal blocks (precond = 1) or not
// (0).  The right-hand side is the device rhs buffer (what compute_rhs [+ the time term] left there), the solution
// overwrites it, so cfb200_update_sol_newton(h, NULL, ...) follows directly.  Rows of non-updatable cells are excluded.
int cfb200_solve_gmres(cfb200_handle* h, const cfb200_gmres_params* prm, int32_t* iterations, double* relResidual) {
  if (!prm || prm->restart < 1 || prm->restart > 200 || prm->maxIter < 1) return CFB200_ERR_INVALID_ARG;
  return guarded(h, [&](cfb::Engine<BackendT>& e) {
    CFB_NEED_MESH(e);
    if (!e.d.bsrVal) throw std::runtime_error("no Jacobian has been computed");
    e.joinHalo();
    const int q = e.q(), m = prm->restart;
    const long long n = (long long)e.d.nCells*q;
    SolverWork& W = solverWork(e);
    if (W.restart < m || W.n != (size_t)n) {
      if (W.V) throw std::runtime_error("solve_gmres: the restart length cannot grow after the first solve");
      W.V = e.template dalloc<double>((size_t)n*(m + 1));
      W.w = e.template dalloc<double>(n); W.z = e.template dalloc<double>(n); W.x = e.template dalloc<double>(n);
      W.b = e.template dalloc<double>(n);
      W.tmp = e.template dalloc<double>((size_t)n + n/1024 + 4096);
      W.dinv = e.template dalloc<double>((size_t)e.d.nCells*q*q);
      W.restart = m; W.n = (size_t)n;
    }
    auto& bk = e.bk;
    auto dot = [&](const double* a, const double* b) {
      launchN(e, "gmresDot", cfb::VecMul{a, b, W.tmp}, n);
      return bk.sumOrdered(W.tmp, (int)n);
    };
    auto spmv = [&](const double* x, double* y) { launchN(e, "gmresSpmv", cfb::BsrSpmv{e.d.bsrRowPtr, e.d.bsrColInd, e.d.bsrVal, x, y, q}, n); };
    auto prec = [&](const double* x, double* y) { launchN(e, "gmresPrec", cfb::ApplyDiagInv{W.dinv, x, y, q, prm->precond}, n); };
    if (prm->precond) bk.launch("gmresInvDiag", cfb::InvertDiag{e.d.bsrRowPtr, e.d.bsrColInd, e.dUpdatableByRow, e.d.bsrVal, W.dinv, q}, e.d.nCells, 128);
    bk.d2d(W.b, e.d.rhs, sizeof(double)*n);
    launchN(e, "gmresMask", cfb::VecZeroMasked{e.dUpdatableByRow, W.b, q}, n);
    bk.memset(W.x, 0, sizeof(double)*n);
    bk.d2d(W.w, W.b, sizeof(double)*n);                 // r0 = b (x0 = 0)
    double beta = std::sqrt(dot(W.w, W.w));
    const double beta0 = beta;
    const double tol = std::max(prm->relTol*beta0, prm->absTol);
    int it = 0;
    std::vector<double> H((size_t)(m + 1)*m), cs(m), sn(m), g(m + 1), y(m);
    bool done = (beta0 == 0.0);
    while (!done) {
      launchN(e, "gmresScale", cfb::VecScaleTo{1.0/beta, W.w, W.V}, n);
      std::fill(g.begin(), g.end(), 0.); g[0] = beta;
      int j = 0;
      for (; j < m; ++j) {
        double* vj = W.V + (size_t)j*n;
        prec(vj, W.z);
        spmv(W.z, W.w);
        for (int i = 0; i <= j; ++i) {                  // modified Gram-Schmidt
          const double hij = dot(W.w, W.V + (size_t)i*n);
          H[(size_t)i*m + j] = hij;
          launchN(e, "gmresAxpy", cfb::VecAxpy{-hij, W.V + (size_t)i*n, W.w}, n);
        }
        const double hn = std::sqrt(dot(W.w, W.w));
        H[(size_t)(j + 1)*m + j] = hn;
        if (hn > 0.) launchN(e, "gmresScale", cfb::VecScaleTo{1.0/hn, W.w, W.V + (size_t)(j + 1)*n}, n);
        for (int i = 0; i < j; ++i) {                   // earlier rotations on the new column
          const double t = cs[i]*H[(size_t)i*m + j] + sn[i]*H[(size_t)(i + 1)*m + j];
          H[(size_t)(i + 1)*m + j] = -sn[i]*H[(size_t)i*m + j] + cs[i]*H[(size_t)(i + 1)*m + j];
          H[(size_t)i*m + j] = t;
        }
        const double a = H[(size_t)j*m + j], b2 = H[(size_t)(j + 1)*m + j], rr = std::sqrt(a*a + b2*b2);
        cs[j] = (rr > 0.) ? a/rr : 1.; sn[j] = (rr > 0.) ? b2/rr : 0.;
        H[(size_t)j*m + j] = rr; H[(size_t)(j + 1)*m + j] = 0.;
        g[j + 1] = -sn[j]*g[j]; g[j] = cs[j]*g[j];
        ++it;
        if (std::abs(g[j + 1]) <= tol || it >= prm->maxIter || hn == 0.) { ++j; break; }
      }
      const int k = j;                                   // columns used
      for (int i = k - 1; i >= 0; --i) {                 // back substitution
        double sum = g[i];
        for (int l = i + 1; l < k; ++l) sum -= H[(size_t)i*m + l]*y[l];
        y[i] = sum/H[(size_t)i*m + i];
      }
      bk.memset(W.w, 0, sizeof(double)*n);
      for (int i = 0; i < k; ++i) launchN(e, "gmresAxpy", cfb::VecAxpy{y[i], W.V + (size_t)i*n, W.w}, n);
      prec(W.w, W.z);
      launchN(e, "gmresAxpy", cfb::VecAxpy{1.0, W.z, W.x}, n);
      // true residual for the restart / the report
      spmv(W.x, W.w);
      launchN(e, "gmresScale", cfb::VecScaleTo{-1.0, W.w, W.w}, n);
      launchN(e, "gmresAxpy", cfb::VecAxpy{1.0, W.b, W.w}, n);
      beta = std::sqrt(dot(W.w, W.w));
      done = (beta <= tol) || it >= prm->maxIter;
    }
    bk.d2d(e.d.rhs, W.x, sizeof(double)*n);
    bk.sync();
    if (iterations) *iterations = it;
    if (relResidual) *relResidual = (beta0 > 0.) ? beta/beta0 : 0.;
    return CFB200_OK;
  });
}

} // extern "C"

// ---- self test of the branch-free divide / square root (cfb_math.cuh) against the compiler's IEEE `/` and sqrt() ----
namespace cfb {
__device__ __forceinline__ unsigned long long splitmix(unsigned long long x) {
  x += 0x9e3779b97f4a7c15ull; x = (x ^ (x >> 30))*0xbf58476d1ce4e5b9ull; x = (x ^ (x >> 27))*0x94d049bb133111ebull; return x ^ (x >> 31);
}
// operand with a random 52-bit significand, random sign and an exponent uniform in [-span, span]
__device__ __forceinline__ double randOperand(unsigned long long h, int span) {
  const unsigned long long mant = h & 0x000fffffffffffffull;
  const int e = (int)((h >> 52) % (unsigned)(2*span + 1)) - span;
  const unsigned long long sign = (h >> 63) << 63;
  return __longlong_as_double((long long)(sign | ((unsigned long long)(1023 + e) << 52) | mant));
}
__global__ void k_selftest_divsqrt(unsigned long long seed, long long n, int span, unsigned long long* bad) {
  const long long i = (long long)blockIdx.x*blockDim.x + threadIdx.x;
  if (i >= n) return;
  const unsigned long long h0 = splitmix(seed + 3ull*i), h1 = splitmix(seed + 3ull*i + 1), h2 = splitmix(seed + 3ull*i + 2);
  double a = randOperand(h0, span), b = randOperand(h1, span);
  if ((h2 & 1023) == 0) a = 0.0;
  if ((h2 & 1023) == 1) a = -0.0;
  if ((h2 & 1023) == 2) a = b;                       // exact quotients
  if ((h2 & 1023) == 3) b = (double)(1 + (h2 >> 40) % 1000);
  unsigned long long nbad = 0;
  const Recip r = recipOf(b);
  const double q0 = a/b, q1 = divBy(a, r);
  if (__double_as_longlong(q0) != __double_as_longlong(q1)) ++nbad;
  const double c = a*1.0000001 + b;                  // a second numerator for the shared reciprocal
  const double q2 = c/b, q3 = divBy(c, r);
  if (__double_as_longlong(q2) != __double_as_longlong(q3)) ++nbad;
  const double o0 = 1.0/b, o1 = ddiv(1.0, b);
  if (__double_as_longlong(o0) != __double_as_longlong(o1)) ++nbad;
  const double x = fabs(a);
  const double s0 = sqrt(x), s1 = dsqrt(x);
  if (__double_as_longlong(s0) != __double_as_longlong(s1)) ++nbad;
  if (nbad) atomicAdd(bad, nbad);
}
}

// ---- allocation-free residual norm (cfb200_l2_norm_noalloc) ----------------------------------------------------------
// No counterpart in the reference: added for callers that cannot call cudaMalloc
// (CUDA graph capture, tight serving loops).  The block partial sums that Engine::sumOrdered keeps behind its
// cudaMalloc'ed `partial` array live here in a file-scope device array instead; the reduction shape is the same
// (1024 consecutive cells per block, 4 strided loads per thread, shared-memory tree, partials summed in index order), so
// the sum of squares equals the one cfb200_update_sol_newton forms, bit for bit.
#define CFB_NOALLOC_L1 16384                      /* level-1 partials: CFB200_NOALLOC_MAX_CELLS / 1024 */
#define CFB_NOALLOC_L2 16                         /* level-2 partials: CFB_NOALLOC_L1 / 1024 */
static_assert((long long)CFB_NOALLOC_L1*1024 == CFB200_NOALLOC_MAX_CELLS, "scratch size and the documented maximum differ");
static __device__ __align__(256) double g_noallocScratch[CFB_NOALLOC_L1 + CFB_NOALLOC_L2];   // 131 200 B per GPU, until exit

namespace cfb {

// squares of one equation's increment, reduced like k_block_sum reduces Dev::partial (UpdateSolNewton: sq = dU*dU for
// updatable cells, 0 for the others)
__global__ void __launch_bounds__(256) k_block_sumsq(const double* __restrict__ dU, const uint8_t* __restrict__ updatable,
                                                     double* __restrict__ out, int n, int q, int var) {
  __shared__ double s[256];
  const int base = blockIdx.x*1024;
  double a = 0.;
  for (int k = 0; k < 4; ++k) {
    const int i = base + k*256 + threadIdx.x;
    if (i < n) {
      double sq = 0.;
      if (!updatable || updatable[i]) { const double v = dU[(size_t)i*q + var]; sq = v*v; }
      a += sq;
    }
  }
  s[threadIdx.x] = a;
  __syncthreads();
  for (int w = 128; w > 0; w >>= 1) { if (threadIdx.x < w) s[threadIdx.x] += s[threadIdx.x + w]; __syncthreads(); }
  if (threadIdx.x == 0) out[blockIdx.x] = s[0];
}
// the host loop of Engine::sumOrdered (partials added in index order) and the log10 of cfb200_update_sol_newton
__global__ void k_noalloc_final(const double* __restrict__ partials, int m, double* __restrict__ out) {
  if (blockIdx.x != 0 || threadIdx.x != 0) return;
  double sum = 0.;
  for (int i = 0; i < m; ++i) sum += partials[i];
  out[0] = sum;
  out[1] = (sum > 0.) ? log10(sqrt(sum)) : -1.7976931348623157e308;
}

// device address of g_noallocScratch per device: in host code the array's name is only a stand-in for the symbol, the
// address a kernel can use comes from cudaGetSymbolAddress, asked once per device with that device current
static std::mutex g_noallocMutex;
static double* g_noallocAddr[64] = {};

static int noallocFail(char* err, int cap, int code, const std::string& what) {
  if (err && cap > 0) { std::snprintf(err, (size_t)cap, "%s", what.c_str()); }
  return code;
}

} // namespace cfb

extern "C" int cfb200_l2_norm_noalloc(int device, const double* d_dU, const uint8_t* d_updatable, int64_t nCells,
                                      int32_t nEqs, int32_t monitorVar, void* cudaStream, double* d_out,
                                      char* err, int32_t errCap) {
  using namespace cfb;
  if (err && errCap > 0) err[0] = 0;
  if (nCells < 0) return noallocFail(err, errCap, CFB200_ERR_INVALID_ARG, "cfb200_l2_norm_noalloc: nCells < 0");
  if (nCells > (int64_t)CFB200_NOALLOC_MAX_CELLS)
    return noallocFail(err, errCap, CFB200_ERR_INVALID_ARG,
                       "cfb200_l2_norm_noalloc: nCells = " + std::to_string((long long)nCells) + " exceeds the fixed scratch of "
                       "CFB200_NOALLOC_MAX_CELLS = " + std::to_string((long long)CFB200_NOALLOC_MAX_CELLS) +
                       " cells (use cfb200_update_sol_newton, whose scratch is sized with the mesh)");
  if (nEqs < 1 || nEqs > 1024 || monitorVar < 0 || monitorVar >= nEqs)
    return noallocFail(err, errCap, CFB200_ERR_INVALID_ARG, "cfb200_l2_norm_noalloc: nEqs / monitorVar out of range");
  if (device < 0 || device >= 64) return noallocFail(err, errCap, CFB200_ERR_INVALID_ARG, "cfb200_l2_norm_noalloc: device out of range");
  int count = 0;
  if (cudaGetDeviceCount(&count) != cudaSuccess || count <= 0) {
    cudaGetLastError();
    return noallocFail(err, errCap, CFB200_ERR_NO_DEVICE, "cfb200_l2_norm_noalloc: no usable CUDA device");
  }
  if (device >= count || cudaSetDevice(device) != cudaSuccess) {
    cudaGetLastError();
    return noallocFail(err, errCap, CFB200_ERR_INVALID_ARG, "cfb200_l2_norm_noalloc: cannot select the device");
  }
  double* scratch = nullptr;
  {
    std::lock_guard<std::mutex> lock(g_noallocMutex);
    if (!g_noallocAddr[device]) {
      double* p = nullptr;
      const cudaError_t e = cudaGetSymbolAddress((void**)&p, g_noallocScratch);
      if (e != cudaSuccess || !p) {
        cudaGetLastError();
        return noallocFail(err, errCap, CFB200_ERR_CUDA,
                           std::string("cfb200_l2_norm_noalloc: cudaGetSymbolAddress: ") + cudaGetErrorString(e));
      }
      g_noallocAddr[device] = p;
    }
    scratch = g_noallocAddr[device];
  }
  if (nCells == 0) return CFB200_OK;                 // set-up call: the address is resolved, nothing is launched
  if (!d_dU || !d_out) return noallocFail(err, errCap, CFB200_ERR_INVALID_ARG, "cfb200_l2_norm_noalloc: null input / output pointer");
  cudaStream_t st = (cudaStream_t)cudaStream;
  const int n = (int)nCells;
  const int nb = (n + 1023)/1024;
  // every partial the final kernel reads is written by this call: nothing has to start at zero
  k_block_sumsq<<<nb, 256, 0, st>>>(d_dU, d_updatable, scratch, n, (int)nEqs, (int)monitorVar);
  if (nb > 1024) {
    const int nb2 = (nb + 1023)/1024;
    k_block_sum<<<nb2, 256, 0, st>>>(scratch, scratch + CFB_NOALLOC_L1, nb);
    k_noalloc_final<<<1, 32, 0, st>>>(scratch + CFB_NOALLOC_L1, nb2, d_out);
  } else {
    k_noalloc_final<<<1, 32, 0, st>>>(scratch, nb, d_out);
  }
  const cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess)
    return noallocFail(err, errCap, CFB200_ERR_CUDA, std::string("cfb200_l2_norm_noalloc: launch: ") + cudaGetErrorString(e));
  return CFB200_OK;
}

extern "C" {

int cfb200_selftest_divsqrt(int device, int64_t n, uint64_t seed, int exponentSpan, int64_t* mismatches) {
  if (!mismatches || n <= 0 || exponentSpan < 0 || exponentSpan > 480) return CFB200_ERR_INVALID_ARG;
  int count = 0;
  if (cudaGetDeviceCount(&count) != cudaSuccess || count <= 0) { cudaGetLastError(); return CFB200_ERR_NO_DEVICE; }
  if (cudaSetDevice(device) != cudaSuccess) return CFB200_ERR_INVALID_ARG;
  unsigned long long* d = nullptr;
  if (cudaMalloc((void**)&d, 8) != cudaSuccess) return CFB200_ERR_CUDA;
  cudaMemset(d, 0, 8);
  cfb::k_selftest_divsqrt<<<(unsigned)((n + 255)/256), 256>>>(seed, n, exponentSpan, d);
  unsigned long long h = 0;
  const cudaError_t e = cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
  cudaFree(d);
  if (e != cudaSuccess) return CFB200_ERR_CUDA;
  *mismatches = (int64_t)h;
  return CFB200_OK;
}

int cfb200_abi_version(void) { return CFB200_ABI_VERSION; }

int cfb200_device_count(void) {
  int count = 0;
  if (cudaGetDeviceCount(&count) != cudaSuccess) { cudaGetLastError(); return 0; }
  return count;
}

} // extern "C"
