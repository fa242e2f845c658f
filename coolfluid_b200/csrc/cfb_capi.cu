// cfb200 — the product library: CUDA backend of cfb::Engine + the extern "C" surface of include/cfb200.h.
// Compiled for sm_100a only (nvcc -gencode arch=compute_100a,code=sm_100a -fmad=false).  There is no CPU
// execution path in this library: without a usable CUDA device cfb200_create returns CFB200_ERR_NO_DEVICE.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <string>
#include <vector>
#include "cfb_engine.hpp"

namespace cfb {

#define CFB_CUDA(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) \
  throw DeviceError(CFB200_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_)); } while (0)

template <class F, int BS, int MINB>
__global__ void __launch_bounds__(BS, MINB) k_run(const F f, const int n) {
  const int i = blockIdx.x*BS + threadIdx.x;
  if (i < n) f(i);
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

struct CudaBackend {
  int dev = 0;
  cudaStream_t stream = nullptr;
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
  }
  ~CudaBackend() {
    cudaSetDevice(dev);
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
    if (n <= 0) return;
    Ev* ev = nullptr;
    if (timing) {
      if (nEv == evs.size()) { Ev e; e.name = name; CFB_CUDA(cudaEventCreate(&e.a)); CFB_CUDA(cudaEventCreate(&e.b)); evs.push_back(e); }
      ev = &evs[nEv++]; ev->name = name;
      CFB_CUDA(cudaEventRecord(ev->a, stream));
    }
    // bs encodes the launch shape: 256 -> 256 threads; 128 -> 128 threads, >= 2 blocks/SM (<= 255 registers);
    // 1283 / 1284 -> 128 threads with >= 3 / 4 resident blocks per SM (<= 168 / 128 registers)
    if (bs == 1283) k_run<F, 128, 3><<<(n + 127)/128, 128, 0, stream>>>(f, n);
    else if (bs == 1284) k_run<F, 128, 4><<<(n + 127)/128, 128, 0, stream>>>(f, n);
    else if (bs == 128) k_run<F, 128, 2><<<(n + 127)/128, 128, 0, stream>>>(f, n);
    else k_run<F, 256, 1><<<(n + 255)/256, 256, 0, stream>>>(f, n);
    CFB_CUDA(cudaGetLastError());
    if (ev) CFB_CUDA(cudaEventRecord(ev->b, stream));
  }

  // sum of x[0..n) with a fixed reduction shape; x must have >= ceil(n/1024) doubles of scratch after position n
  double sumOrdered(double* x, int n) {
    if (n <= 0) return 0.;
    const int nb = (n + 1023)/1024;
    double* out = x + n;                      // Engine allocates nCells + 1024 ... see below
    std::vector<double> h;
    if (nb > 1024) {
      // two levels: reduce the block sums again
      double* tmp = nullptr;
      CFB_CUDA(cudaMalloc((void**)&tmp, sizeof(double)*(nb + (nb + 1023)/1024)));
      k_block_sum<<<nb, 256, 0, stream>>>(x, tmp, n);
      const int nb2 = (nb + 1023)/1024;
      k_block_sum<<<nb2, 256, 0, stream>>>(tmp, tmp + nb, nb);
      h.resize(nb2);
      CFB_CUDA(cudaMemcpyAsync(h.data(), tmp + nb, sizeof(double)*nb2, cudaMemcpyDeviceToHost, stream));
      CFB_CUDA(cudaStreamSynchronize(stream));
      cudaFree(tmp);
    } else {
      k_block_sum<<<nb, 256, 0, stream>>>(x, out, n);
      h.resize(nb);
      CFB_CUDA(cudaMemcpyAsync(h.data(), out, sizeof(double)*nb, cudaMemcpyDeviceToHost, stream));
      CFB_CUDA(cudaStreamSynchronize(stream));
    }
    double s = 0.;
    for (double v : h) s += v;
    return s;
  }
};

} // namespace cfb

typedef cfb::CudaBackend BackendT;
#define CFB_FN(x) cfb200_##x
#include "cfb_capi.inc"

extern "C" {

int cfb200_abi_version(void) { return CFB200_ABI_VERSION; }

int cfb200_device_count(void) {
  int count = 0;
  if (cudaGetDeviceCount(&count) != cudaSuccess) { cudaGetLastError(); return 0; }
  return count;
}

} // extern "C"
