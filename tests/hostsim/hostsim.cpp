// TEST INFRASTRUCTURE ONLY — never linked into libcfb200.so, never imported by coolfluid_b200.
// Instantiates the product's kernel functors (coolfluid_b200/csrc/cfb_kernels.cuh) and engine sequencing
// (cfb_engine.hpp) with a host-loop backend: "thread idx" becomes a for loop.  It exists so that the CPU
// test tier (-m "not gpu", no GPU in the build container) can check the kernel LOGIC against the oracle
// before any GPU time is spent.  The symbols are prefixed hostsim_ so they can never be mistaken for, or
// shadow, the cfb200_ C ABI.
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <string>
#include <vector>
#include "../../coolfluid_b200/csrc/cfb_engine.hpp"

namespace cfb {
struct HostLoopBackend {
  explicit HostLoopBackend(int) {}
  void activate() {}
  void* alloc(size_t bytes) { return std::calloc(1, bytes); }
  void free(void* p) { std::free(p); }
  void h2d(void* d, const void* s, size_t n) { std::memcpy(d, s, n); }
  void d2h(void* d, const void* s, size_t n) { std::memcpy(d, s, n); }
  void d2d(void* d, const void* s, size_t n) { std::memcpy(d, s, n); }
  void d2hOverlapped(void* d, const void* s, size_t n) { std::memcpy(d, s, n); }
  void joinCopies() {}
  int winFirst = 0, winEnd = -1;
  void window(int first, int end) { winFirst = first; winEnd = end; }
  int takeWindow(int& n) { int w0 = 0; if (winEnd >= 0) { w0 = winFirst; if (winEnd < n) n = winEnd; } winFirst = 0; winEnd = -1; return w0; }
  void memset(void* p, int v, size_t n) { std::memset(p, v, n); }
  void sync() {}
  void* streamPtr() { return nullptr; }
  void enableTiming(bool) {}
  void resetTimings() {}
  int timings(const char**, float*, int) { return 0; }
  template <class F> void launch(const char*, const F& f, int n, int) { const int w0 = takeWindow(n); for (int i = w0; i < n; ++i) f(i); }
  template <class F> void launchStaged(const char*, const F& f, int n, int) {
    double stg[F::STAGE];
    const int w0 = takeWindow(n);
    for (int i = w0; i < n; ++i) {
      int s0 = -1, s1 = -1;
      f.run(i, stg, 1); f.slots(i, s0, s1);
      for (int k = 0; k < F::NOUT; ++k) {                   // the warp's block write-out
        const int slot = (k/F::BLK) ? s1 : s0;
        if (slot >= 0) f.out()[(size_t)slot*F::BLK + k%F::BLK] = stg[k];
      }
    }
  }
  template <class F> void launchCellStaged(const char*, const F& f, int n) {
    double stg[F::BLK];
    const int w0 = takeWindow(n);
    for (int i = w0; i < n; ++i) {
      f.run(i, stg, 1);
      const int slot = f.slot(i);
      if (slot >= 0) for (int k = 0; k < F::BLK; ++k) f.out()[(size_t)slot*F::BLK + k] = stg[k];
    }
  }
  template <class F> void launchPhased(const char*, const F& f, int n) {
    double stg[F::STAGE];
    const int w0 = takeWindow(n);
    for (int i = w0; i < n; ++i) {
      typename F::State st;
      const bool active = f.begin(i, st, stg, 1);
      for (int ph = 0; ph < 2; ++ph) {
        if (active) f.phase(ph, st, stg, 1);
        const int slot = f.slot(i, ph);
        if (slot >= 0) for (int k = 0; k < F::BLK; ++k) f.out()[(size_t)slot*F::BLK + k] = stg[k];
      }
    }
  }
  template <class F> void launchStaged2(const char*, const F& f, int n) {
    double stg[F::STAGE];
    const int w0 = takeWindow(n);
    for (int i = w0; i < n; ++i) {
      typename F::State st;
      for (int phase = 0; phase < 2; ++phase) {
        f.run(i, phase, st, stg, 1);
        const int slot = f.slot(i, phase);
        if (slot >= 0) for (int k = 0; k < F::BLK; ++k) f.out()[(size_t)slot*F::BLK + k] = stg[k];
      }
    }
  }
  double sumOrdered(double* x, int n) { double s = 0.; for (int i = 0; i < n; ++i) s += x[i]; return s; }
  double minOrdered(double* x, int n) { double m = x[0]; for (int i = 1; i < n; ++i) m = (x[i] < m) ? x[i] : m; return m; }
};
}

typedef cfb::HostLoopBackend BackendT;
#define CFB_FN(x) hostsim_##x
#include "../../coolfluid_b200/csrc/cfb_capi.inc"
