// cfb200 — the staged launch shape of the face kernels (one item per thread, shared-memory staging, warp-coalesced block
// write-out).  A header of its own because two translation units see it: cfb_capi.cu (every launch) and cfb_hot.cu, which
// holds the explicit instantiation of the interior-face FD-Jacobian kernel of the headline workload.
#pragma once
#include "cfb_kernels.cuh"

namespace cfb {

// One item per thread with a per-thread staging area in shared memory (entry k of thread t at sm[k*(BS+1) + t]: conflict
// free for the thread's own accesses AND for the warp's transposed read-out below).  After the item is evaluated the warp
// writes its 32 x 2 finished blocks to their BSR slots, F::BLK contiguous doubles at a time (coalesced), instead of 8-byte
// scattered stores.  No block-wide barrier: warps stay independent.
template <class F, int BS, int MINB>
__global__ void __launch_bounds__(BS, MINB) k_staged(const F f, const int n, const int first) {
  extern __shared__ double sm[];
  constexpr int STRIDE = BS + 1;
  const int t = threadIdx.x, lane = t & 31, warpBase = t & ~31;
  const int i = first + blockIdx.x*BS + t;
  int s0 = -1, s1 = -1;
  if (i < n) { f.run(i, sm + t, STRIDE); f.slots(i, s0, s1); }
  // the two destination slots of every face of the warp go to the (now free) loop-state rows of the staging area
  sm[(F::NOUT + 0)*STRIDE + t] = __hiloint2double(0, s0);
  sm[(F::NOUT + 1)*STRIDE + t] = __hiloint2double(0, s1);
  __syncwarp();
  double* out = f.out();
  // entry e = lane + 32*it of the warp's 32 x NOUT finished values: face lf = e / NOUT, k = e % NOUT, kept incrementally
  static_assert(F::NOUT >= 32 && F::STAGE >= F::NOUT + 2, "staging layout");
  // batches of WB entries: all shared-memory reads of a batch are issued before its first global store (one round trip
  // per batch instead of one per entry: with 8 resident warps per SM nothing else hides the latency)
  constexpr int WB = (F::NOUT % 10 == 0) ? 10 : 8;
  static_assert(F::NOUT % WB == 0, "write-out batch");
  int k = lane, lf = 0;
#pragma unroll 1
  for (int it = 0; it < F::NOUT; it += WB) {
    double x[WB]; size_t dst[WB]; bool on[WB];
#pragma unroll
    for (int j = 0; j < WB; ++j) {
      const int sel = (k >= F::BLK) ? 1 : 0;
      const int kk = k - sel*F::BLK;
      const int slot = __double2loint(sm[(F::NOUT + sel)*STRIDE + warpBase + lf]);
      x[j] = sm[k*STRIDE + warpBase + lf];
      on[j] = slot >= 0;
      dst[j] = (size_t)slot*F::BLK + kk;
      k += 32;
      if (k >= F::NOUT) { k -= F::NOUT; ++lf; }
    }
#pragma unroll
    for (int j = 0; j < WB; ++j) if (on[j]) out[dst[j]] = x[j];
  }
}

} // namespace cfb
