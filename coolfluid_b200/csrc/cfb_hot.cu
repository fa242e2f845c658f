// cfb200 — the interior-face finite-difference Jacobian kernel of the headline workload (3D Navier-Stokes, Roe, Pvt
// variables, DiamondVolume3D: 70 % of the step) as a translation unit of its own, compiled WITHOUT -split-compile.
// The same source as every other k_staged instance (cfb_staged.cuh, cfb_face.cuh); only where it is compiled differs:
// inside cfb_capi.cu its register allocation followed unrelated edits of that file (238 <-> 252 registers), here it is
// 251 registers whatever else the library holds.
#include <type_traits>
#include <cuda_runtime.h>
#include "cfb_staged.cuh"

namespace cfb {
template __global__ void k_staged<InnerFaceK<3, true, 0, 0, 1>, 128, 2>(const InnerFaceK<3, true, 0, 0, 1>, const int, const int);
} // namespace cfb
