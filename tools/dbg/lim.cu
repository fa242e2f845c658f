#include <cstdio>
#include <cstdlib>
#include <cmath>
#include "../../coolfluid_b200/csrc/cfb_math.cuh"
using namespace cfb;
__global__ void k(const double* in, double* out, int n) {
  int i = blockIdx.x*blockDim.x + threadIdx.x; if (i >= n) return;
  const double deltaMin = in[5*i], dpmax = in[5*i+1], dpmin = in[5*i+2], mag = in[5*i+3], eps2 = in[5*i+4];
  const Recip rmag = recipOf(mag);
  const bool pos = deltaMin > 0.0;
  const double dMinStar = ((deltaMin < 0.0) ? -1.0 : 1.0)*(divBy(fabs(deltaMin), rmag) + 1e-24);
  const double dPlus = divBy(pos ? dpmax : dpmin, rmag);
  const double dPlus2 = dPlus*dPlus;
  const double dPlusMin = dPlus*dMinStar;
  const double Num = (dPlus2 + eps2)*dMinStar + ((2*dMinStar)*dMinStar)*dPlus;
  const double Den = ((dPlus2 + (2*dMinStar)*dMinStar) + dPlusMin) + eps2;
  out[4*i] = ddiv(1., dMinStar)*ddiv(Num, Den);
  out[4*i+1] = dMinStar; out[4*i+2] = Num; out[4*i+3] = Den;
}
int main() {
  const int n = 1 << 16;
  double* h = (double*)malloc(sizeof(double)*5*n); double* o = (double*)malloc(sizeof(double)*4*n);
  srand(1);
  for (int i = 0; i < n; ++i) {
    h[5*i] = (rand()/(double)RAND_MAX - 0.5)*0.2; h[5*i+1] = rand()/(double)RAND_MAX*0.3; h[5*i+2] = -rand()/(double)RAND_MAX*0.3;
    h[5*i+3] = 0.2 + rand()/(double)RAND_MAX*2.5; h[5*i+4] = rand()/(double)RAND_MAX*1e-3;
  }
  double *di, *dout; cudaMalloc(&di, sizeof(double)*5*n); cudaMalloc(&dout, sizeof(double)*4*n);
  cudaMemcpy(di, h, sizeof(double)*5*n, cudaMemcpyHostToDevice);
  k<<<n/256, 256>>>(di, dout, n); cudaMemcpy(o, dout, sizeof(double)*4*n, cudaMemcpyDeviceToHost);
  int bad[4] = {0,0,0,0};
  for (int i = 0; i < n; ++i) {
    const double deltaMin = h[5*i], dpmax = h[5*i+1], dpmin = h[5*i+2], mag = h[5*i+3], eps2 = h[5*i+4];
    const double dMinStar = ((deltaMin < 0.0) ? -1.0 : 1.0)*(std::abs(deltaMin)/mag + 1e-24);
    const double dPlus = ((deltaMin > 0.0) ? dpmax : dpmin)/mag;
    const double dPlus2 = dPlus*dPlus; const double dPlusMin = dPlus*dMinStar;
    const double Num = (dPlus2 + eps2)*dMinStar + 2*dMinStar*dMinStar*dPlus;
    const double Den = (dPlus2 + 2*dMinStar*dMinStar + dPlusMin + eps2);
    const double psi = 1./dMinStar*(Num/Den);
    const double ref[4] = {psi, dMinStar, Num, Den};
    for (int j = 0; j < 4; ++j) if (ref[j] != o[4*i+j]) { if (bad[j] < 3) printf("i %d j %d gpu %.17g host %.17g (deltaMin %.17g mag %.17g)\n", i, j, o[4*i+j], ref[j], deltaMin, mag); ++bad[j]; }
  }
  printf("mismatches psi %d dMinStar %d Num %d Den %d of %d\n", bad[0], bad[1], bad[2], bad[3], n);
  return 0;
}
