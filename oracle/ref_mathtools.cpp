// ORACLE — TEST INFRASTRUCTURE ONLY.
// Thin extern "C" harness around the REFERENCE's own header-only MathTools expression templates,
// compiled in place from /root/reference/src (never copied) into oracle/_ref/libcfref.so by
// oracle/Makefile.  It pins the evaluation order of the lazy expressions the FVMCC path relies on:
//   RoeFlux.cxx:162   result = 0.5*(_sumFlux - k*(_rightEv*(_absEvalues*_leftEv))*(stateR - stateL))
//   RoeFlux.cxx:200   _absEvalues = abs(_eValues)
//   NavierStokes3DVarSet.cxx:39  _flux.slice(_uID, dim) = _tau*normal
//   CFMat::determ3 / determ4 (VolumeCalculator)
// and lets tests compare the oracle's explicit loops against them bit for bit.
#include "MathTools/RealVector.hh"
#include "MathTools/RealMatrix.hh"
#include "MathTools/MathFunctions.hh"

using namespace COOLFluiD;
using namespace COOLFluiD::MathTools;

extern "C" {

void ref_roe_dissipation(int q, const double* R, const double* L, const double* ev, double kred,
                         const double* sumFlux, const double* sL, const double* sR, double* out) {
  RealMatrix _rightEv(q, q), _leftEv(q, q);
  RealVector _eValues(q), _absEvalues(q), _sumFlux(q), stateL(q), stateR(q), result(q);
  for (int i = 0; i < q; ++i) {
    for (int j = 0; j < q; ++j) { _rightEv(i, j) = R[i*q+j]; _leftEv(i, j) = L[i*q+j]; }
    _eValues[i] = ev[i]; _sumFlux[i] = sumFlux[i]; stateL[i] = sL[i]; stateR[i] = sR[i];
  }
  _absEvalues = abs(_eValues);
  result = 0.5*(_sumFlux - kred*(_rightEv*(_absEvalues*_leftEv))*(stateR - stateL));
  for (int i = 0; i < q; ++i) out[i] = result[i];
}

void ref_matvec(int n, const double* A, const double* x, double* y) {
  RealMatrix _tau(n, n); RealVector normal(n), r(n);
  for (int i = 0; i < n; ++i) { for (int j = 0; j < n; ++j) _tau(i, j) = A[i*n+j]; normal[i] = x[i]; }
  r = _tau*normal;
  for (int i = 0; i < n; ++i) y[i] = r[i];
}

double ref_det3(const double* m) { RealMatrix a(3, 3); for (int i = 0; i < 9; ++i) a[i] = m[i]; return a.determ3(); }
double ref_det4(const double* m) { RealMatrix a(4, 4); for (int i = 0; i < 16; ++i) a[i] = m[i]; return a.determ4(); }

double ref_inner_prod(int n, const double* a, const double* b) {
  RealVector x(n), y(n);
  for (int i = 0; i < n; ++i) { x[i] = a[i]; y[i] = b[i]; }
  return MathFunctions::innerProd(x, y);
}
double ref_distance(int n, const double* a, const double* b) {
  RealVector x(n), y(n);
  for (int i = 0; i < n; ++i) { x[i] = a[i]; y[i] = b[i]; }
  return MathFunctions::getDistance(x, y);
}
void ref_cross(const double* a, const double* b, double* c) {
  RealVector x(3), y(3), z(3);
  for (int i = 0; i < 3; ++i) { x[i] = a[i]; y[i] = b[i]; }
  MathFunctions::crossProd(x, y, z);
  for (int i = 0; i < 3; ++i) c[i] = z[i];
}
double ref_norm2(int n, const double* a) { RealVector x(n); for (int i = 0; i < n; ++i) x[i] = a[i]; return x.norm2(); }

} // extern "C"
