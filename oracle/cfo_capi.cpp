// ORACLE — TEST INFRASTRUCTURE ONLY (see cfo_types.hpp header).
// extern "C" surface of the CPU oracle, loaded with ctypes by tests/, smoke() and
// bench.py's cpu_baseline / --impl reference legs.
#include "cfo_solver.hpp"
#include "cfo_mesh.hpp"
#include <cstring>
#include <map>
#ifdef _OPENMP
#include <omp.h>
#endif

using namespace cfo;

namespace {
template <typename T> void cpy(void* dst, const std::vector<T>& v) { if (!v.empty()) std::memcpy(dst, v.data(), v.size()*sizeof(T)); }
}

extern "C" {

struct CfoOutputs {
  double* rhs; double* updateCoeff; double* uX; double* uY; double* uZ; double* limiter; double* nstates;
  const int32_t* bsrRowPtr; const int32_t* bsrColInd; double* bsrVal; int32_t status; double* divB;
};

void* cfo_mesh_build(int dim, int nNodes, const double* nodes, int nCells, const int32_t* cellNodePtr,
                     const int32_t* cellNodes, const int32_t* cellShape, int nTRS, const int32_t* trsFacePtr,
                     const int32_t* trsFaceCell, const int32_t* trsFaceNodePtr, const int32_t* trsFaceNodes) {
  MeshData* M = new MeshData();
  try {
    buildMesh(dim, nNodes, nodes, nCells, cellNodePtr, cellNodes, cellShape, nTRS, trsFacePtr, trsFaceCell,
              trsFaceNodePtr, trsFaceNodes, *M);
  } catch (const std::exception& e) { delete M; return nullptr; }
  return M;
}

// sizes: nFaces, nInner, nGhosts, len(faceNodes), len(cellFaces), len(stencil), len(nodeNbr), nnzb
void cfo_mesh_sizes(void* h, int64_t* out) {
  MeshData* M = (MeshData*)h;
  out[0] = M->nFaces; out[1] = M->nInner; out[2] = M->nGhosts; out[3] = (int64_t)M->faceNodes.size();
  out[4] = (int64_t)M->cellFaces.size(); out[5] = (int64_t)M->stencil.size(); out[6] = (int64_t)M->nodeNbr.size();
  out[7] = (int64_t)M->bsrColInd.size();
}

int cfo_mesh_get(void* h, const char* name, void* dst) {
  MeshData* M = (MeshData*)h;
  const std::string n(name);
#define G(x) if (n == #x) { cpy(dst, M->x); return 0; }
  G(centres) G(normals) G(areas) G(volumes) G(nodeW) G(cvol) G(faceCentres)
  G(faceL) G(faceR) G(faceNodePtr) G(faceNodes) G(faceGeoID) G(faceBC) G(cellFacePtr) G(cellFaces)
  G(stencilPtr) G(stencil) G(nodeNbrPtr) G(nodeNbr) G(trsPtr) G(cvSign) G(updatable) G(bsrRowPtr) G(bsrColInd)
#undef G
  return 1;
}

// Setup1.stencil = FaceVertex (1) | FaceEdge (2) | Face (3); the build leaves FaceVertexPlusGhost
int cfo_mesh_set_stencil(void* h, int kind) { return rebuildStencil(*(MeshData*)h, kind); }

void cfo_mesh_free(void* h) { delete (MeshData*)h; }

int cfo_apply_bc(const Config* c, const MeshView* m, const BC* bcs, double* U) {
  try {
    if (c->physicalModel == PM_MHD3D_PROJECTION) Solver<3, MhdPhys>(*c, *m, bcs).applyBC(U);
    else if (c->dim == 2) Solver<2>(*c, *m, bcs).applyBC(U); else Solver<3>(*c, *m, bcs).applyBC(U);
  } catch (const std::exception&) { return 2; }
  return 0;
}

// One ComputeRHS (+Jacobian) evaluation.  Ghost states in U must have been set (cfo_apply_bc).
int cfo_compute(const Config* c, const MeshView* m, const BC* bcs, double* U, CfoOutputs* co) {
  Outputs o{co->rhs, co->updateCoeff, co->uX, co->uY, co->uZ, co->limiter, co->nstates,
            co->bsrRowPtr, co->bsrColInd, co->bsrVal, 0, co->divB};
  try {
    if (c->physicalModel == PM_MHD3D_PROJECTION) {
      if (c->dim != 3 || c->nEqs != 9 || !(c->fluxSplitter == FS_LAXFRIED || c->fluxSplitter == FS_HLL || c->fluxSplitter == FS_CENTRED) ||
          c->diffusive != DIFF_NONE) { co->status = 2; return 2; }
      Solver<3, MhdPhys> s(*c, *m, bcs); s.execute(U, o);
    }
    else if (c->computeJacobian == 2 && (c->fluxSplitter != FS_AUSMPLUSUP ||
                                         (c->diffusive != DIFF_NONE && c->updateVar != UV_PVT) ||
                                         (c->diffusive != DIFF_NONE && c->derivativeStrategy == 1 && c->dim != 2) ||
                                         (c->diffusive != DIFF_NONE && c->derivativeStrategy == 2))) {
      co->status = 2; return 2;
    }
    else if (c->dim == 2) { Solver<2> s(*c, *m, bcs); s.execute(U, o); }
    else { Solver<3> s(*c, *m, bcs); s.execute(U, o); }
  } catch (const std::exception&) { co->status = 2; return 2; }
  co->status = o.status;
  return o.status;
}

int cfo_time_rhs(const Config* c, const MeshView* m, const BC* bcs, double* U, const double* pastU, CfoOutputs* co) {
  Outputs o{co->rhs, co->updateCoeff, co->uX, co->uY, co->uZ, co->limiter, co->nstates,
            co->bsrRowPtr, co->bsrColInd, co->bsrVal, 0, co->divB};
  try {
    if (c->physicalModel == PM_MHD3D_PROJECTION) Solver<3, MhdPhys>(*c, *m, bcs).pseudoSteadyTimeRhs(U, pastU, o);
    else if (c->dim == 2) Solver<2>(*c, *m, bcs).pseudoSteadyTimeRhs(U, pastU, o);
    else Solver<3>(*c, *m, bcs).pseudoSteadyTimeRhs(U, pastU, o);
  } catch (const std::exception&) { return 2; }
  return 0;
}

// Test hook: analytic AUSM+up flux Jacobian (Cons) of one face from the two CONSERVATIVE face states.
int cfo_ausm_jacob_cons(const Config* c, const double* uL, const double* uR, const double* n, int side, double* J) {
  if (c->dim == 2) { Phys<2>::Data a, b; Phys<2>::computePhysicalData(*c, uL, a); Phys<2>::computePhysicalData(*c, uR, b);
                     Phys<2>::ausmPlusUpJacobCons(*c, a, b, n, side, J); }
  else { Phys<3>::Data a, b; Phys<3>::computePhysicalData(*c, uL, a); Phys<3>::computePhysicalData(*c, uR, b);
         Phys<3>::ausmPlusUpJacobCons(*c, a, b, n, side, J); }
  return 0;
}
// Same in Puvt / Pvt variables (c->updateVar == UV_PVT): the two face states in UPDATE variables.
int cfo_ausm_jacob_pvt(const Config* c, const double* uL, const double* uR, const double* n, int side, double* J) {
  if (c->dim == 2) { Phys<2>::Data a, b; Phys<2>::computePhysicalData(*c, uL, a); Phys<2>::computePhysicalData(*c, uR, b);
                     Phys<2>::ausmPlusUpJacobPvt(*c, a, b, n, side, J); }
  else { Phys<3>::Data a, b; Phys<3>::computePhysicalData(*c, uL, a); Phys<3>::computePhysicalData(*c, uR, b);
         Phys<3>::ausmPlusUpJacobPvt(*c, a, b, n, side, J); }
  return 0;
}
// Test hook: analytic diffusive flux Jacobian of one side from the average state, d grad / d state, mu and lambda.
int cfo_ns_flux_jacob(const Config* c, const double* avState, const double* gj, const double* n, double mu, double tc,
                      double* J) {
  if (c->dim == 2) Phys<2>::nsFluxJacobian(*c, avState, gj, n, mu, tc, J);
  else Phys<3>::nsFluxJacobian(*c, avState, gj, n, mu, tc, J);
  return 0;
}
int cfo_ausm_flux(const Config* c, const double* uL, const double* uR, const double* n, double* F) {
  if (c->dim == 2) { Phys<2>::Data a, b; Phys<2>::computePhysicalData(*c, uL, a); Phys<2>::computePhysicalData(*c, uR, b);
                     Phys<2>::ausmFlux(*c, a, b, n, F); }
  else { Phys<3>::Data a, b; Phys<3>::computePhysicalData(*c, uL, a); Phys<3>::computePhysicalData(*c, uR, b);
         Phys<3>::ausmFlux(*c, a, b, n, F); }
  return 0;
}

// ForwardEuler driver (FwdEuler.cxx:146-220 + UpdateSol.cxx:83-193 + ComputeL2Norm.cxx:81-143 +
// NormCondition.cxx:55-65): iterate applyBC -> ComputeRHS -> U += CFL/updateCoeff * rhs -> norm until
// log10(L2(rhs[:,monitorVar])) < normTarget or maxIter.  history[k] = residual after iteration k+1.
// Returns the number of iterations done (negative on failure).
int cfo_run_fwd_euler(Config* c, const MeshView* m, const BC* bcs, double* U, CfoOutputs* co,
                      double cfl, int monitorVar, double normTarget, int maxIter, double* history) {
  const int q = c->nEqs, nc = m->nCells;
  c->residual = 0.0;             // SubSystemStatus.cxx:75
  int it = 0;
  for (; it < maxIter; ) {
    c->iter = it + 1;            // subSysStatus->updateNbIter() (FwdEuler.cxx:153)
    if (cfo_apply_bc(c, m, bcs, U)) return -2;
    if (cfo_compute(c, m, bcs, U, co)) return -1;
    double value = 0.0;
    for (int i = 0; i < nc; ++i) {
      if (!m->updatable[i]) { for (int j = 0; j < q; ++j) co->rhs[(size_t)i*q+j] = 0.; co->updateCoeff[i] = 0.0; continue; }
      const double uc = co->updateCoeff[i];
      const bool isZero = std::abs(uc) <= std::numeric_limits<double>::min();
      const double dt = isZero ? 0. : cfl/uc;
      for (int j = 0; j < q; ++j) { double rhsState = co->rhs[(size_t)i*q+j]; U[(size_t)i*q+j] += rhsState *= dt; }
      co->updateCoeff[i] = 0.0;
      const double tmp = co->rhs[(size_t)i*q + monitorVar];
      value += tmp*tmp;
    }
    const double res = (value > 0.) ? std::log10(std::sqrt(value)) : -std::numeric_limits<double>::max();
    history[it] = res;
    c->residual = res;
    ++it;
    if (it > 1 && res < normTarget) break;   // NormCondition: the first isAchieved() call always returns false
  }
  return it;
}

int cfo_num_threads() {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

// sizeof checks for the ctypes mirror
int cfo_sizeof_config() { return (int)sizeof(Config); }
int cfo_sizeof_bc() { return (int)sizeof(BC); }
int cfo_sizeof_meshview() { return (int)sizeof(MeshView); }

// exposed leaf functions for unit checks against oracle/_ref
void cfo_roe_dissipation(int q, const double* R, const double* L, const double* aev, double kred,
                         const double* sumFlux, const double* sL, const double* sR, double* result) {
  for (int i = 0; i < q; ++i) {
    double res = 0.;
    for (int k = 0; k < q; ++k) {
      double mm = 0.;
      for (int j = 0; j < q; ++j) mm += R[i*q+j]*(aev[j]*L[j*q+k]);
      res += (kred*mm)*(sR[k] - sL[k]);
    }
    result[i] = 0.5*(sumFlux[i] - res);
  }
}
double cfo_det3(const double* m) { return det3(m); }
double cfo_det4(const double* m) { return det4(m); }
double cfo_pow15(double x) { return cr_pow15(x); }
double cfo_pow3(double x) { return cr_pow3(x); }

} // extern "C"
