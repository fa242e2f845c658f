// cfb200 — the FVMCC ComputeRHS / ComputeRhsJacob kernels, written as per-item functors
// (`operator()(int idx)`: one cell, node, face or face-side per call) over the device
// structure-of-arrays `Dev`.  cfb_engine.cu launches each functor as a CUDA kernel with one
// thread per item.  Design:
//   * no floating-point atomics anywhere: faces write their flux / Jacobian blocks to face-owned
//     slots, cells GATHER them in the reference's face-loop order, so every sum is formed in exactly
//     the order of the sequential CPU loop (FVMCC_ComputeRHS.cxx:202-319) and is run-to-run deterministic;
//   * off-diagonal Jacobian blocks have exactly one writer (the face between the two cells) and are
//     stored straight into the BSR value array; the diagonal block of a cell is minus the sum of the
//     blocks in its COLUMN (block(n,c) for each face neighbour n) plus its boundary-face blocks, again
//     gathered in face order (FVMCC_ComputeRhsJacob.cxx:347-365: row R receives -1 x what row L receives);
//   * all reals FP64, SoA by component so that one-thread-per-item accesses coalesce.
#pragma once
#include "cfb_math.cuh"

namespace cfb {

struct Dev {
  int dim, nCells, nGhosts, nStates, nNodes, nFaces, nInner;
  // ---- static mesh data (uploaded once) ----
  const double* nodes;        // [d][nNodes]
  const double* centres;      // [d][nStates]   cells then ghosts
  const int* faceL; const int* faceR;
  const int* faceNodePtr; const int* faceNodes;
  const double* normals;      // [d][nFaces]
  const double* areas;        // [nFaces]
  const double* faceCentres;  // [d][nFaces]    face quadrature point = node average (FVMCC_PolyRec.hh:114-123)
  const double* volumes;
  const int* cellFacePtr; const int* cellFaces;    // reference local face order (limiter loop)
  const int* cellFaceSorted;  // same CSR; faces ascending in loop order; ~f when the cell is the face's R state
  const int* stencilPtr; const int* stencil;
  const int* lsqPtr; const int* lsqNbr;            // per-cell neighbour list in the reference's edge order
  const int* nodeNbrPtr; const int* nodeNbr; const double* nodeW;
  const double* cvol; const uint8_t* cvSign; const int* faceBC;
  const uint8_t* updatable;
  const uint8_t* cellActive;  // cell touches at least one face with an updatable side
  const BCond* bcs;
  double* lsqL;               // [6][nCells]  l11 l12 l13 l22 l23 l33
  // ---- block Jacobian ----
  const int* bsrRowPtr; const int* bsrColInd;
  const int* slotLR;          // [nInner] block (L,R)
  const int* slotRL;          // [nInner] block (R,L)
  const int* diagSlot;        // [nCells]
  double* bsrVal;             // [nnzb][q][q]
  double* bndBlocks;          // [nGhosts][q][q]  boundary-face contribution to the diagonal of the inner cell
  // ---- fields ----
  double* U;                  // [q][nStates]
  double* pastU;              // [q][nCells]
  double* uX; double* uY; double* uZ; double* lim;   // [q][nCells]
  double* nstates;            // [q][nNodes]
  double* faceFlux;           // [q][nFaces]  resFactor is NOT applied here
  double* ucL; double* ucR; double* ucD; double* faceMu;   // [nFaces]
  double* rhs;                // AoS [nCells][q] (consumer layout: PETSc/Paralution block vectors)
  double* updateCoeff;        // [nCells]
  double* aos;                // staging [max(nStates,nNodes)*q]
  double* partial;            // reduction scratch
  int* status;
};

// ------------------------------------------------------------------------------------------------
// layout conversion between the reference's AoS sockets and the device SoA
struct AosToSoa {   // idx over n*q
  const double* src; double* dst; int n, q, ld;
  CFB_HD void operator()(int idx) const { const int i = idx/q, v = idx - i*q; dst[(size_t)v*ld + i] = src[idx]; }
};
struct SoaToAos {
  const double* src; double* dst; int n, q, ld;
  CFB_HD void operator()(int idx) const { const int i = idx/q, v = idx - i*q; dst[idx] = src[(size_t)v*ld + i]; }
};
struct FaceCentres {   // idx over faces
  Dev d;
  CFB_HD void operator()(int f) const {
    const int b = d.faceNodePtr[f], e = d.faceNodePtr[f+1];
    const double ov = 1./(e - b);
    for (int k = 0; k < d.dim; ++k) {
      double x = 0.0;
      for (int j = b; j < e; ++j) x += d.nodes[(size_t)k*d.nNodes + d.faceNodes[j]]*ov;
      const_cast<double*>(d.faceCentres)[(size_t)k*d.nFaces + f] = x;
    }
  }
};

// ------------------------------------------------------------------------------------------------
// a16: boundary conditions, FVMCC_BaseBC.ci:144-187 executeOnTrs -> setGhostState(face)
template <int D>
CFB_HD void ghostState(const Cfg& c, const BCond& b, const double* nr /*area-weighted normal*/,
                       const double* inner, double* ghost) {
  constexpr int Q = D + 2;
  typedef Phys<D> P;
  switch (b.type) {
    case CFB200_BC_SUPER_OUTLET:
      for (int i = 0; i < Q; ++i) ghost[i] = inner[i];
      break;
    case CFB200_BC_DIRICHLET:
      for (int i = 0; i < Q; ++i) ghost[i] = b.mask[i] ? (2.*b.value[i] - inner[i]) : inner[i];
      break;
    case CFB200_BC_MIRROR_VELOCITY: {     // MirrorVelocity.cxx:71-111
      double vn = 0., area2 = 0.;
      for (int i = 0; i < D; ++i) { const double nxComp = nr[i]; vn += nxComp*inner[1+i]; area2 += nxComp*nxComp; }
      ghost[0] = inner[0];
      for (int i = 0; i < D; ++i) ghost[1+i] = inner[1+i] - 2.0*vn*nr[i]/area2;
      ghost[Q-1] = inner[Q-1];
      break;
    }
    default: {                            // CFB200_BC_MIRROR_EULER_CONS, MirrorEuler2D.cxx:56-94
      double n[3] = {0., 0., 0.};
      double s2;
      if (D == 2) s2 = nr[0]*nr[0] + nr[1]*nr[1]; else s2 = nr[0]*nr[0] + nr[1]*nr[1] + nr[2]*nr[2];
      const double invFaceLength = 1./sqrt(s2);
      for (int i = 0; i < D; ++i) n[i] = nr[i]*invFaceLength;
      PData di, dg;
      P::physicalData(c, inner, di);
      const double vn = P::vn(di, n);
      const double gamma = c.gamma;
      const double gammaDivGammaMinus1 = gamma/(gamma - 1.0);
      dg.rho = di.rho;
      dg.v[2] = 0.;
      for (int i = 0; i < D; ++i) dg.v[i] = di.v[i] - 2.0*vn*n[i];
      dg.V = di.V;
      dg.p = di.p;
      dg.H = (gammaDivGammaMinus1*dg.p + 0.5*dg.rho*dg.V*dg.V)/dg.rho;
      dg.A = sqrt(gamma*dg.p/dg.rho);
      dg.T = di.T;
      P::consFromPhysicalData(dg, ghost);
      break;
    }
  }
}

template <int D>
struct GhostFill {   // idx over boundary faces
  Dev d; Cfg c;
  CFB_HD void operator()(int g) const {
    constexpr int Q = D + 2;
    const int f = d.nInner + g;
    const int L = d.faceL[f];
    double inner[Q], ghost[Q], nr[D];
    for (int v = 0; v < Q; ++v) inner[v] = d.U[(size_t)v*d.nStates + L];
    for (int k = 0; k < D; ++k) nr[k] = d.normals[(size_t)k*d.nFaces + f];
    ghostState<D>(c, d.bcs[d.faceBC[f]], nr, inner, ghost);
    for (int v = 0; v < Q; ++v) d.U[(size_t)v*d.nStates + d.nCells + g] = ghost[v];
  }
};

// ------------------------------------------------------------------------------------------------
// a4/a5: least-squares gradients.  The reference accumulates edge by edge, each edge visited once from
// its lower-ID cell (LeastSquareP1PolyRec2D.cxx:100-134, 3D.cxx:104-160).  `lsqNbr` lists, per cell, the
// other end of every edge that touches it in exactly that global edge order, so the per-cell gather below
// forms the same sums term by term ((-dx)*(-du) == dx*du exactly, so the edge orientation is irrelevant).
template <int D>
CFB_HD double edgeWeight(const Dev& d, int c, int nb, double* dx) {
  double dist = 0.;
  for (int k = 0; k < D; ++k) {
    const double xc = d.centres[(size_t)k*d.nStates + c], xn = d.centres[(size_t)k*d.nStates + nb];
    const double diff = xc - xn;
    dist += diff*diff;
    dx[k] = xn - xc;
  }
  const double w = 1.0/sqrt(dist);
  for (int k = 0; k < D; ++k) dx[k] = w*dx[k];
  return w;
}

template <int D>
struct LsqSetup {   // LeastSquareP1PolyRec{2,3}D::updateWeights; idx over cells
  Dev d;
  CFB_HD void operator()(int c) const {
    double l[6] = {0., 0., 0., 0., 0., 0.};
    for (int j = d.lsqPtr[c]; j < d.lsqPtr[c+1]; ++j) {
      double dx[3] = {0., 0., 0.};
      edgeWeight<D>(d, c, d.lsqNbr[j], dx);
      if (D == 2) { l[0] += dx[0]*dx[0]; l[1] += dx[0]*dx[1]; l[3] += dx[1]*dx[1]; }
      else {
        l[0] += dx[0]*dx[0]; l[1] += dx[0]*dx[1]; l[2] += dx[0]*dx[2];
        l[3] += dx[1]*dx[1]; l[4] += dx[1]*dx[2]; l[5] += dx[2]*dx[2];
      }
    }
    for (int k = 0; k < 6; ++k) d.lsqL[(size_t)k*d.nCells + c] = l[k];
  }
};

template <int D>
struct Gradients {   // idx over cells
  Dev d;
  CFB_HD void operator()(int c) const {
    constexpr int Q = D + 2;
    double lf[3][Q];
    double u0[Q];
    for (int v = 0; v < Q; ++v) { lf[0][v] = lf[1][v] = lf[2][v] = 0.; u0[v] = d.U[(size_t)v*d.nStates + c]; }
    for (int j = d.lsqPtr[c]; j < d.lsqPtr[c+1]; ++j) {
      const int nb = d.lsqNbr[j];
      double dx[3] = {0., 0., 0.};
      const double w = edgeWeight<D>(d, c, nb, dx);
      for (int v = 0; v < Q; ++v) {
        const double du = w*(d.U[(size_t)v*d.nStates + nb] - u0[v]);
        lf[0][v] += dx[0]*du; lf[1][v] += dx[1]*du;
        if (D == 3) lf[2][v] += dx[2]*du;
      }
    }
    if (D == 2) {
      const double l11 = d.lsqL[c], l12 = d.lsqL[(size_t)d.nCells + c], l22 = d.lsqL[(size_t)3*d.nCells + c];
      const double invDet = 1./(l11*l22 - l12*l12);
      for (int v = 0; v < Q; ++v) {
        d.uX[(size_t)v*d.nCells + c] = (l22*lf[0][v] - l12*lf[1][v])*invDet;
        d.uY[(size_t)v*d.nCells + c] = (l11*lf[1][v] - l12*lf[0][v])*invDet;
      }
    } else {
      const size_t n = d.nCells;
      const double l11 = d.lsqL[c], l12 = d.lsqL[n + c], l13 = d.lsqL[2*n + c];
      const double l22 = d.lsqL[3*n + c], l23 = d.lsqL[4*n + c], l33 = d.lsqL[5*n + c];
      const double det = l11*l22*l33 - l11*l23*l23 - l12*l12*l33 + l12*l13*l23 + l13*l12*l23 - l13*l13*l22;
      const double linv11 = l22*l33 - l23*l23;
      const double linv22 = l11*l33 - l13*l13;
      const double linv33 = l11*l22 - l12*l12;
      const double linv12 = -(l12*l33 - l13*l23);
      const double linv13 = l12*l23 - l13*l22;
      const double linv23 = -(l11*l23 - l13*l12);
      const bool ok = !(fabs(det) <= 2.2250738585072014e-308);   // !MathChecks::isZero(det)
      for (int v = 0; v < Q; ++v) {
        double gx = 0.0, gy = 0.0, gz = 0.0;
        if (ok) {
          gx = (linv11*lf[0][v] + linv12*lf[1][v] + linv13*lf[2][v])/det;
          gy = (linv12*lf[0][v] + linv22*lf[1][v] + linv23*lf[2][v])/det;
          gz = (linv13*lf[0][v] + linv23*lf[1][v] + linv33*lf[2][v])/det;
        }
        d.uX[(size_t)v*n + c] = gx; d.uY[(size_t)v*n + c] = gy; d.uZ[(size_t)v*n + c] = gz;
      }
    }
  }
};

// ------------------------------------------------------------------------------------------------
// a6: nodal states, DistanceBasedExtrapolator.ci:105-149; idx over nodes
template <int D>
struct NodalExtrapolate {
  Dev d;
  CFB_HD void operator()(int n) const {
    constexpr int Q = D + 2;
    double ns[Q];
    for (int v = 0; v < Q; ++v) ns[v] = 0.;
    for (int k = d.nodeNbrPtr[n]; k < d.nodeNbrPtr[n+1]; ++k) {
      const double w = d.nodeW[k];
      const int s = d.nodeNbr[k];
      for (int v = 0; v < Q; ++v) ns[v] += d.U[(size_t)v*d.nStates + s]*w;
    }
    for (int v = 0; v < Q; ++v) d.nstates[(size_t)v*d.nNodes + n] = ns[v];
  }
};

// ------------------------------------------------------------------------------------------------
// a7: Venkatakrishnan / Barth-Jespersen limiter, evaluated eagerly per cell (the reference evaluates it at
// the first face that touches the cell, FVMCC_PolyRec.cxx:120-167; all its inputs are frozen before the face
// loop, so the value is the same).  Venktn2D.cxx:111-260, Venktn3D.cxx:44-190, BarthJesp.cxx:60-155.
template <int D>
struct LimiterK {   // idx over cells
  Dev d; Cfg c;
  CFB_HD void operator()(int cell) const {
    constexpr int Q = D + 2;
    if (!d.cellActive[cell]) return;
    const bool fresh = (c.residual > c.limitRes) && (c.limitIter > 0 && c.iter < c.limitIter);
    if (!fresh && c.freezeLimiter) return;
    const int fb = d.cellFacePtr[cell], fe = d.cellFacePtr[cell+1];
    const int nbFaces = fe - fb;
    double xc[3] = {0., 0., 0.};
    for (int k = 0; k < D; ++k) xc[k] = d.centres[(size_t)k*d.nStates + cell];
    double dxq[6][3];       // face quadrature point - cell centre
    for (int j = 0; j < nbFaces; ++j) {
      const int f = d.cellFaces[fb + j];
      for (int k = 0; k < D; ++k) dxq[j][k] = d.faceCentres[(size_t)k*d.nFaces + f] - xc[k];
    }
    for (int v = 0; v < Q; ++v) {
      const double u0 = d.U[(size_t)v*d.nStates + cell];
      double min0 = u0, max0 = u0, avgDistance = 0.0;
      if (c.useFullStencil) {
        const int sb = d.stencilPtr[cell], se = d.stencilPtr[cell+1];
        for (int is = sb; is < se; ++is) {
          const int s = d.stencil[is];
          if (s != cell) {
            const double us = d.U[(size_t)v*d.nStates + s];
            min0 = dmin(min0, us); max0 = dmax(max0, us);
            double dist = 0.;
            for (int k = 0; k < D; ++k) { const double diff = xc[k] - d.centres[(size_t)k*d.nStates + s]; dist += diff*diff; }
            avgDistance += sqrt(dist);
          }
        }
        avgDistance /= (se - sb);
      } else {
        for (int j = 0; j < nbFaces; ++j) {
          const int f = d.cellFaces[fb + j];
          const int fl = d.faceL[f];
          const int s = (fl == cell) ? d.faceR[f] : fl;
          const double us = d.U[(size_t)v*d.nStates + s];
          min0 = dmin(min0, us); max0 = dmax(max0, us);
          double dist = 0.;
          for (int k = 0; k < D; ++k) { const double diff = xc[k] - d.centres[(size_t)k*d.nStates + s]; dist += diff*diff; }
          avgDistance += sqrt(dist);
        }
        avgDistance /= nbFaces;
      }
      const double deltaPlusMax = max0 - u0;
      const double deltaPlusMin = min0 - u0;
      double psi = 1.0;       // carried over from face to face when deltaMin == 0 (reference quirk)
      double psimin = 1.1;
      const double ux = d.uX[(size_t)v*d.nCells + cell], uy = d.uY[(size_t)v*d.nCells + cell];
      const double uz = (D == 3) ? d.uZ[(size_t)v*d.nCells + cell] : 0.;
      if (c.limiter == CFB200_LIM_VENKTN) {
        const double epsilon = c.coeffEps*c.magnitudeValues[v]*c.magnitudeValues[v]*pow3(avgDistance/c.length);
        for (int j = 0; j < nbFaces; ++j) {
          double deltaMin;
          if (D == 2) deltaMin = ux*dxq[j][0] + uy*dxq[j][1];
          else deltaMin = ux*dxq[j][0] + uy*dxq[j][1] + uz*dxq[j][2];
          if (deltaMin > 0.0) {
            const double dPlusMax2 = deltaPlusMax*deltaPlusMax;
            const double dPlusMaxMin = deltaPlusMax*deltaMin;
            psi = (dPlusMax2 + 2.0*dPlusMaxMin + epsilon)/(dPlusMax2 + dPlusMaxMin + 2.0*deltaMin*deltaMin + epsilon);
          }
          if (deltaMin < 0.0) {
            const double y = deltaPlusMin/deltaMin;
            psi = (y*y + 2*y)/(y*y + y + 2.);
          }
          psimin = dmin(psi, psimin);
        }
      } else {
        for (int j = 0; j < nbFaces; ++j) {
          double deltaMin;
          if (D == 2) deltaMin = ux*dxq[j][0] + uy*dxq[j][1];
          else deltaMin = ux*dxq[j][0] + uy*dxq[j][1] + uz*dxq[j][2];
          if (deltaMin > 0.0) psi = dmin(1., deltaPlusMax/deltaMin);
          if (deltaMin < 0.0) psi = dmin(1., deltaPlusMin/deltaMin);
          psimin = dmin(psi, psimin);
        }
      }
      double* out = &d.lim[(size_t)v*d.nCells + cell];
      *out = fresh ? psimin : dmin(psimin, *out);
    }
  }
};

// ------------------------------------------------------------------------------------------------
// Per-face working set shared by the flux kernel and the finite-difference Jacobian kernels.
template <int D>
struct FaceWork {
  static constexpr int Q = D + 2;
  typedef Phys<D> P;
  int f, L, R, nbN; bool isB;
  double n[3], area, xq[3], xL[3], xR[3];
  double gL[Q], gR[Q];      // limiter * gradientFactor * grad . (xq - x): the frozen reconstruction increments
  // diamond control volume (DiamondVolume{2,3}DDerivative)
  double cvn[8][3];         // 2D: 4 edge normals; 3D quad: 2x4 pyramid normals; 3D tri: 6 tetra normals
  double volume, mu, lambda;
  double val[Q][6];         // values(i,j) = state j of the control volume, variable i
  double avState[Q];
  double grad[Q][3];

  CFB_HD void loadGeom(const Dev& d, int face) {
    f = face; L = d.faceL[f]; R = d.faceR[f]; isB = (R >= d.nCells);
    area = d.areas[f];
    const double invArea = 1./area;                     // FVMCC_ComputeRHS.cxx:518-534
    for (int k = 0; k < D; ++k) {
      n[k] = d.normals[(size_t)k*d.nFaces + f]*invArea;
      xq[k] = d.faceCentres[(size_t)k*d.nFaces + f];
      xL[k] = d.centres[(size_t)k*d.nStates + L];
      xR[k] = d.centres[(size_t)k*d.nStates + R];
    }
    if (D == 2) { n[2] = 0.; xq[2] = 0.; xL[2] = 0.; xR[2] = 0.; }
  }

  // a8: LeastSquareP1PolyRec2D.cxx:158-217, 3D.cxx:204-268.  Boundary faces use the INNER cell's gradient
  // and limiter for the ghost side (reference behaviour).
  CFB_HD void loadReconstruction(const Dev& d, const Cfg& c) {
    if (c.polyRec == CFB200_PR_CONSTANT) { for (int v = 0; v < Q; ++v) { gL[v] = 0.; gR[v] = 0.; } return; }
    const int gradR = isB ? L : R;
    const int32_t* zg = isB ? d.bcs[d.faceBC[f]].zeroGrad : nullptr;
    for (int v = 0; v < Q; ++v) {
      const bool z = zg && zg[v];
      double gcS = 0.0, gcN = 0.0;
      if (!z) {
        const size_t iL = (size_t)v*d.nCells + L, iR = (size_t)v*d.nCells + gradR;
        if (D == 2) {
          gcS = c.gradientFactor*(d.uX[iL]*(xq[0] - xL[0]) + d.uY[iL]*(xq[1] - xL[1]));
          gcN = c.gradientFactor*(d.uX[iR]*(xq[0] - xR[0]) + d.uY[iR]*(xq[1] - xR[1]));
        } else {
          gcS = c.gradientFactor*(d.uX[iL]*(xq[0] - xL[0]) + d.uY[iL]*(xq[1] - xL[1]) + d.uZ[iL]*(xq[2] - xL[2]));
          gcN = c.gradientFactor*(d.uX[iR]*(xq[0] - xR[0]) + d.uY[iR]*(xq[1] - xR[1]) + d.uZ[iR]*(xq[2] - xR[2]));
        }
      }
      const double lValue = (c.stopLimiting == 0) ? d.lim[(size_t)v*d.nCells + L] : 1.;
      const double rValue = (c.stopLimiting == 0) ? d.lim[(size_t)v*d.nCells + gradR] : 1.;
      gL[v] = lValue*gcS;
      gR[v] = rValue*gcN;
    }
  }
  CFB_HD void reconstruct(const Cfg& c, const double* sL, const double* sR, double* vL, double* vR) const {
    if (c.polyRec == CFB200_PR_CONSTANT) { for (int v = 0; v < Q; ++v) { vL[v] = sL[v]; vR[v] = sR[v]; } return; }
    for (int v = 0; v < Q; ++v) { vL[v] = sL[v] + gL[v]; vR[v] = sR[v] + gR[v]; }
  }

  // a15: control volume of the face, DiamondVolume2DDerivative.cxx:91-137, DiamondVolume3DDerivative.cxx:336-411,
  // and the nodal part of the control-volume states.
  CFB_HD void loadDiamond(const Dev& d) {
    const int b = d.faceNodePtr[f];
    nbN = d.faceNodePtr[f+1] - b;
    int fn[4];
    double xn[4][3];
    for (int j = 0; j < nbN; ++j) {
      fn[j] = d.faceNodes[b + j];
      for (int k = 0; k < D; ++k) xn[j][k] = d.nodes[(size_t)k*d.nNodes + fn[j]];
    }
    if (D == 2) {
      const double* node0 = xn[0]; const double* node1 = xL; const double* node2 = xn[1]; const double* node3 = xR;
      cvn[0][0] = -node0[1] + node1[1]; cvn[0][1] = -node1[0] + node0[0];
      cvn[1][0] = -node1[1] + node2[1]; cvn[1][1] = -node2[0] + node1[0];
      cvn[2][0] = -node2[1] + node3[1]; cvn[2][1] = -node3[0] + node2[0];
      cvn[3][0] = -node3[1] + node0[1]; cvn[3][1] = -node0[0] + node3[0];
      const double det = (node2[0] - node0[0])*(node3[1] - node1[1]) - (node3[0] - node1[0])*(node2[1] - node0[1]);
      if (det < 0.0) for (int k = 0; k < 4; ++k) { cvn[k][0] *= -1.; cvn[k][1] *= -1.; }
      volume = 0.5*fabs(det);
      // control-volume order: node0, L, node1, R
      for (int v = 0; v < Q; ++v) {
        val[v][0] = d.nstates[(size_t)v*d.nNodes + fn[0]];
        val[v][2] = d.nstates[(size_t)v*d.nNodes + fn[1]];
        avState[v] = 0.5*(val[v][0] + val[v][2]);
      }
    } else if (nbN == 3) {
      // compute2TetraGradient normals, DiamondVolume3DDerivative.cxx:190-263
      double factor = (!d.cvSign[f]) ? -0.5 : 0.5;
      tetN(xL, xn[0], xn[1], xn[0], factor, cvn[0]);     // n031
      tetN(xL, xn[1], xn[2], xn[1], factor, cvn[1]);     // n132
      tetN(xn[2], xn[0], xL, xn[0], factor, cvn[2]);     // n023
      factor *= -1;
      tetN(xR, xn[0], xn[1], xn[0], factor, cvn[3]);     // n041
      tetN(xR, xn[1], xn[2], xn[1], factor, cvn[4]);     // n142
      tetN(xn[2], xn[0], xR, xn[0], factor, cvn[5]);     // n024
      volume = d.cvol[f];
      for (int v = 0; v < Q; ++v) {
        for (int j = 0; j < 3; ++j) val[v][j] = d.nstates[(size_t)v*d.nNodes + fn[j]];
        avState[v] = (1./3.)*(val[v][0] + val[v][1] + val[v][2]);
      }
    } else {
      // compute2PyramidNormals, DiamondVolume3DDerivative.cxx:336-383
      double factor = (!d.cvSign[f]) ? 1.0 : -1.0;
      triNormal(xn[0], xn[1], xL, factor, cvn[0]); triNormal(xn[1], xn[2], xL, factor, cvn[1]);
      triNormal(xn[2], xn[3], xL, factor, cvn[2]); triNormal(xn[0], xL, xn[3], factor, cvn[3]);
      factor *= -1.;
      triNormal(xn[0], xn[1], xR, factor, cvn[4]); triNormal(xn[1], xn[2], xR, factor, cvn[5]);
      triNormal(xn[2], xn[3], xR, factor, cvn[6]); triNormal(xn[0], xR, xn[3], factor, cvn[7]);
      volume = d.cvol[f];
      for (int v = 0; v < Q; ++v) {
        for (int j = 0; j < 4; ++j) val[v][j] = d.nstates[(size_t)v*d.nNodes + fn[j]];
        avState[v] = 0.25*(val[v][0] + val[v][1] + val[v][2] + val[v][3]);
      }
    }
  }
  static CFB_HD void tetN(const double* a, const double* b, const double* c2, const double* dd, double fct, double* out) {
    double v1[3], v2[3], v3[3];
    for (int k = 0; k < 3; ++k) { v1[k] = a[k] - b[k]; v2[k] = c2[k] - dd[k]; }
    cross3(v1, v2, v3);
    for (int k = 0; k < 3; ++k) out[k] = fct*v3[k];
  }
  // positions of the L / R cell states in the control volume
  CFB_HD int slotL() const { return (D == 2) ? 1 : nbN; }
  CFB_HD int slotR() const { return (D == 2) ? 3 : nbN + 1; }

  // Green-Gauss gradient of variable v over the diamond: DiamondVolume2DDerivative.cxx:139-176,
  // DiamondVolume3DDerivative.cxx:190-298
  CFB_HD void diamondGrad(int v) {
    const double* a = val[v];
    if (D == 2) {
      const double invVolume = 0.5/volume;
      const double v01 = a[0] + a[1], v12 = a[1] + a[2], v23 = a[2] + a[3], v30 = a[3] + a[0];
      for (int ix = 0; ix < 2; ++ix)
        grad[v][ix] = invVolume*(cvn[0][ix]*v01 + cvn[1][ix]*v12 + cvn[2][ix]*v23 + cvn[3][ix]*v30);
      grad[v][2] = 0.;
    } else if (nbN == 3) {
      const double invVolume = (1./3.)/volume;
      const double v031 = a[0] + a[3] + a[1], v132 = a[1] + a[3] + a[2], v023 = a[0] + a[2] + a[3];
      const double v041 = a[0] + a[4] + a[1], v142 = a[1] + a[4] + a[2], v024 = a[0] + a[2] + a[4];
      for (int ix = 0; ix < 3; ++ix)
        grad[v][ix] = invVolume*(cvn[0][ix]*v031 + cvn[1][ix]*v132 + cvn[2][ix]*v023 +
                                 cvn[3][ix]*v041 + cvn[4][ix]*v142 + cvn[5][ix]*v024);
    } else {
      const double invVolume = (1./3.)/volume;
      const double v014 = a[0] + a[1] + a[4], v124 = a[1] + a[2] + a[4], v234 = a[2] + a[3] + a[4], v043 = a[0] + a[4] + a[3];
      const double v015 = a[0] + a[1] + a[5], v125 = a[1] + a[2] + a[5], v235 = a[2] + a[3] + a[5], v053 = a[0] + a[5] + a[3];
      for (int ix = 0; ix < 3; ++ix)
        grad[v][ix] = invVolume*(cvn[0][ix]*v014 + cvn[1][ix]*v124 + cvn[2][ix]*v234 + cvn[3][ix]*v043 +
                                 cvn[4][ix]*v015 + cvn[5][ix]*v125 + cvn[6][ix]*v235 + cvn[7][ix]*v053);
    }
  }

  // a14: NSFlux.ci:59-156: flux integrated over the face.  mu/lambda depend on the NODAL average state only,
  // which the perturbation loop never touches, so they are evaluated once per face.
  CFB_HD void transport(const Cfg& c) {
    mu = P::dynViscosity(c, avState)*c.artDiffCoeff;
    lambda = P::thermConductivity(c, mu);
  }
  CFB_HD void diffFlux(const Cfg& c, double* dflux) const {
    P::nsFlux(c, avState, grad, n, mu, lambda, dflux);
    for (int v = 0; v < Q; ++v) dflux[v] *= area;
  }
};

// ------------------------------------------------------------------------------------------------
// a11/a14/a17: one thread per face: reconstruct, Roe flux, Navier-Stokes flux, update coefficients.
// Writes face-owned slots only (no scatter).
template <int D>
struct FaceFlux {   // idx over faces
  Dev d; Cfg c;
  CFB_HD void operator()(int f) const {
    constexpr int Q = D + 2;
    typedef Phys<D> P;
    FaceWork<D> w;
    w.loadGeom(d, f);
    const bool upL = d.updatable[w.L];
    const bool upR = (!w.isB) && d.updatable[w.R];
    if (!(upL || upR)) {                                   // FVMCC_ComputeRHS.cxx:212-213
      for (int v = 0; v < Q; ++v) d.faceFlux[(size_t)v*d.nFaces + f] = 0.;
      d.ucL[f] = 0.; d.ucR[f] = 0.; d.ucD[f] = 0.;
      return;
    }
    double sL[Q], sR[Q], vL[Q], vR[Q];
    for (int v = 0; v < Q; ++v) { sL[v] = d.U[(size_t)v*d.nStates + w.L]; sR[v] = d.U[(size_t)v*d.nStates + w.R]; }
    w.loadReconstruction(d, c);
    w.reconstruct(c, sL, sR, vL, vR);
    typename P::Side a, b;
    P::side(c, vL, w.n, a);
    P::side(c, vR, w.n, b);
    double flux[Q];
    if (!P::roe(c, a, b, w.n, flux)) { *d.status = CFB200_ERR_BAD_VALUE; for (int v = 0; v < Q; ++v) flux[v] = 0.; }
    // RoeFlux.cxx:165-185 update coefficients (one quadrature point per face)
    const double faceArea = w.area/1;
    d.ucL[f] = dmax(P::maxEigenValue(a.pd, w.n), 0.)*faceArea;
    double tn[3];
    for (int k = 0; k < 3; ++k) tn[k] = -1.0*w.n[k];
    d.ucR[f] = dmax(P::maxEigenValue(b.pd, tn), 0.)*faceArea;
    for (int v = 0; v < Q; ++v) flux[v] *= w.area;         // FVMCC_FluxSplitter.hh:79-83
    double ucD = 0.;
    if (c.diffusive == CFB200_DIFF_NS) {
      w.loadDiamond(d);
      for (int v = 0; v < Q; ++v) { w.val[v][w.slotL()] = sL[v]; w.val[v][w.slotR()] = sR[v]; }
      for (int v = 0; v < Q; ++v) w.diamondGrad(v);
      w.transport(c);
      double dflux[Q];
      w.diffFlux(c, dflux);
      for (int v = 0; v < Q; ++v) flux[v] -= dflux[v];     // FVMCC_ComputeRHS.cxx:260-276
      const double nu = w.mu/P::density(c, w.avState);     // NSFlux.ci:127-152
      ucD = nu*w.area*w.area/w.volume;
      d.faceMu[f] = w.mu;
    }
    d.ucD[f] = ucD;
    for (int v = 0; v < Q; ++v) d.faceFlux[(size_t)v*d.nFaces + f] = flux[v];
  }
};

// a17: FVMCC_ComputeRHS::updateRHS (FVMCC_ComputeRHS.cxx:455-514) as a gather: for each cell, its faces in
// ascending loop order; rhs starts from 0, updateCoeff accumulates onto the incoming value
// (it is zeroed by the updater, UpdateSol.cxx:192).
template <int D>
struct CellGather {   // idx over cells
  Dev d; Cfg c;
  CFB_HD void operator()(int cell) const {
    constexpr int Q = D + 2;
    double r[Q];
    for (int v = 0; v < Q; ++v) r[v] = 0.;
    double uc = d.updateCoeff[cell];
    const double coeff = c.resFactor*1.0;
    const bool hasDiff = (c.diffusive == CFB200_DIFF_NS);
    for (int j = d.cellFacePtr[cell]; j < d.cellFacePtr[cell+1]; ++j) {
      const int e = d.cellFaceSorted[j];
      const bool isR = e < 0;
      const int f = isR ? ~e : e;
      const int oL = d.faceL[f];
      const bool skip = !(d.updatable[oL] || (f < d.nInner && d.updatable[d.faceR[f]]));
      if (skip) continue;
      for (int v = 0; v < Q; ++v) {
        const double rFlux = coeff*d.faceFlux[(size_t)v*d.nFaces + f];
        if (isR) r[v] += rFlux*1.0; else r[v] -= rFlux*1.0;
      }
      uc += isR ? d.ucR[f] : d.ucL[f];
      if (hasDiff) uc += d.ucD[f];
    }
    for (int v = 0; v < Q; ++v) d.rhs[(size_t)cell*Q + v] = r[v];
    d.updateCoeff[cell] = uc;
  }
};

// ------------------------------------------------------------------------------------------------
// a18: finite-difference block Jacobian of an interior face, FVMCC_ComputeRhsJacob.cxx:126-236,321-365.
// One thread per (face, side): perturbs the q variables of that side's cell in turn, re-evaluates the
// convective + diffusive flux with frozen gradient/limiter, and writes the q x q block
//   side L: block(R,L) = -resFactor * dF/dU_L      side R: block(L,R) = +resFactor * dF/dU_R
// (the diagonal contributions are the negatives of these and are gathered by DiagGather).
template <int D>
struct JacInner {   // idx over 2*nInner
  Dev d; Cfg c;
  CFB_HD void operator()(int idx) const {
    constexpr int Q = D + 2;
    typedef Phys<D> P;
    const int f = idx >> 1, side = idx & 1;
    FaceWork<D> w;
    w.loadGeom(d, f);
    double* blk = &d.bsrVal[(size_t)(side == 0 ? d.slotRL[f] : d.slotLR[f])*Q*Q];
    const bool upL = d.updatable[w.L], upR = d.updatable[w.R];
    if (!(upL || upR)) { for (int k = 0; k < Q*Q; ++k) blk[k] = 0.; return; }
    double sL[Q], sR[Q], vL[Q], vR[Q], flux[Q];
    for (int v = 0; v < Q; ++v) {
      sL[v] = d.U[(size_t)v*d.nStates + w.L]; sR[v] = d.U[(size_t)v*d.nStates + w.R];
      flux[v] = d.faceFlux[(size_t)v*d.nFaces + f];
    }
    w.loadReconstruction(d, c);
    w.reconstruct(c, sL, sR, vL, vR);
    const bool hasDiff = (c.diffusive == CFB200_DIFF_NS);
    if (hasDiff) {
      w.loadDiamond(d);
      for (int v = 0; v < Q; ++v) { w.val[v][w.slotL()] = sL[v]; w.val[v][w.slotR()] = sR[v]; }
      for (int v = 0; v < Q; ++v) w.diamondGrad(v);
      w.mu = d.faceMu[f];
      w.lambda = P::thermConductivity(c, w.mu);
    }
    typename P::Side fixed, pert;
    P::side(c, side == 0 ? vR : vL, w.n, fixed);
    double* vP = side == 0 ? vL : vR;
    const double* sP = side == 0 ? sL : sR;
    const double* gP = side == 0 ? w.gL : w.gR;
    const int slot = side == 0 ? w.slotL() : w.slotR();
    // both updatable: rows L and R get +rf and -rf; one side only: row idx gets (-1)^idx * rf
    // (FVMCC_ComputeRhsJacob.cxx:184-236).  Stored block = what row "other" receives.
    const double rf = c.resFactor;
    for (int iVar = 0; iVar < Q; ++iVar) {
      const double orig = sP[iVar];
      const double eps = fdEps(c, iVar, orig);
      const double pv = orig + eps;
      const double keep = vP[iVar];
      vP[iVar] = (c.polyRec == CFB200_PR_CONSTANT) ? pv : pv + gP[iVar];   // LeastSquareP1PolyRec2D.cxx:221-262
      P::side(c, vP, w.n, pert);
      double pf[Q];
      bool ok = (side == 0) ? P::roe(c, pert, fixed, w.n, pf) : P::roe(c, fixed, pert, w.n, pf);
      if (!ok) { *d.status = CFB200_ERR_BAD_VALUE; for (int v = 0; v < Q; ++v) pf[v] = 0.; }
      for (int v = 0; v < Q; ++v) pf[v] *= w.area;
      if (hasDiff) {
        const double keepVal = w.val[iVar][slot];
        double keepGrad[3];
        for (int k = 0; k < 3; ++k) keepGrad[k] = w.grad[iVar][k];
        w.val[iVar][slot] = pv;
        w.diamondGrad(iVar);
        double dflux[Q];
        w.diffFlux(c, dflux);
        for (int v = 0; v < Q; ++v) pf[v] -= dflux[v];
        w.val[iVar][slot] = keepVal;
        for (int k = 0; k < 3; ++k) w.grad[iVar][k] = keepGrad[k];
      }
      const double invEps = 1./eps;
      for (int v = 0; v < Q; ++v) {
        double fd = (pf[v] - flux[v])*invEps;
        fd *= rf;                              // what row L receives
        blk[v*Q + iVar] = (side == 0) ? fd*(-1.0) : fd;
      }
      vP[iVar] = keep;
    }
  }
};

// Boundary faces: FVMCC_ComputeRhsJacob.cxx:240-317.  The inner state is perturbed, the ghost state is rebuilt
// by the boundary condition, both sides are re-extrapolated, and resFactor * dF/dU_L goes to the diagonal block.
template <int D>
struct JacBoundary {   // idx over boundary faces
  Dev d; Cfg c;
  CFB_HD void operator()(int g) const {
    constexpr int Q = D + 2;
    typedef Phys<D> P;
    const int f = d.nInner + g;
    FaceWork<D> w;
    w.loadGeom(d, f);
    double* blk = &d.bndBlocks[(size_t)g*Q*Q];
    if (!d.updatable[w.L]) { for (int k = 0; k < Q*Q; ++k) blk[k] = 0.; return; }
    double sL[Q], sR[Q], vL[Q], vR[Q], flux[Q], nr[D];
    for (int v = 0; v < Q; ++v) { sL[v] = d.U[(size_t)v*d.nStates + w.L]; flux[v] = d.faceFlux[(size_t)v*d.nFaces + f]; }
    for (int k = 0; k < D; ++k) nr[k] = d.normals[(size_t)k*d.nFaces + f];
    const BCond& bc = d.bcs[d.faceBC[f]];
    const bool hasDiff = (c.diffusive == CFB200_DIFF_NS);
    if (hasDiff) {
      w.loadDiamond(d);
      w.mu = d.faceMu[f];
      w.lambda = P::thermConductivity(c, w.mu);
    }
    // the reconstruction increments use the inner cell's gradient for both sides and do not depend on the states
    w.loadReconstruction(d, c);
    const double rf = c.resFactor;
    for (int iVar = 0; iVar < Q; ++iVar) {
      const double orig = sL[iVar];
      const double eps = fdEps(c, iVar, orig);
      sL[iVar] = orig + eps;
      ghostState<D>(c, bc, nr, sL, sR);
      w.reconstruct(c, sL, sR, vL, vR);
      typename P::Side a, b;
      P::side(c, vL, w.n, a);
      P::side(c, vR, w.n, b);
      double pf[Q];
      if (!P::roe(c, a, b, w.n, pf)) { *d.status = CFB200_ERR_BAD_VALUE; for (int v = 0; v < Q; ++v) pf[v] = 0.; }
      for (int v = 0; v < Q; ++v) pf[v] *= w.area;
      if (hasDiff) {
        for (int v = 0; v < Q; ++v) { w.val[v][w.slotL()] = sL[v]; w.val[v][w.slotR()] = sR[v]; }
        for (int v = 0; v < Q; ++v) w.diamondGrad(v);
        double dflux[Q];
        w.diffFlux(c, dflux);
        for (int v = 0; v < Q; ++v) pf[v] -= dflux[v];
      }
      const double invEps = 1./eps;
      for (int v = 0; v < Q; ++v) {
        double fd = (pf[v] - flux[v])*invEps;
        fd *= rf;
        blk[v*Q + iVar] = fd;
      }
      sL[iVar] = orig;
    }
  }
};

// Diagonal blocks: one thread per (cell, block entry).  diag(c) = sum over the cell's faces in loop order of
//   inner face, c == L:  -block(R,c)      inner face, c == R:  -block(L,c)      boundary face: +bnd block
// Rows of non-updatable cells are staging only and are zeroed by ZeroRows afterwards.
template <int D>
struct DiagGather {   // idx over nCells*32 (one warp-sized group per cell, Q*Q <= 25 lanes active)
  Dev d;
  CFB_HD void operator()(int idx) const {
    constexpr int Q = D + 2;
    const int cell = idx >> 5, e = idx & 31;
    if (e >= Q*Q) return;
    double acc = 0.;
    if (d.updatable[cell]) {
      for (int j = d.cellFacePtr[cell]; j < d.cellFacePtr[cell+1]; ++j) {
        const int s = d.cellFaceSorted[j];
        const bool isR = s < 0;
        const int f = isR ? ~s : s;
        if (f < d.nInner) {
          const int slot = isR ? d.slotLR[f] : d.slotRL[f];
          acc -= d.bsrVal[(size_t)slot*Q*Q + e];
        } else {
          acc += d.bndBlocks[(size_t)(f - d.nInner)*Q*Q + e];
        }
      }
    }
    d.bsrVal[(size_t)d.diagSlot[cell]*Q*Q + e] = acc;
  }
};

template <int D>
struct ZeroRows {   // idx over (non-updatable cell list) * 32
  Dev d; const int* cells;
  CFB_HD void operator()(int idx) const {
    constexpr int Q = D + 2;
    const int cell = cells[idx >> 5], e = idx & 31;
    if (e >= Q*Q) return;
    for (int k = d.bsrRowPtr[cell]; k < d.bsrRowPtr[cell+1]; ++k) d.bsrVal[(size_t)k*Q*Q + e] = 0.;
  }
};

// ------------------------------------------------------------------------------------------------
// a21: FVMCC_PseudoSteadyTimeRhs.cxx:106-262, steady branch (dt == 0): diag = updateCoeff/CFL,
// rhs -= (U - U_past) * diag * resFactor in SOLUTION variables, diagonal block += diag * dU/dP (numerical)
template <int D>
struct TimeRhs {   // idx over cells
  Dev d; Cfg c; int doJac;
  CFB_HD void operator()(int cell) const {
    constexpr int Q = D + 2;
    typedef Phys<D> P;
    if (!d.updatable[cell]) return;
    const double diagValue = d.updateCoeff[cell]/c.cfl;
    double u[Q], up[Q], tempState[Q], tPast[Q], tPert[Q];
    for (int v = 0; v < Q; ++v) { u[v] = d.U[(size_t)v*d.nStates + cell]; up[v] = d.pastU[(size_t)v*d.nCells + cell]; }
    P::toCons(c, u, tempState);
    P::toCons(c, up, tPast);
    for (int v = 0; v < Q; ++v) {
      const double dU = tempState[v] - tPast[v];
      d.rhs[(size_t)cell*Q + v] -= dU*diagValue*c.resFactor;
    }
    if (!doJac) return;
    double* blk = &d.bsrVal[(size_t)d.diagSlot[cell]*Q*Q];
    for (int iVar = 0; iVar < Q; ++iVar) {
      const double orig = u[iVar];
      const double eps = fdEps(c, iVar, orig);
      u[iVar] = orig + eps;
      P::toCons(c, u, tPert);
      const double invEps = 1./eps;
      for (int v = 0; v < Q; ++v) {
        double fd = (tPert[v] - tempState[v])*invEps;
        fd *= diagValue*c.resFactor;
        blk[v*Q + iVar] += fd;
      }
      u[iVar] = orig;
    }
  }
};

// ------------------------------------------------------------------------------------------------
// §8f-1: ForwardEuler UpdateSol.cxx:83-193 on the device: dt = CFL/updateCoeff, rhs *= dt (in place), U += rhs,
// updateCoeff = 0; per-cell squared dU[monitorVar] goes to `partial` for the ordered norm reduction.
template <int D>
struct UpdateSolFwdEuler {   // idx over cells
  Dev d; double cfl; int monitorVar;
  CFB_HD void operator()(int cell) const {
    constexpr int Q = D + 2;
    double sq = 0.;
    if (!d.updatable[cell]) {
      for (int v = 0; v < Q; ++v) d.rhs[(size_t)cell*Q + v] = 0.;
    } else {
      const double uc = d.updateCoeff[cell];
      const bool isZero = fabs(uc) <= 2.2250738585072014e-308;
      const double dt = isZero ? 0. : cfl/uc;
      for (int v = 0; v < Q; ++v) {
        double r = d.rhs[(size_t)cell*Q + v];
        r *= dt;
        d.rhs[(size_t)cell*Q + v] = r;
        d.U[(size_t)v*d.nStates + cell] += r;
      }
      const double t = d.rhs[(size_t)cell*Q + monitorVar];
      sq = t*t;
    }
    d.updateCoeff[cell] = 0.0;
    d.partial[cell] = sq;
  }
};

} // namespace cfb
