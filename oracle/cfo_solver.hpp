// ORACLE — TEST INFRASTRUCTURE ONLY (see cfo_types.hpp header).
// The FVMCC ComputeRHS / ComputeRhsJacob command restated with the reference's
// own loop structure: edge-wise LSQ accumulation, nodal extrapolation, a single
// loop over faces in TRS order with first-touch limiter evaluation, scatter of
// the flux into rhs, and per-face finite-difference Jacobian blocks added into
// a BSR matrix in face order.
#pragma once
#include "cfo_physics.hpp"

namespace cfo {

struct Outputs {
  double* rhs;          // [nCells*q]   (zeroed here, FVMCC_ComputeRHS.cxx:692)
  double* updateCoeff;  // [nCells]     (accumulated, NOT zeroed: UpdateSol.cxx:192 does that)
  double* uX;           // [nCells*q]
  double* uY;
  double* uZ;
  double* limiter;      // [nCells*q]   in/out (historical limiter)
  double* nstates;      // [nNodes*q]
  // BSR Jacobian (may be null): pattern from FVMCCSparsity.cxx:47-106 (diag + face neighbours)
  const int32_t* bsrRowPtr;
  const int32_t* bsrColInd;
  double* bsrVal;       // [nnzb*q*q], block row-major [ib][jb]  (BlockAccumulator.hh:289-296)
  int32_t status;       // 0 ok, 1 = a2<0 in the Roe linearisation (BadValueException)
  double* divB;         // [nCells] or null: FVMCC_ComputeRHSMHD.ci:160-203 diagnostic (zeroed here, MHD only)
};

template <int D, class PHYS = Phys<D> >
class Solver {
public:
  typedef PHYS P;                       // Phys<D>: Euler / Navier-Stokes (Q = D + 2);  MhdPhys: MHD3DProjection (Q = 9)
  static constexpr int Q = P::Q;
  typedef typename P::Data PData;

  Solver(const Config& c, const MeshView& m, const BC* bcs) : cfg(c), mesh(m), bc(bcs) {}

  // ---- boundary conditions (a16) -------------------------------------------
  // FVMCC_BaseBC.ci:144-187 executeOnTrs -> setGhostState(face) for every boundary face.
  void setGhostState(int f, const double* inner, double* ghost) const {
    const BC& b = bc[mesh.faceBC[f]];
    switch (b.type) {
      case BC_SUPER_OUTLET:
        for (int i = 0; i < Q; ++i) ghost[i] = inner[i];
        break;
      case BC_DIRICHLET: {
        const int g = f - mesh.nInner;
        const double* value = (mesh.bcFaceFlag && mesh.bcFaceFlag[g]) ? &mesh.bcFaceVal[(size_t)g*Q] : b.value;
        for (int i = 0; i < Q; ++i) ghost[i] = b.mask[i] ? (2.*value[i] - inner[i]) : inner[i];
        break;
      }
      case BC_MIRROR_VELOCITY: {       // MirrorVelocity.cxx:71-111 (velocity IDs = 1..D)
        const double* nr = &mesh.normals[(size_t)f*D];
        double vn = 0., area2 = 0.;
        for (int i = 0; i < D; ++i) { const double nxComp = nr[i]; vn += nxComp*inner[1+i]; area2 += nxComp*nxComp; }
        ghost[0] = inner[0];
        for (int i = 0; i < D; ++i) ghost[1+i] = inner[1+i] - 2.0*vn*nr[i]/area2;
        ghost[Q-1] = inner[Q-1];
        break;
      }
      case BC_MIRROR_EULER_CONS:       // MirrorEuler2D.cxx:56-94 / MirrorMHD3DProjection.cxx:56-96
        P::mirrorGhost(cfg, &mesh.normals[(size_t)f*D], inner, ghost);
        break;
      case BC_FARFIELD_EULER:          // FarFieldEuler2D.cxx:127-300, FarFieldEuler3D.cxx:62-196
        P::farFieldGhost(cfg, &mesh.normals[(size_t)f*D], b.value, inner, ghost);
        break;
      case BC_NOSLIP_ISOTHERMAL: {     // NoSlipWallIsothermalNSvt.ci:178-243 (not adiabatic, FixTWallInGhost = false)
        // computeGhostPosition (FVMCC_BaseBC.ci:271-350): mirror of the inner node through the face plane, the wall
        // point half way, the two distances
        const double* n = &mesh.normals[(size_t)f*D];
        const int L = mesh.faceL[f];
        const double* xI = &mesh.centres[(size_t)L*D];
        double mid[3] = {0., 0., 0.};
        const int nb = mesh.faceNodePtr[f+1] - mesh.faceNodePtr[f];
        for (int j = mesh.faceNodePtr[f]; j < mesh.faceNodePtr[f+1]; ++j)
          for (int k = 0; k < D; ++k) mid[k] += mesh.nodes[(size_t)mesh.faceNodes[j]*D + k];
        for (int k = 0; k < D; ++k) mid[k] /= (double)nb;
        double nm = 0., n2 = 0., nx = 0.;
        for (int k = 0; k < D; ++k) { nm += n[k]*mid[k]; n2 += n[k]*n[k]; nx += n[k]*xI[k]; }
        const double kk = -nm;
        const double t = (nx + kk)/n2;
        double xG[3], xM[3];
        for (int k = 0; k < D; ++k) { xG[k] = xI[k] - 2.*t*n[k]; xM[k] = 0.5*(xI[k] + xG[k]); }
        const double drXiXw = getDistance<D>(xI, xM);
        const double drXiXg = getDistance<D>(xI, xG);
        // repositionNode (FVMCC_BaseBC.hh:290-340): ghostValue = 2 bValue - innerValue; the loop that moves the ghost
        // node when it is below minValue is NOT restated: status 8
        const int TID = Q - 1;
        const double ghostT = 2.*b.value[TID] - inner[TID];
        if (!(ghostT >= b.value[0])) throw std::runtime_error("ghost reposition");
        for (int i = 0; i < Q; ++i) {
          if (i >= 1 && i <= D) ghost[i] = inner[i] - (inner[i] - b.value[i])*(drXiXg/drXiXw);     // linearInterpolate
          else if (i != TID) ghost[i] = inner[i];
          else ghost[i] = ghostT;
        }
        break;
      }
      case BC_FARFIELD_EULER_CHAR:     // FarFieldEulerChar2D.cxx:36-254, FarFieldEulerChar3D.cxx:40-200
        P::farFieldCharGhost(cfg, &mesh.normals[(size_t)f*D], b.value, inner, ghost);
        break;
      case BC_SUPER_OUTLET_MHD:        // SuperOutletMHD3DProjection.cxx:76-239 (inner / ghost state coordinates)
        P::superOutletGhost(cfg, b.mask[0] != 0, b.value[0], &mesh.centres[(size_t)mesh.faceL[f]*D],
                            &mesh.centres[(size_t)mesh.faceR[f]*D], inner, ghost);
        break;
      default: throw std::runtime_error("unknown BC type");
    }
  }

  void applyBC(double* U) const {
    for (int f = mesh.nInner; f < mesh.nFaces; ++f) {
      setGhostState(f, &U[(size_t)mesh.faceL[f]*Q], &U[(size_t)mesh.faceR[f]*Q]);
    }
  }

  // ---- LSQ weights / normal matrix (a4): LeastSquareP1PolyRec{2,3}D::updateWeights
  // (LeastSquareP1PolyRec2D.cxx:266-330, 3D.cxx:333-411)
  void updateWeights() {
    const int nc = mesh.nCells;
    for (int k = 0; k < 6; ++k) l[k].assign(nc, 0.0);
    weights.clear();
    for (int iState = 0; iState < nc; ++iState) {
      const double* xf = &mesh.centres[(size_t)iState*D];
      for (int in = mesh.stencilPtr[iState]; in < mesh.stencilPtr[iState+1]; ++in) {
        const int last = mesh.stencil[in];
        const bool ghost = last >= nc;
        if (ghost || last > iState) {
          const double* xl = &mesh.centres[(size_t)last*D];
          const double deltaR = getDistance<D>(xf, xl);
          const double w = 1.0/deltaR;
          weights.push_back(w);
          double d[3] = {0,0,0};
          for (int i = 0; i < D; ++i) d[i] = w*(xl[i] - xf[i]);
          if (D == 2) {
            l[0][iState] += d[0]*d[0]; l[1][iState] += d[0]*d[1]; l[3][iState] += d[1]*d[1];
            if (!ghost) { l[0][last] += d[0]*d[0]; l[1][last] += d[0]*d[1]; l[3][last] += d[1]*d[1]; }
          } else {
            // l11 l12 l13 l22 l23 l33 -> l[0] l[1] l[2] l[3] l[4] l[5]
            l[0][iState] += d[0]*d[0]; l[1][iState] += d[0]*d[1]; l[2][iState] += d[0]*d[2];
            l[3][iState] += d[1]*d[1]; l[4][iState] += d[1]*d[2]; l[5][iState] += d[2]*d[2];
            if (!ghost) {
              l[0][last] += d[0]*d[0]; l[1][last] += d[0]*d[1]; l[2][last] += d[0]*d[2];
              l[3][last] += d[1]*d[1]; l[4][last] += d[1]*d[2]; l[5][last] += d[2]*d[2];
            }
          }
        }
      }
    }
  }

  // ---- LSQ gradients (a5): LeastSquareP1PolyRec2D.cxx:80-154, 3D.cxx:80-200
  void computeGradients(const double* U, Outputs& o) {
    const int nc = mesh.nCells;
    std::vector<double> lf1(nc), lf2(nc), lf3(nc);
    for (int iVar = 0; iVar < Q; ++iVar) {
      std::fill(lf1.begin(), lf1.end(), 0.0);
      std::fill(lf2.begin(), lf2.end(), 0.0);
      std::fill(lf3.begin(), lf3.end(), 0.0);
      size_t iEdge = 0;
      for (int iState = 0; iState < nc; ++iState) {
        const double* xf = &mesh.centres[(size_t)iState*D];
        for (int in = mesh.stencilPtr[iState]; in < mesh.stencilPtr[iState+1]; ++in) {
          const int last = mesh.stencil[in];
          const bool ghost = last >= nc;
          if (ghost || last > iState) {
            const double* xl = &mesh.centres[(size_t)last*D];
            const double w = weights[iEdge];
            const double dx = w*(xl[0] - xf[0]);
            const double dy = w*(xl[1] - xf[1]);
            const double dz = (D == 3) ? w*(xl[2] - xf[2]) : 0.;
            const double du = w*(U[(size_t)last*Q + iVar] - U[(size_t)iState*Q + iVar]);
            lf1[iState] += dx*du; lf2[iState] += dy*du;
            if (D == 3) lf3[iState] += dz*du;
            if (!ghost) {
              lf1[last] += dx*du; lf2[last] += dy*du;
              if (D == 3) lf3[last] += dz*du;
            }
            ++iEdge;
          }
        }
      }
      for (int i = 0; i < nc; ++i) {
        if (D == 2) {
          const double l11 = l[0][i], l12 = l[1][i], l22 = l[3][i];
          const double invDet = 1./(l11*l22 - l12*l12);
          o.uX[(size_t)i*Q + iVar] = (l22*lf1[i] - l12*lf2[i])*invDet;
          o.uY[(size_t)i*Q + iVar] = (l11*lf2[i] - l12*lf1[i])*invDet;
        } else {
          const double l11 = l[0][i], l12 = l[1][i], l13 = l[2][i], l22 = l[3][i], l23 = l[4][i], l33 = l[5][i];
          const double det = l11*l22*l33 - l11*l23*l23 - l12*l12*l33 + l12*l13*l23 + l13*l12*l23 - l13*l13*l22;
          const double linv11 = l22*l33 - l23*l23;
          const double linv22 = l11*l33 - l13*l13;
          const double linv33 = l11*l22 - l12*l12;
          const double linv12 = -(l12*l33 - l13*l23);
          const double linv13 = l12*l23 - l13*l22;
          const double linv23 = -(l11*l23 - l13*l12);
          if (!(std::abs(det) <= std::numeric_limits<double>::min())) {   // !MathChecks::isZero(det)
            o.uX[(size_t)i*Q + iVar] = (linv11*lf1[i] + linv12*lf2[i] + linv13*lf3[i])/det;
            o.uY[(size_t)i*Q + iVar] = (linv12*lf1[i] + linv22*lf2[i] + linv23*lf3[i])/det;
            o.uZ[(size_t)i*Q + iVar] = (linv13*lf1[i] + linv23*lf2[i] + linv33*lf3[i])/det;
          } else {
            o.uX[(size_t)i*Q + iVar] = 0.0; o.uY[(size_t)i*Q + iVar] = 0.0; o.uZ[(size_t)i*Q + iVar] = 0.0;
          }
        }
      }
    }
  }

  // ---- nodal extrapolation (a6): DistanceBasedExtrapolator.ci:35-149
  void extrapolateInAllNodes(const double* U, Outputs& o) const {
    for (int n = 0; n < mesh.nNodes; ++n) {
      double* ns = &o.nstates[(size_t)n*Q];
      for (int i = 0; i < Q; ++i) ns[i] = 0.;
      for (int k = mesh.nodeNbrPtr[n]; k < mesh.nodeNbrPtr[n+1]; ++k) {
        const double w = mesh.nodeW[k];
        const double* s = &U[(size_t)mesh.nodeNbr[k]*Q];
        for (int i = 0; i < Q; ++i) ns[i] += s[i]*w;
      }
      // DistanceBasedExtrapolatorGMove.cxx:162-173 extrapolate(): applyInner(), then applyBC() on the prescribed nodes;
      // :140-160 applyBC() with constant Values (idx < _radTID = 10000000 always; no file-interpolated nodal values)
      if (mesh.nodePresc && mesh.nodePresc[n])
        for (int i = 0; i < mesh.nNodePresc; ++i) ns[mesh.nodePrescIdx[i]] = mesh.nodePrescVal[i];
    }
  }

  // FVMCC_PolyRec.hh:114-123
  void faceMidPoint(int f, double* xq) const {
    const int b = mesh.faceNodePtr[f], e = mesh.faceNodePtr[f+1];
    const int nb = e - b;
    const double ov = 1./nb;
    for (int d = 0; d < D; ++d) xq[d] = 0.0;
    for (int k = b; k < e; ++k) {
      const double* xn = &mesh.nodes[(size_t)mesh.faceNodes[k]*D];
      for (int d = 0; d < D; ++d) xq[d] += xn[d]*ov;
    }
  }

  // ---- limiter (a7): Venktn2D.cxx:111-260, Venktn3D.cxx:44-190 ("OLD IMPLEMENTATION" branch,
  // isMFMHD = false), BarthJesp.cxx:60-155
  void limitCell(int c, const double* U, const Outputs& o, double* lim) const {
    const int fb = mesh.cellFacePtr[c], fe = mesh.cellFacePtr[c+1];
    const int nbFaces = fe - fb;
    const double* xc = &mesh.centres[(size_t)c*D];
    double xq[6][3];
    for (int k = 0; k < nbFaces; ++k) faceMidPoint(mesh.cellFaces[fb+k], xq[k]);
    double weight = 0.0;        // Venktn3DStrict.cxx:87: set at iVar == 0, kept for the other variables
    for (int iVar = 0; iVar < Q; ++iVar) {
      const double u0 = U[(size_t)c*Q + iVar];
      double min0 = u0, max0 = u0, avgDistance = 0.0;
      if (cfg.useFullStencil) {
        const int sb = mesh.stencilPtr[c], se = mesh.stencilPtr[c+1];
        const int nbNeighbors = se - sb;
        for (int is = sb; is < se; ++is) {
          const int s = mesh.stencil[is];
          if (s != c) {
            min0 = std::min(min0, U[(size_t)s*Q + iVar]);
            max0 = std::max(max0, U[(size_t)s*Q + iVar]);
            avgDistance += getDistance<D>(xc, &mesh.centres[(size_t)s*D]);
          }
        }
        avgDistance /= nbNeighbors;
      } else {
        for (int k = 0; k < nbFaces; ++k) {
          const int f = mesh.cellFaces[fb+k];
          const int nb = (mesh.faceL[f] == c) ? mesh.faceR[f] : mesh.faceL[f];
          min0 = std::min(min0, U[(size_t)nb*Q + iVar]);
          max0 = std::max(max0, U[(size_t)nb*Q + iVar]);
          avgDistance += getDistance<D>(xc, &mesh.centres[(size_t)nb*D]);
        }
        avgDistance /= nbFaces;
      }
      const double deltaPlusMax = max0 - u0;
      const double deltaPlusMin = min0 - u0;
      double psi = 1.0;
      double psimin = 1.1;
      const double ux = o.uX[(size_t)c*Q + iVar], uy = o.uY[(size_t)c*Q + iVar];
      const double uz = (D == 3) ? o.uZ[(size_t)c*Q + iVar] : 0.;
      if (cfg.limiter == LIM_VENKTN_STRICT && iVar == 0)      // Venktn3DStrict.cxx:150-151 (pow(x, 1.0) == x)
        weight = cfg.strictCoeff*std::abs((min0 + 1e-10)/(max0 + 1e-10)) + (1.0 - cfg.strictCoeff);
      if ((cfg.limiter == LIM_VENKTN && cfg.venktnIsMFMHD) || cfg.limiter == LIM_VENKTN_STRICT) {
        // "implementation from the Venkatakrishnan paper": Venktn3D.cxx:125-156, Venktn3DStrict.cxx:166-203 (the strict
        // limiter has only this branch: without isMFMHD it returns 1.1 and complains)
        if (cfg.limiter == LIM_VENKTN_STRICT && !cfg.venktnIsMFMHD) { lim[iVar] = psimin; continue; }
        const double mag = cfg.magnitudeValues[iVar];
        const double epsilon2 = cr_pow3(cfg.coeffEps*avgDistance/cfg.length);
        for (int k = 0; k < nbFaces; ++k) {
          double deltaMin;
          if (D == 2) deltaMin = ux*(xq[k][0] - xc[0]) + uy*(xq[k][1] - xc[1]);
          else deltaMin = ux*(xq[k][0] - xc[0]) + uy*(xq[k][1] - xc[1]) + uz*(xq[k][2] - xc[2]);
          for (int sgn = 0; sgn < 2; ++sgn) {
            if (sgn == 0 ? !(deltaMin > 0.0) : !(deltaMin < 0.0)) continue;
            const double dMinStar = ((deltaMin < 0.0) ? -1.0 : 1.0)*(std::abs(deltaMin)/mag + 1e-24);
            const double dPlus = (sgn == 0 ? deltaPlusMax : deltaPlusMin)/mag;
            const double dPlus2 = dPlus*dPlus;
            const double dPlusMin = dPlus*dMinStar;
            const double Num = (dPlus2 + epsilon2)*dMinStar + 2*dMinStar*dMinStar*dPlus;
            const double Den = (dPlus2 + 2*dMinStar*dMinStar + dPlusMin + epsilon2);
            psi = 1./dMinStar*(Num/Den);
          }
          if (cfg.limiter == LIM_VENKTN_STRICT) {
            if (iVar < 4 || iVar == 7) psimin = std::min(psi*weight, psimin);
            else psimin = (!cfg.psiMinEqual1) ? std::min(psi, psimin) : 1.0;
          } else psimin = std::min(psi, psimin);
        }
      } else if (cfg.limiter == LIM_VENKTN) {
        const double epsilon = cfg.coeffEps*cfg.magnitudeValues[iVar]*cfg.magnitudeValues[iVar]*
          cr_pow3(avgDistance/cfg.length);
        for (int k = 0; k < nbFaces; ++k) {
          double deltaMin;
          if (D == 2) deltaMin = ux*(xq[k][0] - xc[0]) + uy*(xq[k][1] - xc[1]);
          else deltaMin = ux*(xq[k][0] - xc[0]) + uy*(xq[k][1] - xc[1]) + uz*(xq[k][2] - xc[2]);
          if (deltaMin > 0.0) {
            const double dPlusMax2 = deltaPlusMax*deltaPlusMax;
            const double dPlusMaxMin = deltaPlusMax*deltaMin;
            psi = (dPlusMax2 + 2.0*dPlusMaxMin + epsilon)/
              (dPlusMax2 + dPlusMaxMin + 2.0*deltaMin*deltaMin + epsilon);
          }
          if (deltaMin < 0.0) {
            const double y = deltaPlusMin/deltaMin;
            psi = (y*y + 2*y)/(y*y + y + 2.);
          }
          psimin = std::min(psi, psimin);
        }
      } else { // LIM_BARTHJESP  BarthJesp.cxx:121-148
        for (int k = 0; k < nbFaces; ++k) {
          double deltaMin;
          if (D == 2) deltaMin = ux*(xq[k][0] - xc[0]) + uy*(xq[k][1] - xc[1]);
          else deltaMin = ux*(xq[k][0] - xc[0]) + uy*(xq[k][1] - xc[1]) + uz*(xq[k][2] - xc[2]);
          if (deltaMin > 0.0) psi = std::min(1., deltaPlusMax/deltaMin);
          if (deltaMin < 0.0) psi = std::min(1., deltaPlusMin/deltaMin);
          psimin = std::min(psi, psimin);
        }
      }
      lim[iVar] = psimin;
    }
    // Venktn2D option "2DNoLimiterID" (Venktn2D.cxx:253-258): cancels the limiter of the listed variables
    if (D == 2 && cfg.limiter == LIM_VENKTN)
      for (int iVar = 0; iVar < Q; ++iVar) if (cfg.noLimiterVar[iVar]) lim[iVar] = 1.0;
  }

  // FVMCC_PolyRec.cxx:120-167 computeFaceLimiter (first touch per cell through cellFlag)
  void computeFaceLimiter(int f, const double* U, Outputs& o) {
    if (cfg.limiter == LIM_NONE) return;
    for (int iCell = 0; iCell < 2; ++iCell) {
      const int c = (iCell == 0) ? mesh.faceL[f] : mesh.faceR[f];
      if (c >= mesh.nCells) continue;   // ghost
      if (cellFlag[c]) continue;
      if (cfg.residual > cfg.limitRes && (cfg.limitIter > 0 && cfg.iter < cfg.limitIter)) {
        for (int i = 0; i < Q; ++i) o.limiter[(size_t)c*Q + i] = 1.0;
        limitCell(c, U, o, &o.limiter[(size_t)c*Q]);
      } else if (!cfg.freezeLimiter) {
        double tmp[Q];
        limitCell(c, U, o, tmp);
        for (int i = 0; i < Q; ++i)
          o.limiter[(size_t)c*Q + i] = std::min(tmp[i], o.limiter[(size_t)c*Q + i]);
      }
    }
  }

  // ---- reconstruction (a8) -------------------------------------------------
  struct FaceCtx {
    int f, L, R; bool isB;
    const double* U;      // the state array of the evaluation (RoeSA reads the neighbours' states)
    double xq[3];
    double n[3];          // unit normal (FVMCC_ComputeRHS.cxx:518-534)
    double area;
    double vL[Q], vR[Q];  // extrapolated values
    double bL[Q], bR[Q];  // backups
    PData pL, pR;
    // diffusive part
    int nbCV;             // vertices in the control volume
    const double* cvState[6];
    double muBkp, lambdaBkp;
    double pn1[5][3], pn2[5][3];  // pyramid normals (rows 1..4 used) / tetra normals
    double n2d[4][2];
    double volume;
    // Corrected{2,3}D derivative (CorrectedDerivative2D.hh:150-190)
    double dr, eRLdotNN[3], oEminusNN[3][3];
    int nCorr, corrStencil[128];   // _stencil of a boundary face (state indices)
    // a19: d grad / d state of the left / right cell (DerivativeComputer::getGradientsJacob) and the average state of
    // the unperturbed evaluation, what NSFluxAndJacob::computeFlux (NSFluxAndJacob.cxx:106-118) differentiates
    double gradJacob[2][3], avState[Q];
  };

  double gradCoeff(int stateID, int iVar, const double* xq, const double* xs, const Outputs& o) const {
    if (D == 2)
      return cfg.gradientFactor*(o.uX[(size_t)stateID*Q+iVar]*(xq[0] - xs[0]) +
                                 o.uY[(size_t)stateID*Q+iVar]*(xq[1] - xs[1]));
    return cfg.gradientFactor*(o.uX[(size_t)stateID*Q+iVar]*(xq[0] - xs[0]) +
                               o.uY[(size_t)stateID*Q+iVar]*(xq[1] - xs[1]) +
                               o.uZ[(size_t)stateID*Q+iVar]*(xq[2] - xs[2]));
  }

  // LeastSquareP1PolyRec2D.cxx:158-217 / 3D.cxx:204-268 / ConstantPolyRec.cxx:43-47
  void extrapolate(FaceCtx& x, const double* U, Outputs& o, bool withLimiter) {
    faceMidPoint(x.f, x.xq);
    const double* sL = &U[(size_t)x.L*Q];
    const double* sR = &U[(size_t)x.R*Q];
    if (cfg.polyRec == PR_CONSTANT) {
      for (int i = 0; i < Q; ++i) { x.vL[i] = sL[i]; x.vR[i] = sR[i]; x.bL[i] = sL[i]; x.bR[i] = sR[i]; }
      return;
    }
    if (withLimiter) computeFaceLimiter(x.f, U, o);
    const int gradR = x.isB ? x.L : x.R;
    const double* xL = &mesh.centres[(size_t)x.L*D];
    const double* xR = &mesh.centres[(size_t)x.R*D];
    const int32_t* zg = x.isB ? bc[mesh.faceBC[x.f]].zeroGrad : nullptr;
    for (int iVar = 0; iVar < Q; ++iVar) {
      const bool z = zg && zg[iVar];
      const double gcS = (!z) ? gradCoeff(x.L, iVar, x.xq, xL, o) : 0.0;
      const double gcN = (!z) ? gradCoeff(gradR, iVar, x.xq, xR, o) : 0.0;
      const double lValue = (cfg.stopLimiting == 0) ? o.limiter[(size_t)x.L*Q + iVar] : 1.;
      x.vL[iVar] = sL[iVar] + lValue*gcS;
      const double rValue = (cfg.stopLimiting == 0) ? o.limiter[(size_t)gradR*Q + iVar] : 1.;
      x.vR[iVar] = sR[iVar] + rValue*gcN;
      x.bL[iVar] = x.vL[iVar]; x.bR[iVar] = x.vR[iVar];
    }
  }

  // LeastSquareP1PolyRec2D.cxx:221-262 / 3D.cxx:272-318 / FVMCC_PolyRec.hh:149-157
  void extrapolateVar(FaceCtx& x, const double* U, const Outputs& o, int iVar, int leftOrRight) {
    for (int i = 0; i < Q; ++i) { x.vL[i] = x.bL[i]; x.vR[i] = x.bR[i]; }   // copyBackupValues
    const int st = leftOrRight == 0 ? x.L : x.R;
    double* v = leftOrRight == 0 ? x.vL : x.vR;
    if (cfg.polyRec == PR_CONSTANT) { v[iVar] = U[(size_t)st*Q + iVar]; return; }
    const int stateID = (leftOrRight == 0) ? x.L : (x.isB ? x.L : x.R);
    const double* xs = &mesh.centres[(size_t)st*D];
    const int32_t* zg = x.isB ? bc[mesh.faceBC[x.f]].zeroGrad : nullptr;
    const bool z = zg && zg[iVar];
    const double gc = (!z) ? gradCoeff(stateID, iVar, x.xq, xs, o) : 0.0;
    const double lValue = (cfg.stopLimiting == 0) ? o.limiter[(size_t)stateID*Q + iVar] : 1.;
    v[iVar] = U[(size_t)st*Q + iVar] + lValue*gc;
  }

  // RoeSAFlux::updateEtaPA (RoeSAFlux.cxx:146-190).  No partition faces in a single domain (`_isPartitionFace` all false).
  template <class PD>
  void updateEtaPAImpl(const FaceCtx& x, int cell, const double* eValues, double& etaSA, PD*) {
    const int nbC = mesh.faceNodePtr[x.f+1] - mesh.faceNodePtr[x.f];
    const int32_t* fnC = &mesh.faceNodes[mesh.faceNodePtr[x.f]];
    for (int k = mesh.cellFacePtr[cell]; k < mesh.cellFacePtr[cell+1]; ++k) {
      const int nf = mesh.cellFaces[k];
      bool shares = false;                                 // GeometricEntity::sharesNodesWith
      const int32_t* fnN = &mesh.faceNodes[mesh.faceNodePtr[nf]];
      const int nbN = mesh.faceNodePtr[nf+1] - mesh.faceNodePtr[nf];
      for (int a = 0; a < nbN && !shares; ++a)
        for (int b = 0; b < nbC; ++b) if (fnN[a] == fnC[b]) { shares = true; break; }
      if ((shares && !cfg.roeSaOldStencil) || cfg.roeSaOldStencil) {
        const int rs = (mesh.faceL[nf] == cell) ? mesh.faceR[nf] : mesh.faceL[nf];
        const double invArea = 1./mesh.areas[nf];
        double nrm[3] = {0., 0., 0.};
        for (int i = 0; i < D; ++i) nrm[i] = mesh.normals[(size_t)nf*D + i]*invArea;
        PD dataR;
        P::computePhysicalData(cfg, &x.U[(size_t)rs*Q], dataR);
        double tmpEv[Q];
        P::eigenValues(dataR, nrm, tmpEv);
        for (int i = 0; i < Q; ++i) etaSA = std::max(etaSA, std::abs(tmpEv[i] - eValues[i]));
      }
    }
  }
  void updateEtaPA(const FaceCtx& x, int cell, const double* eValues, double& etaSA) {
    updateEtaPAImpl(x, cell, eValues, etaSA, (PData*)nullptr);
  }

  // ---- convective flux incl. face integration (FVMCC_FluxSplitter.hh:79-83) and, when
  // !isPerturb, the update coefficient (RoeFlux.cxx:165-185)
  bool convectiveFlux(FaceCtx& x, bool isPerturb, Outputs& o, double* flux) {
    P::computePhysicalData(cfg, x.vL, x.pL);   // computeStatesData, FVMCC_ComputeRHS.hh:112-118
    P::computePhysicalData(cfg, x.vR, x.pR);
    // both extrapolated states sit at the face quadrature point (FVMCC_PolyRec::baseExtrapolateImpl, FVMCC_PolyRec.cxx:170-185)
    P::setCoordinates(x.pL, x.xq); P::setCoordinates(x.pR, x.xq);
    for (int i = 0; i < Q; ++i) flux[i] = 0.;
    if (cfg.fluxSplitter == FS_ROE) {
      if (!P::roeFlux(cfg, x.vL, x.vR, x.pL, x.pR, x.n, flux)) { o.status = 1; return false; }
    } else if (cfg.fluxSplitter == FS_ROESA) {
      // RoeSAFlux::setAbsEigenValues (RoeSAFlux.cxx:66-98): eigenvalues of the two reconstructed states against those of
      // the states across the faces of the left and of the right cell; the states are read where they live (x.U: a
      // perturbed cell state is seen through the current face's own entry, as in the reference)
      double evL[Q], evR[Q];
      P::eigenValues(x.pL, x.n, evL);
      P::eigenValues(x.pR, x.n, evR);
      double etaSA = 0.0;
      updateEtaPA(x, x.L, evL, etaSA);
      if (!x.isB) updateEtaPA(x, x.R, evR, etaSA);
      etaSA *= 0.5;
      if (!(std::abs(etaSA) > 0.)) { o.status = 1; return false; }     // assert(std::abs(etaSA) > 0.), :97
      if (!P::roeFlux(cfg, x.vL, x.vR, x.pL, x.pR, x.n, flux, etaSA)) { o.status = 1; return false; }
    } else if (cfg.fluxSplitter == FS_LAXFRIED || cfg.fluxSplitter == FS_HLL || cfg.fluxSplitter == FS_CENTRED) {
      // update coefficient (LaxFriedFlux.cxx:120-188, HLLFlux.cxx:135-150): _leftEv.max() == un + a, as Roe's
      P::centredFlux(cfg, x.vL, x.vR, x.pL, x.pR, x.n, flux);
    } else {
      P::ausmFlux(cfg, x.pL, x.pR, x.n, flux);      // update coefficient: AUSMFlux.ci:410-440, same form as Roe's
    }
    if (!isPerturb) {
      const double faceArea = x.area/1;       // polyRec->nbQPoints() == 1
      // left side: getMaxEigenValue (Roe, AUSM) or _leftEv.max() (LaxFriedFlux.cxx:131, HLLFlux.cxx:141); right side:
      // getMaxEigenValue with the reversed normal in all splitters; Centred: getMaxEigenValue on both sides (CentredFlux.cxx:75-87)
      const bool centred = (cfg.fluxSplitter == FS_LAXFRIED || cfg.fluxSplitter == FS_HLL);
      double maxEV = centred ? P::evMaxOfVector(cfg, x.pL, x.n) : P::maxEigenValue(cfg, x.pL, x.n);
      o.updateCoeff[x.L] += std::max(maxEV, 0.)*faceArea;
      if (!x.isB) {
        double tn[3];
        for (int i = 0; i < D; ++i) tn[i] = -1.0*x.n[i];
        maxEV = P::maxEigenValue(cfg, x.pR, tn);
        o.updateCoeff[x.R] += std::max(maxEV, 0.)*faceArea;
      }
    }
    for (int i = 0; i < Q; ++i) flux[i] *= x.area;
    return true;
  }

  // ---- diffusive flux (a14,a15): NSFlux.ci:59-156 with DiamondVolume{2,3}DDerivative
  // ---- a15, second DerivativeComputer: CorrectedDerivative2D.cxx:91-430 / CorrectedDerivative3D.cxx:54-220 inside
  // NSFlux.ci:59-156.  The face gradient is the mean of the two cell gradients (the least-squares uX/uY/uZ of the
  // update variables, frozen during the perturbations) with its normal part replaced by the two-point difference:
  //   grad = (vR - vL)/|xR - xL| * (e.n) n + (I - n n) 0.5 (gradL + gradR)
  // On a boundary face gradR is a local least-squares gradient at the ghost centre over the left cell, the other ghosts
  // of the left cell's stencil and those stencil cells that share a node with the face.
  void diffusiveFluxCorrected(FaceCtx& x, const double* U, bool isPerturb, int iPerturbVar, Outputs& o, double* dflux) {
    const int nb = mesh.faceNodePtr[x.f+1] - mesh.faceNodePtr[x.f];
    const int32_t* fn = &mesh.faceNodes[mesh.faceNodePtr[x.f]];
    const double* cL = &mesh.centres[(size_t)x.L*D];
    const double* cR = &mesh.centres[(size_t)x.R*D];
    if (!isPerturb) {
      // computeControlVolume, CorrectedDerivative2D.cxx:262-331 (+ setControlVolumeStates :352-430)
      if (x.isB) {
        x.nCorr = 0;
        x.corrStencil[x.nCorr++] = x.L;
        for (int in = mesh.stencilPtr[x.L]; in < mesh.stencilPtr[x.L+1]; ++in) {
          const int last = mesh.stencil[in];
          bool take = false;
          if (last >= mesh.nCells) take = (last != x.R);
          else {
            for (int j = 0; j < nb && !take; ++j)        // cell->containNode(face->getNode(iNode))
              for (int k = mesh.nodeNbrPtr[fn[j]]; k < mesh.nodeNbrPtr[fn[j]+1]; ++k)
                if (mesh.nodeNbr[k] == last) { take = true; break; }
          }
          if (take) {
            if (x.nCorr >= 128) throw std::runtime_error("Corrected derivative: boundary stencil larger than 128");
            x.corrStencil[x.nCorr++] = last;
          }
        }
      }
      x.dr = getDistance<D>(cL, cR);
      double eRL[3] = {0., 0., 0.};
      for (int k = 0; k < D; ++k) eRL[k] = (cR[k] - cL[k])/x.dr;
      double eDotN = 0.0;
      for (int k = 0; k < D; ++k) eDotN += eRL[k]*x.n[k];
      for (int k = 0; k < D; ++k) x.eRLdotNN[k] = eDotN*x.n[k];
      // computeOEminusNN, CorrectedDerivative2D.cxx:498-504, 3D.cxx:190-211
      for (int a = 0; a < D; ++a)
        for (int b = 0; b < D; ++b)
          x.oEminusNN[a][b] = (a == b) ? 1.0 - x.n[a]*x.n[a] : -(x.n[std::min(a, b)]*x.n[std::max(a, b)]);
      x.volume = x.isB ? mesh.volumes[x.L] : 0.5*(mesh.volumes[x.L] + mesh.volumes[x.R]);
      // getGradientsJacob, CorrectedDerivative2D.cxx:461-469 (CorrectedDerivative3D has none: rejected at set_config)
      const double ovDr = 1./x.dr;
      for (int k = 0; k < D; ++k) { x.gradJacob[0][k] = (-ovDr)*x.eRLdotNN[k]; x.gradJacob[1][k] = ovDr*x.eRLdotNN[k]; }
    }
    // setGradientVars: values(i,j) = (*states[j])[i]; states = L, R, face nodes, boundary stencil
    const double* sL = &U[(size_t)x.L*Q];
    const double* sR = &U[(size_t)x.R*Q];
    double grad[Q][3];
    double uR[Q][3];
    if (x.isB) {
      // computeBoundaryGradients, CorrectedDerivative2D.cxx:155-257, 3D.cxx:88-176
      double l11 = 0.0, l12 = 0.0, l13 = 0.0, l22 = 0.0, l23 = 0.0, l33 = 0.0;
      double lf1[Q], lf2[Q], lf3[Q];
      for (int i = 0; i < Q; ++i) { lf1[i] = 0.0; lf2[i] = 0.0; lf3[i] = 0.0; }
      for (int in = 0; in < x.nCorr; ++in) {
        const double* nodeLast = &mesh.centres[(size_t)x.corrStencil[in]*D];
        const double deltaR = getDistance<D>(cR, nodeLast);
        const double weight = 1.0/deltaR;
        const double dx = weight*(nodeLast[0] - cR[0]);
        const double dy = weight*(nodeLast[1] - cR[1]);
        const double dz = (D == 3) ? weight*(nodeLast[D-1] - cR[D-1]) : 0.0;
        l11 += dx*dx; l12 += dx*dy; l22 += dy*dy;
        if (D == 3) { l13 += dx*dz; l23 += dy*dz; l33 += dz*dz; }
        const double* sv = &U[(size_t)x.corrStencil[in]*Q];
        for (int i = 0; i < Q; ++i) {
          const double du = weight*(sv[i] - sR[i]);
          lf1[i] += dx*du; lf2[i] += dy*du;
          if (D == 3) lf3[i] += dz*du;
        }
      }
      if (D == 2) {
        const double invDet = 1./(l11*l22 - l12*l12);
        for (int i = 0; i < Q; ++i) {
          uR[i][0] = (l22*lf1[i] - l12*lf2[i])*invDet;
          uR[i][1] = (l11*lf2[i] - l12*lf1[i])*invDet;
          uR[i][2] = 0.;
        }
      } else {
        const double det = (l11*l22*l33 - l11*l23*l23 - l12*l12*l33 + l12*l13*l23 + l13*l12*l23 - l13*l13*l22);
        const double invDet = 1./det;
        const double linv11 = l22*l33 - l23*l23, linv22 = l11*l33 - l13*l13, linv33 = l11*l22 - l12*l12;
        const double linv12 = -(l12*l33 - l13*l23), linv13 = l12*l23 - l13*l22, linv23 = -(l11*l23 - l13*l12);
        for (int i = 0; i < Q; ++i) {
          uR[i][0] = (linv11*lf1[i] + linv12*lf2[i] + linv13*lf3[i])*invDet;
          uR[i][1] = (linv12*lf1[i] + linv22*lf2[i] + linv23*lf3[i])*invDet;
          uR[i][2] = (linv13*lf1[i] + linv23*lf2[i] + linv33*lf3[i])*invDet;
        }
      }
    } else {
      for (int i = 0; i < Q; ++i) {
        uR[i][0] = o.uX[(size_t)x.R*Q + i]; uR[i][1] = o.uY[(size_t)x.R*Q + i];
        uR[i][2] = (D == 3) ? o.uZ[(size_t)x.R*Q + i] : 0.;
      }
    }
    for (int i = 0; i < Q; ++i) {
      const double dv = (sR[i] - sL[i])/x.dr;
      double gradLR[3];
      gradLR[0] = 0.5*(o.uX[(size_t)x.L*Q + i] + uR[i][0]);
      gradLR[1] = 0.5*(o.uY[(size_t)x.L*Q + i] + uR[i][1]);
      gradLR[2] = (D == 3) ? 0.5*(o.uZ[(size_t)x.L*Q + i] + uR[i][2]) : 0.;
      // correctGradient (CorrectedDerivative2D.hh:127-130): grad*_eRLdotNN + _oEminusNN*gradLR, the matrix-vector
      // product accumulating from 0 (MatExprT.hh:356-363)
      for (int a = 0; a < D; ++a) {
        double res = 0.0;
        for (int k = 0; k < D; ++k) res += x.oEminusNN[a][k]*gradLR[k];
        grad[i][a] = dv*x.eRLdotNN[a] + res;
      }
      if (D == 2) grad[i][2] = 0.;
    }
    // computeAverageValues, CorrectedDerivative2D.cxx:335-349: mean of the nodal states of the face
    double avState[Q];
    const double invNbNodes = 1./static_cast<double>(nb);
    for (int i = 0; i < Q; ++i) {
      avState[i] = 0.0;
      for (int j = 0; j < nb; ++j) avState[i] += invNbNodes*o.nstates[(size_t)fn[j]*Q + i];
    }
    bool useBackUp = false, setBackUp = false;
    if (isPerturb && iPerturbVar != P::TID) useBackUp = true;
    else if (!isPerturb) setBackUp = true;
    double mu, lambda;
    if (useBackUp || freezeCoeff) { mu = x.muBkp; lambda = x.lambdaBkp; }
    else {
      mu = P::dynViscosity(cfg, avState)*cfg.artDiffCoeff;
      lambda = P::thermConductivity(cfg, mu);
      if (setBackUp) { x.muBkp = mu; x.lambdaBkp = lambda; }
    }
    P::nsFluxUnit(cfg, avState, grad, x.n, mu, lambda, dflux);
    for (int i = 0; i < Q; ++i) dflux[i] *= x.area;
    if (!isPerturb) {
      for (int i = 0; i < Q; ++i) x.avState[i] = avState[i];
      const double nu = mu/P::nsDensity(cfg, avState);
      const double diffUpdateCoeff = nu*x.area*x.area/x.volume;
      o.updateCoeff[x.L] += diffUpdateCoeff;
      if (!x.isB) o.updateCoeff[x.R] += diffUpdateCoeff;
    }
  }

  // ---- a15, third DerivativeComputer: "StateDiff" (StateDiffDerivative.cxx:56-131) inside NSFlux.ci:59-156: the
  // two-point difference along the line of centres, grad = (vR - vL)/|xR - xL| * (xR - xL)/|xR - xL|; the average state is
  // the mean of the nodal states of the first two face nodes (also in 3D); volume = mean of the cell volumes.  The flat
  // mesh orients every face from its L state (isOutward == L), so `_vectorLR` is never flipped (:98-102).
  void diffusiveFluxStateDiff(FaceCtx& x, const double* U, bool isPerturb, int iPerturbVar, Outputs& o, double* dflux) {
    const int32_t* fn = &mesh.faceNodes[mesh.faceNodePtr[x.f]];
    const double* cL = &mesh.centres[(size_t)x.L*D];
    const double* cR = &mesh.centres[(size_t)x.R*D];
    if (!isPerturb) {
      x.dr = getDistance<D>(cL, cR);
      for (int k = 0; k < D; ++k) x.eRLdotNN[k] = cR[k] - cL[k];       // _vectorLR
      for (int k = 0; k < D; ++k) x.eRLdotNN[k] /= x.dr;
      x.volume = x.isB ? mesh.volumes[x.L] : 0.5*(mesh.volumes[x.L] + mesh.volumes[x.R]);
    }
    const double* sL = &U[(size_t)x.L*Q];
    const double* sR = &U[(size_t)x.R*Q];
    double grad[Q][3];
    for (int i = 0; i < Q; ++i) {
      const double dv = (sR[i] - sL[i])/x.dr;
      for (int ix = 0; ix < D; ++ix) grad[i][ix] = dv*x.eRLdotNN[ix];
      if (D == 2) grad[i][2] = 0.;
    }
    double avState[Q];
    for (int i = 0; i < Q; ++i) avState[i] = 0.5*(o.nstates[(size_t)fn[0]*Q + i] + o.nstates[(size_t)fn[1]*Q + i]);
    bool useBackUp = false, setBackUp = false;
    if (isPerturb && iPerturbVar != P::TID) useBackUp = true;
    else if (!isPerturb) setBackUp = true;
    double mu, lambda;
    if (useBackUp || freezeCoeff) { mu = x.muBkp; lambda = x.lambdaBkp; }
    else {
      mu = P::dynViscosity(cfg, avState)*cfg.artDiffCoeff;
      lambda = P::thermConductivity(cfg, mu);
      if (setBackUp) { x.muBkp = mu; x.lambdaBkp = lambda; }
    }
    P::nsFluxUnit(cfg, avState, grad, x.n, mu, lambda, dflux);
    for (int i = 0; i < Q; ++i) dflux[i] *= x.area;
    if (!isPerturb) {
      const double nu = mu/P::nsDensity(cfg, avState);
      const double diffUpdateCoeff = nu*x.area*x.area/x.volume;
      o.updateCoeff[x.L] += diffUpdateCoeff;
      if (!x.isB) o.updateCoeff[x.R] += diffUpdateCoeff;
    }
  }

  // ---- a15, fourth DerivativeComputer: "CorrectedGG2D" / "CorrectedGG3D" (CorrectedDerivativeGG2D.cxx:108-487,
  // CorrectedDerivativeGG3D.cxx:56-150) inside NSFlux.ci:59-156.  The Corrected face gradient
  //   grad = (vR - vL)/|xR - xL| * (e.n) n + (I - n n) 0.5 (gradL + gradR)
  // with gradL / gradR the Green-Gauss gradients of the two cells over the NODAL states (sum over the cell's faces of
  // the face-averaged nodal state times the outward area normal, divided by the cell volume).  On a boundary face gradR is
  // a least-squares gradient at the ghost centre over the states across the other faces of the left cell that share a node
  // with the face, unweighted unless "useWeights".
  // The control-volume list is states = [L, R, nodal states of the L cell (cell-local order), nodal states of the R cell |
  // boundary stencil].  _currFaceNodeIDs collects `counter` AFTER `states[counter++] = ...` (:317-321), so the "average
  // state of the face nodes" of computeAverageValues (:277-296) reads the entry FOLLOWING each face node of the L cell: the
  // next L-cell node, or the first R-cell node / stencil state after the last one.  Restated as it is.
  static const int* ggFaceDof(int nbCellNodes, int iFace, int& nb) {     // LocalConnectionDataBuilder (as cfo_mesh.hpp faceDof)
    static const int tri[3][2] = {{0,1},{1,2},{2,0}};
    static const int quad[4][2] = {{0,1},{1,2},{2,3},{3,0}};
    static const int tet[4][3] = {{0,1,2},{0,3,1},{1,3,2},{0,2,3}};
    static const int hex[6][4] = {{0,3,2,1},{4,5,6,7},{0,1,5,4},{1,2,6,5},{2,3,7,6},{0,4,7,3}};
    if (D == 2) { nb = 2; return (nbCellNodes == 3) ? tri[iFace] : quad[iFace]; }
    if (nbCellNodes == 4) { nb = 3; return tet[iFace]; }
    if (nbCellNodes == 8) { nb = 4; return hex[iFace]; }
    throw std::runtime_error("CorrectedGG: prisms and pyramids are not provided");
  }
  // computeGradientsGG (CorrectedDerivativeGG2D.cxx:437-476): ux[iDim][iVar]
  void gradientsGG(int cell, const Outputs& o, double ux[3][Q]) const {
    if (!mesh.cellNodePtr) throw std::runtime_error("CorrectedGG needs the cellNodes connectivity");
    const int32_t* cn = &mesh.cellNodes[mesh.cellNodePtr[cell]];
    const int nbCellNodes = mesh.cellNodePtr[cell+1] - mesh.cellNodePtr[cell];
    for (int a = 0; a < 3; ++a) for (int i = 0; i < Q; ++i) ux[a][i] = 0.0;
    const int fb = mesh.cellFacePtr[cell], fe = mesh.cellFacePtr[cell+1];
    for (int iFace = 0; iFace < fe - fb; ++iFace) {
      const int faceID = mesh.cellFaces[fb + iFace];
      double fnormal[3] = {0., 0., 0.};
      for (int iDim = 0; iDim < D; ++iDim) {
        fnormal[iDim] = mesh.normals[(size_t)faceID*D + iDim];
        if (mesh.faceL[faceID] != cell) fnormal[iDim] *= -1.;      // isOutward[faceID] != state->getLocalID()
      }
      int nbNodesInFace;
      const int* dof = ggFaceDof(nbCellNodes, iFace, nbNodesInFace);
      const double invNodesInFace = 1./nbNodesInFace;
      for (int i = 0; i < Q; ++i)
        for (int iNode = 0; iNode < nbNodesInFace; ++iNode) {
          const double vOvN = o.nstates[(size_t)cn[dof[iNode]]*Q + i]*invNodesInFace;
          for (int iDim = 0; iDim < D; ++iDim) ux[iDim][i] += vOvN*fnormal[iDim];
        }
    }
    for (int a = 0; a < D; ++a) for (int i = 0; i < Q; ++i) ux[a][i] /= mesh.volumes[cell];
  }
  void diffusiveFluxCorrectedGG(FaceCtx& x, const double* U, bool isPerturb, int iPerturbVar, Outputs& o, double* dflux) {
    const int nb = mesh.faceNodePtr[x.f+1] - mesh.faceNodePtr[x.f];
    const int32_t* fn = &mesh.faceNodes[mesh.faceNodePtr[x.f]];
    const double* cL = &mesh.centres[(size_t)x.L*D];
    const double* cR = &mesh.centres[(size_t)x.R*D];
    if (!mesh.cellNodePtr) throw std::runtime_error("CorrectedGG needs the cellNodes connectivity");
    const int32_t* cnL = &mesh.cellNodes[mesh.cellNodePtr[x.L]];
    const int nL = mesh.cellNodePtr[x.L+1] - mesh.cellNodePtr[x.L];
    if (!isPerturb) {
      // setControlVolumeStates (:300-397): the boundary stencil
      x.nCorr = 0;
      if (x.isB) {
        for (int k = mesh.cellFacePtr[x.L]; k < mesh.cellFacePtr[x.L+1]; ++k) {
          const int cf = mesh.cellFaces[k];
          if (cf == x.f) continue;
          bool sharesNode = false;
          const int32_t* cfn = &mesh.faceNodes[mesh.faceNodePtr[cf]];
          const int cnb = mesh.faceNodePtr[cf+1] - mesh.faceNodePtr[cf];
          for (int j = 0; j < nb && !sharesNode; ++j)
            for (int m = 0; m < cnb; ++m) if (cfn[m] == fn[j]) { sharesNode = true; break; }
          if (sharesNode) x.corrStencil[x.nCorr++] = (mesh.faceL[cf] != x.L) ? mesh.faceL[cf] : mesh.faceR[cf];
        }
      }
      // computeControlVolume (:222-263)
      x.dr = getDistance<D>(cL, cR);
      double eRL[3] = {0., 0., 0.};
      for (int k = 0; k < D; ++k) eRL[k] = cR[k] - cL[k];
      for (int k = 0; k < D; ++k) eRL[k] /= x.dr;
      double eDotN = 0.0;
      for (int k = 0; k < D; ++k) eDotN += eRL[k]*x.n[k];
      for (int k = 0; k < D; ++k) x.eRLdotNN[k] = eDotN*x.n[k];
      for (int a = 0; a < D; ++a)
        for (int b = 0; b < D; ++b)
          x.oEminusNN[a][b] = (a == b) ? 1.0 - x.n[a]*x.n[a] : -(x.n[std::min(a, b)]*x.n[std::max(a, b)]);
      x.volume = x.isB ? mesh.volumes[x.L] : 0.5*(mesh.volumes[x.L] + mesh.volumes[x.R]);
    }
    const double* sL = &U[(size_t)x.L*Q];
    const double* sR = &U[(size_t)x.R*Q];
    double uxL[3][Q], uxR[3][Q];
    gradientsGG(x.L, o, uxL);
    if (!x.isB) gradientsGG(x.R, o, uxR);
    else {
      // computeBoundaryGradients (2D :146-217, 3D :56-150)
      double l11 = 0.0, l12 = 0.0, l13 = 0.0, l22 = 0.0, l23 = 0.0, l33 = 0.0;
      double lf1[Q], lf2[Q], lf3[Q];
      for (int i = 0; i < Q; ++i) { lf1[i] = 0.0; lf2[i] = 0.0; lf3[i] = 0.0; }
      for (int in = 0; in < x.nCorr; ++in) {
        const double* nodeLast = &mesh.centres[(size_t)x.corrStencil[in]*D];
        const double weight = (!cfg.correctedGGUseWeights) ? 1.0 : 1.0/getDistance<D>(cR, nodeLast);
        const double dx = weight*(nodeLast[0] - cR[0]);
        const double dy = weight*(nodeLast[1] - cR[1]);
        const double dz = (D == 3) ? weight*(nodeLast[D-1] - cR[D-1]) : 0.0;
        l11 += dx*dx; l12 += dx*dy; l22 += dy*dy;
        if (D == 3) { l13 += dx*dz; l23 += dy*dz; l33 += dz*dz; }
        const double* sv = &U[(size_t)x.corrStencil[in]*Q];
        for (int i = 0; i < Q; ++i) {
          const double du = weight*(sv[i] - sR[i]);
          lf1[i] += dx*du; lf2[i] += dy*du;
          if (D == 3) lf3[i] += dz*du;
        }
      }
      if (D == 2) {
        const double invDet = 1./(l11*l22 - l12*l12);
        for (int i = 0; i < Q; ++i) {
          uxR[0][i] = (l22*lf1[i] - l12*lf2[i])*invDet;
          uxR[1][i] = (l11*lf2[i] - l12*lf1[i])*invDet;
          uxR[2][i] = 0.;
        }
      } else {
        const double invDet = 1./(l11*l22*l33 - l11*l23*l23 - l12*l12*l33 + l12*l13*l23 + l13*l12*l23 - l13*l13*l22);
        const double linv11 = l22*l33 - l23*l23, linv22 = l11*l33 - l13*l13, linv33 = l11*l22 - l12*l12;
        const double linv12 = -(l12*l33 - l13*l23), linv13 = l12*l23 - l13*l22, linv23 = -(l11*l23 - l13*l12);
        for (int i = 0; i < Q; ++i) {
          uxR[0][i] = (linv11*lf1[i] + linv12*lf2[i] + linv13*lf3[i])*invDet;
          uxR[1][i] = (linv12*lf1[i] + linv22*lf2[i] + linv23*lf3[i])*invDet;
          uxR[2][i] = (linv13*lf1[i] + linv23*lf2[i] + linv33*lf3[i])*invDet;
        }
      }
    }
    double grad[Q][3];
    for (int i = 0; i < Q; ++i) {
      const double dv = (sR[i] - sL[i])/x.dr;
      double gradLR[3] = {0., 0., 0.};
      for (int a = 0; a < D; ++a) gradLR[a] = 0.5*(uxL[a][i] + uxR[a][i]);
      for (int a = 0; a < D; ++a) {
        double res = 0.0;
        for (int k = 0; k < D; ++k) res += x.oEminusNN[a][k]*gradLR[k];
        grad[i][a] = dv*x.eRLdotNN[a] + res;
      }
      if (D == 2) grad[i][2] = 0.;
    }
    // computeAverageValues (:277-296) with the shifted ids (see above): entry j of the control-volume list
    auto cvEntry = [&](int j) -> const double* {
      if (j == 0) return sL;
      if (j == 1) return sR;
      if (j < 2 + nL) return &o.nstates[(size_t)cnL[j - 2]*Q];
      const int k = j - 2 - nL;
      if (!x.isB) {
        const int nR = mesh.cellNodePtr[x.R+1] - mesh.cellNodePtr[x.R];
        if (k >= nR) throw std::runtime_error("CorrectedGG: average state reads past the control volume");
        return &o.nstates[(size_t)mesh.cellNodes[mesh.cellNodePtr[x.R] + k]*Q];
      }
      if (k >= x.nCorr) throw std::runtime_error("CorrectedGG: average state reads past the control volume");
      return &U[(size_t)x.corrStencil[k]*Q];
    };
    double avState[Q];
    const double invNbNodes = 1./static_cast<double>(nb);
    for (int i = 0; i < Q; ++i) {
      avState[i] = 0.0;
      for (int iNode = 0; iNode < nL; ++iNode) {
        bool inFace = false;
        for (int j = 0; j < nb; ++j) if (fn[j] == cnL[iNode]) inFace = true;    // face->containNode(nodesL[iNode])
        if (inFace) avState[i] += cvEntry(2 + iNode + 1)[i];
      }
      avState[i] *= invNbNodes;
    }
    bool useBackUp = false, setBackUp = false;
    if (isPerturb && iPerturbVar != P::TID) useBackUp = true;
    else if (!isPerturb) setBackUp = true;
    double mu, lambda;
    if (useBackUp || freezeCoeff) { mu = x.muBkp; lambda = x.lambdaBkp; }
    else {
      mu = P::dynViscosity(cfg, avState)*cfg.artDiffCoeff;
      lambda = P::thermConductivity(cfg, mu);
      if (setBackUp) { x.muBkp = mu; x.lambdaBkp = lambda; }
    }
    P::nsFluxUnit(cfg, avState, grad, x.n, mu, lambda, dflux);
    for (int i = 0; i < Q; ++i) dflux[i] *= x.area;
    if (!isPerturb) {
      const double nu = mu/P::nsDensity(cfg, avState);
      const double diffUpdateCoeff = nu*x.area*x.area/x.volume;
      o.updateCoeff[x.L] += diffUpdateCoeff;
      if (!x.isB) o.updateCoeff[x.R] += diffUpdateCoeff;
    }
  }

  void diffusiveFlux(FaceCtx& x, const double* U, bool isPerturb, int iPerturbVar, Outputs& o, double* dflux) {
    if (cfg.derivativeStrategy == 1) { diffusiveFluxCorrected(x, U, isPerturb, iPerturbVar, o, dflux); return; }
    if (cfg.derivativeStrategy == 3) { diffusiveFluxCorrectedGG(x, U, isPerturb, iPerturbVar, o, dflux); return; }
    if (cfg.derivativeStrategy == 2) { diffusiveFluxStateDiff(x, U, isPerturb, iPerturbVar, o, dflux); return; }
    const int nb = mesh.faceNodePtr[x.f+1] - mesh.faceNodePtr[x.f];
    const int32_t* fn = &mesh.faceNodes[mesh.faceNodePtr[x.f]];
    const double* cL = &mesh.centres[(size_t)x.L*D];
    const double* cR = &mesh.centres[(size_t)x.R*D];
    if (!isPerturb) {
      // computeControlVolume: DiamondVolume2DDerivative.cxx:91-137, 3D.cxx:386-411
      if (D == 2) {
        x.nbCV = 4;
        x.cvState[0] = &o.nstates[(size_t)fn[0]*Q]; x.cvState[1] = &U[(size_t)x.L*Q];
        x.cvState[2] = &o.nstates[(size_t)fn[1]*Q]; x.cvState[3] = &U[(size_t)x.R*Q];
        const double* node0 = &mesh.nodes[(size_t)fn[0]*D]; const double* node1 = cL;
        const double* node2 = &mesh.nodes[(size_t)fn[1]*D]; const double* node3 = cR;
        x.n2d[0][0] = -node0[1] + node1[1]; x.n2d[0][1] = -node1[0] + node0[0];
        x.n2d[1][0] = -node1[1] + node2[1]; x.n2d[1][1] = -node2[0] + node1[0];
        x.n2d[2][0] = -node2[1] + node3[1]; x.n2d[2][1] = -node3[0] + node2[0];
        x.n2d[3][0] = -node3[1] + node0[1]; x.n2d[3][1] = -node0[0] + node3[0];
        const double det = (node2[0] - node0[0])*(node3[1] - node1[1]) - (node3[0] - node1[0])*(node2[1] - node0[1]);
        if (det < 0.0) for (int k = 0; k < 4; ++k) { x.n2d[k][0] *= -1.; x.n2d[k][1] *= -1.; }
        x.volume = 0.5*std::abs(det);
        // getGradientsJacob, DiamondVolume2DDerivative.cxx:153-159
        for (int k = 0; k < 2; ++k) {
          x.gradJacob[0][k] = (0.5/x.volume)*(x.n2d[0][k] + x.n2d[1][k]);
          x.gradJacob[1][k] = (0.5/x.volume)*(x.n2d[3][k] + x.n2d[2][k]);
        }
      } else if (nb == 3) {
        x.nbCV = 5;
        for (int k = 0; k < 3; ++k) x.cvState[k] = &o.nstates[(size_t)fn[k]*Q];
        x.cvState[3] = &U[(size_t)x.L*Q]; x.cvState[4] = &U[(size_t)x.R*Q];
        x.volume = mesh.cvol[x.f];
      } else {
        x.nbCV = 6;
        for (int k = 0; k < 4; ++k) x.cvState[k] = &o.nstates[(size_t)fn[k]*Q];
        x.cvState[4] = &U[(size_t)x.L*Q]; x.cvState[5] = &U[(size_t)x.R*Q];
        // compute2PyramidNormals: DiamondVolume3DDerivative.cxx:336-383
        const double* c0 = &mesh.nodes[(size_t)fn[0]*3]; const double* c1 = &mesh.nodes[(size_t)fn[1]*3];
        const double* c2 = &mesh.nodes[(size_t)fn[2]*3]; const double* c3 = &mesh.nodes[(size_t)fn[3]*3];
        double factor = (!mesh.cvSign[x.f]) ? 1.0 : -1.0;
        faceTriagNormal(c0, c1, cL, x.pn1[1]); faceTriagNormal(c1, c2, cL, x.pn1[2]);
        faceTriagNormal(c2, c3, cL, x.pn1[3]); faceTriagNormal(c0, cL, c3, x.pn1[4]);
        for (int r = 1; r < 5; ++r) for (int d = 0; d < 3; ++d) x.pn1[r][d] *= factor;
        faceTriagNormal(c0, c1, cR, x.pn2[1]); faceTriagNormal(c1, c2, cR, x.pn2[2]);
        faceTriagNormal(c2, c3, cR, x.pn2[3]); faceTriagNormal(c0, cR, c3, x.pn2[4]);
        factor *= -1.;
        for (int r = 1; r < 5; ++r) for (int d = 0; d < 3; ++d) x.pn2[r][d] *= factor;
        x.volume = mesh.cvol[x.f];
        // getGradientsJacob, DiamondVolume3DDerivative.cxx:607-625
        const double invVolume = (1./3.)/x.volume;
        for (int ix = 0; ix < 3; ++ix) {
          x.gradJacob[0][ix] = invVolume*(x.pn1[1][ix] + x.pn1[2][ix] + x.pn1[3][ix] + x.pn1[4][ix]);
          x.gradJacob[1][ix] = invVolume*(x.pn2[1][ix] + x.pn2[2][ix] + x.pn2[3][ix] + x.pn2[4][ix]);
        }
      }
    }
    // setGradientVars (NavierStokes3DPvt.cxx:52-63): values(i,j) = (*states[j])[i]
    double values[Q][6];
    for (int i = 0; i < Q; ++i) for (int j = 0; j < x.nbCV; ++j) values[i][j] = x.cvState[j][i];
    double grad[Q][3];
    double avState[Q];
    if (D == 2) {
      const double invVolume = 0.5/x.volume;
      for (int i = 0; i < Q; ++i) {
        const double v01 = values[i][0] + values[i][1];
        const double v12 = values[i][1] + values[i][2];
        const double v23 = values[i][2] + values[i][3];
        const double v30 = values[i][3] + values[i][0];
        for (int ix = 0; ix < 2; ++ix)
          grad[i][ix] = invVolume*(x.n2d[0][ix]*v01 + x.n2d[1][ix]*v12 + x.n2d[2][ix]*v23 + x.n2d[3][ix]*v30);
        grad[i][2] = 0.;
      }
      for (int i = 0; i < Q; ++i) avState[i] = 0.5*(x.cvState[0][i] + x.cvState[2][i]);
    } else if (nb == 3) {
      // compute2TetraGradient: DiamondVolume3DDerivative.cxx:190-263 (normals recomputed on every call)
      const double* node0 = &mesh.nodes[(size_t)fn[0]*3]; const double* node1 = &mesh.nodes[(size_t)fn[1]*3];
      const double* node2 = &mesh.nodes[(size_t)fn[2]*3]; const double* node3 = cL; const double* node4 = cR;
      double factor = (!mesh.cvSign[x.f]) ? -0.5 : 0.5;
      double v1[3], v2[3], v3[3], n031[3], n132[3], n023[3], n041[3], n142[3], n024[3];
      auto nrm = [&](const double* a, const double* b, const double* c2_, const double* d, double fct, double* out) {
        for (int k = 0; k < 3; ++k) { v1[k] = a[k] - b[k]; v2[k] = c2_[k] - d[k]; }
        crossProd(v1, v2, v3);
        for (int k = 0; k < 3; ++k) out[k] = fct*v3[k];
      };
      nrm(node3, node0, node1, node0, factor, n031);
      nrm(node3, node1, node2, node1, factor, n132);
      nrm(node2, node0, node3, node0, factor, n023);
      factor *= -1;
      nrm(node4, node0, node1, node0, factor, n041);
      nrm(node4, node1, node2, node1, factor, n142);
      nrm(node2, node0, node4, node0, factor, n024);
      const double invVolume = (1./3.)/x.volume;
      if (!isPerturb)    // getGradientsJacob, DiamondVolume3DDerivative.cxx:611-614 (the normals of the last evaluation)
        for (int ix = 0; ix < 3; ++ix) {
          x.gradJacob[0][ix] = invVolume*(n031[ix] + n132[ix] + n023[ix]);
          x.gradJacob[1][ix] = invVolume*(n041[ix] + n142[ix] + n024[ix]);
        }
      for (int i = 0; i < Q; ++i) {
        const double v031 = values[i][0] + values[i][3] + values[i][1];
        const double v132 = values[i][1] + values[i][3] + values[i][2];
        const double v023 = values[i][0] + values[i][2] + values[i][3];
        const double v041 = values[i][0] + values[i][4] + values[i][1];
        const double v142 = values[i][1] + values[i][4] + values[i][2];
        const double v024 = values[i][0] + values[i][2] + values[i][4];
        for (int ix = 0; ix < 3; ++ix)
          grad[i][ix] = invVolume*(n031[ix]*v031 + n132[ix]*v132 + n023[ix]*v023 +
                                   n041[ix]*v041 + n142[ix]*v142 + n024[ix]*v024);
      }
      for (int i = 0; i < Q; ++i) avState[i] = (1./3.)*(x.cvState[0][i] + x.cvState[1][i] + x.cvState[2][i]);
    } else {
      // compute2PyramidGradient: DiamondVolume3DDerivative.cxx:267-298
      const double invVolume = (1./3.)/x.volume;
      for (int i = 0; i < Q; ++i) {
        const double v014 = values[i][0] + values[i][1] + values[i][4];
        const double v124 = values[i][1] + values[i][2] + values[i][4];
        const double v234 = values[i][2] + values[i][3] + values[i][4];
        const double v043 = values[i][0] + values[i][4] + values[i][3];
        const double v015 = values[i][0] + values[i][1] + values[i][5];
        const double v125 = values[i][1] + values[i][2] + values[i][5];
        const double v235 = values[i][2] + values[i][3] + values[i][5];
        const double v053 = values[i][0] + values[i][5] + values[i][3];
        for (int ix = 0; ix < 3; ++ix)
          grad[i][ix] = invVolume*(x.pn1[1][ix]*v014 + x.pn1[2][ix]*v124 + x.pn1[3][ix]*v234 + x.pn1[4][ix]*v043 +
                                   x.pn2[1][ix]*v015 + x.pn2[2][ix]*v125 + x.pn2[3][ix]*v235 + x.pn2[4][ix]*v053);
      }
      for (int i = 0; i < Q; ++i)
        avState[i] = 0.25*(x.cvState[0][i] + x.cvState[1][i] + x.cvState[2][i] + x.cvState[3][i]);
    }
    // setComposition (NavierStokes3DPvt.cxx:206-223, no chemistry library) +
    // computeTransportProperties (NavierStokesVarSet.cxx:70-92)
    bool useBackUp = false, setBackUp = false;
    if (isPerturb && iPerturbVar != P::TID) useBackUp = true;
    else if (!isPerturb) setBackUp = true;
    double mu, lambda;
    if (useBackUp || freezeCoeff) { mu = x.muBkp; lambda = x.lambdaBkp; }
    else {
      mu = P::dynViscosity(cfg, avState)*cfg.artDiffCoeff;
      lambda = P::thermConductivity(cfg, mu);
      if (setBackUp) { x.muBkp = mu; x.lambdaBkp = lambda; }
    }
    P::nsFluxUnit(cfg, avState, grad, x.n, mu, lambda, dflux);
    for (int i = 0; i < Q; ++i) dflux[i] *= x.area;
    if (!isPerturb) {
      for (int i = 0; i < Q; ++i) x.avState[i] = avState[i];
      const double nu = mu/P::nsDensity(cfg, avState);
      const double diffUpdateCoeff = nu*x.area*x.area/x.volume;
      o.updateCoeff[x.L] += diffUpdateCoeff;
      if (!x.isB) o.updateCoeff[x.R] += diffUpdateCoeff;
    }
  }

  // NumericalJacobian.hh:131-140,224-228
  double computeEps(int iVar, double value) const {
    const double sgn = (value < 0.0) ? -1.0 : 1.0;
    return cfg.fdTol*sgn*std::max(std::abs(value), std::abs(cfg.refValues[iVar]));
  }

  int findSlot(const Outputs& o, int row, int col) const {
    for (int k = o.bsrRowPtr[row]; k < o.bsrRowPtr[row+1]; ++k) if (o.bsrColInd[k] == col) return k;
    throw std::runtime_error("BSR slot not found");
  }
  // BlockAccumulator::addValues(i,j,jb,values) -> column jb of block (row,col): [ib][jb]
  void addColumn(Outputs& o, int row, int col, int jb, const double* v) const {
    double* blk = &o.bsrVal[(size_t)findSlot(o, row, col)*Q*Q];
    for (int ib = 0; ib < Q; ++ib) blk[ib*Q + jb] += v[ib];
  }

  // ---- source term (Data.SourceTerm = MHDConsACAST): MHDConsACASourceTerm.cxx:164-211, the part that runs when
  // gravity = Manchester = RadiativeLossTerm = 0 (:211 returns before the solar terms).  Reads the CELL state.
  void computeSource(int cell, const double* U, double* source) const {
    for (int i = 0; i < Q - 1; ++i) source[i] = 0.0;
    const double refSpeedSq = cfg.refSpeed*cfg.refSpeed;
    if (cfg.mhdDissipCoeff != 0.0) {                       // correctionType == "Mixed"
      const double dissipCoeffSq = cfg.mhdDissipCoeff*cfg.mhdDissipCoeff;
      source[Q-1] = -(refSpeedSq/dissipCoeffSq)*U[(size_t)cell*Q + Q - 1]*mesh.volumes[cell];
    } else {
      source[Q-1] = 0.0;                                    // "Hyperbolic"
    }
  }
  // FVMCC_ComputeRhsJacob.cxx:359-364 / :400-405 with addSourceTermNumJacob (.hh:96-105): the state of `cell` is perturbed
  // in place by the caller; _upStFactor = -resFactor (FVMCC_ComputeRHS.hh:143)
  void addSourceJacobColumn(Outputs& o, int cell, const double* U, const double* source, int iVar, double eps) const {
    double pertSource[Q], sourceDiffSum[Q];
    computeSource(cell, U, pertSource);
    const double invEps = 1./eps;
    for (int i = 0; i < Q; ++i) sourceDiffSum[i] = 0.0;
    for (int i = 0; i < Q; ++i) sourceDiffSum[i] += (pertSource[i] - source[i])*invEps;
    for (int i = 0; i < Q; ++i) sourceDiffSum[i] *= -cfg.resFactor;
    addColumn(o, cell, cell, iVar, sourceDiffSum);
  }

  // ---- the command (a17, a18): FVMCC_ComputeRHS.cxx:139-325, FVMCC_ComputeRhsJacob.cxx:126-443
  // U holds cell states followed by ghost states ([nCells+nGhosts][q]); ghost states must already
  // be set (CellCenterFVM.cxx:464 applyBC before execute); U is perturbed and restored in place.
  void execute(double* U, Outputs& o) {
    const int nc = mesh.nCells;
    o.status = 0;
    // initializeComputationRHS (FVMCC_ComputeRHS.cxx:690-705 / FVMCC_ComputeRhsJacob.cxx:410-429)
    std::fill(o.rhs, o.rhs + (size_t)nc*Q, 0.0);
    cellFlag.assign(nc, 0);
    const bool doJac = cfg.computeJacobian && o.bsrVal;
    if (doJac) std::fill(o.bsrVal, o.bsrVal + (size_t)o.bsrRowPtr[nc]*Q*Q, 0.0);
    if (cfg.polyRec == PR_LINEARLS) {
      if (weights.empty()) updateWeights();
      computeGradients(U, o);
    }
    extrapolateInAllNodes(U, o);

    if (o.divB) for (int i = 0; i < nc; ++i) o.divB[i] = 0.;
    FaceCtx x;
    double flux[Q], dFlux[Q], pertFlux[Q], fluxDiff[Q];
    for (int f = 0; f < mesh.nFaces; ++f) {
      x.f = f; x.L = mesh.faceL[f]; x.R = mesh.faceR[f]; x.isB = (x.R >= nc); x.U = U;
      const bool upL = mesh.updatable[x.L];
      const bool upR = (!x.isB) && mesh.updatable[x.R];
      if (!(upL || upR)) continue;                     // FVMCC_ComputeRHS.cxx:212-213
      // setFaceIntegratorData
      const double invArea = 1./mesh.areas[f];
      for (int i = 0; i < D; ++i) x.n[i] = mesh.normals[(size_t)f*D + i]*invArea;
      x.area = mesh.areas[f];
      extrapolate(x, U, o, true);
      if (!convectiveFlux(x, false, o, flux)) return;
      const bool hasDiff = (cfg.diffusive == DIFF_NS);
      if (hasDiff) {
        freezeCoeff = false;
        diffusiveFlux(x, U, false, 0, o, dFlux);
        for (int i = 0; i < Q; ++i) flux[i] -= dFlux[i];
      }
      if (o.divB && P::Q == 9) {
        // a22: FVMCC_ComputeRHSMHD<BASE>::computeInterConvDiff (FiniteVolumeMHD/FVMCC_ComputeRHSMHD.ci:160-203)
        const double refSpeedSq = cfg.refSpeed*cfg.refSpeed;
        o.divB[x.L] += flux[Q-1]/(refSpeedSq*mesh.volumes[x.L]);
        if (!x.isB) o.divB[x.R] -= flux[Q-1]/(refSpeedSq*mesh.volumes[x.R]);
      }
      // computeSourceTerm (FVMCC_ComputeRHS.cxx:538-600): at the first face that touches the cell (cellFlag), before this
      // face's flux is added; invR = 1 (non axisymmetric)
      bool sourceJacobOnCell[2] = {false, false};
      double source[2][Q];
      if (cfg.sourceTerm) {
        for (int iCell = 0; iCell < 2; ++iCell) {
          if (iCell == 1 && x.isB) continue;
          const int cellID = (iCell == 0) ? x.L : x.R;
          if (cellFlag[cellID]) continue;
          computeSource(cellID, U, source[iCell]);
          for (int iEq = 0; iEq < Q; ++iEq) o.rhs[(size_t)cellID*Q + iEq] += cfg.resFactor*source[iCell][iEq]*1.0;
          cellFlag[cellID] = 1;
          sourceJacobOnCell[iCell] = true;
        }
      }
      // updateRHS (FVMCC_ComputeRHS.cxx:455-514), non axisymmetric: _rMid = _invr = 1
      {
        const double coeff = cfg.resFactor*1.0;
        double rFlux[Q];
        for (int i = 0; i < Q; ++i) rFlux[i] = coeff*flux[i];
        for (int i = 0; i < Q; ++i) o.rhs[(size_t)x.L*Q + i] -= rFlux[i]*1.0;
        if (!x.isB) for (int i = 0; i < Q; ++i) o.rhs[(size_t)x.R*Q + i] += rFlux[i]*1.0;
        cellFlag[x.L] = 1; if (!x.isB) cellFlag[x.R] = 1;
      }
      if (!doJac) continue;
      freezeCoeff = cfg.freezeDiffCoeff != 0;
      if (!x.isB && cfg.computeJacobian == 2) {
        // a19: FVMCC_ComputeRhsJacobAnalytic.cxx:68-246 (non axisymmetric, no diffusive term, no source term) with the
        // flux Jacobians integrated over the face by FVMCC_FluxSplitter::integrateFluxAndJacob (FVMCC_FluxSplitter.hh:88-96)
        double JL[Q*Q], JR[Q*Q], colv[Q];
        if (cfg.updateVar == UV_PVT) {          // AUSMPlusUpFlux.ci:1238-1273
          P::ausmPlusUpJacobPvt(cfg, x.pL, x.pR, x.n, 0, JL);
          P::ausmPlusUpJacobPvt(cfg, x.pL, x.pR, x.n, 1, JR);
        } else {
          P::ausmPlusUpJacobCons(cfg, x.pL, x.pR, x.n, 0, JL);
          P::ausmPlusUpJacobCons(cfg, x.pL, x.pR, x.n, 1, JR);
        }
        for (int k = 0; k < Q*Q; ++k) { JL[k] *= x.area; JR[k] *= x.area; }
        if (cfg.diffusive != DIFF_NONE) {
          // NSFluxAndJacob.cxx:106-118, then "note the opposite sign" FVMCC_ComputeRhsJacobAnalytic.cxx:83-91
          double DL[Q*Q], DR[Q*Q];
          P::nsFluxJacobian(cfg, x.avState, x.gradJacob[0], x.n, x.muBkp, x.lambdaBkp, DL);
          P::nsFluxJacobian(cfg, x.avState, x.gradJacob[1], x.n, x.muBkp, x.lambdaBkp, DR);
          for (int k = 0; k < Q*Q; ++k) { DL[k] *= x.area; DR[k] *= x.area; }
          for (int k = 0; k < Q*Q; ++k) { JL[k] -= DL[k]; JR[k] -= DR[k]; }
        }
        auto addBlock = [&](int rowCell, int colCell, const double* Jm) {
          for (int j = 0; j < Q; ++j) {
            for (int i = 0; i < Q; ++i) colv[i] = Jm[i*Q + j];
            addColumn(o, rowCell, colCell, j, colv);
          }
        };
        if (upL && upR) {                       // noaxiBothJacobTerms
          for (int k = 0; k < Q*Q; ++k) { JL[k] *= cfg.resFactor; JR[k] *= cfg.resFactor; }
          addBlock(x.L, x.L, JL); addBlock(x.L, x.R, JR);
          for (int k = 0; k < Q*Q; ++k) { JL[k] *= -1.; JR[k] *= -1.; }
          addBlock(x.R, x.L, JL); addBlock(x.R, x.R, JR);
        } else {                                // computeJacobTerm(idx) -> noaxiJacobTerm
          const int idx = upL ? 0 : 1;
          const double factor = std::pow(-1., (double)idx)*cfg.resFactor;
          for (int k = 0; k < Q*Q; ++k) { JL[k] *= factor; JR[k] *= factor; }
          const int rowCell = idx == 0 ? x.L : x.R;
          addBlock(rowCell, x.L, JL); addBlock(rowCell, x.R, JR);
        }
        continue;
      }
      if (!x.isB) {
        // computeBothJacobTerms / computeJacobTerm (FVMCC_ComputeRhsJacob.cxx:126-236)
        double upFactor[2];
        if (upL && upR) { upFactor[0] = cfg.resFactor; upFactor[1] = -1.0; }
        else {
          const int idx = upL ? 0 : 1;
          upFactor[idx] = std::pow(-1., (double)idx)*cfg.resFactor;
        }
        for (int iCell = 0; iCell < 2; ++iCell) {
          const int st = (iCell == 0) ? x.L : x.R;
          for (int iVar = 0; iVar < Q; ++iVar) {
            double& val = U[(size_t)st*Q + iVar];
            const double orig = val;
            const double eps = computeEps(iVar, val);
            val += eps;
            // computeConvDiffFluxes (FVMCC_ComputeRhsJacob.cxx:321-343)
            extrapolateVar(x, U, o, iVar, iCell);
            if (!convectiveFlux(x, true, o, pertFlux)) { val = orig; return; }
            if (hasDiff) {
              diffusiveFlux(x, U, true, iVar, o, dFlux);
              for (int i = 0; i < Q; ++i) pertFlux[i] -= dFlux[i];
            }
            const double invEps = 1./eps;
            for (int i = 0; i < Q; ++i) fluxDiff[i] = (pertFlux[i] - flux[i])*invEps;
            if (upL && upR) {
              for (int i = 0; i < Q; ++i) fluxDiff[i] *= upFactor[0];
              addColumn(o, x.L, st, iVar, fluxDiff);
              for (int i = 0; i < Q; ++i) fluxDiff[i] *= upFactor[1];
              addColumn(o, x.R, st, iVar, fluxDiff);
              if (sourceJacobOnCell[iCell]) addSourceJacobColumn(o, st, U, source[iCell], iVar, eps);
            } else {
              const int idx = upL ? 0 : 1;
              for (int i = 0; i < Q; ++i) fluxDiff[i] *= upFactor[idx];
              addColumn(o, idx == 0 ? x.L : x.R, st, iVar, fluxDiff);
              if (idx == iCell && sourceJacobOnCell[idx]) addSourceJacobColumn(o, st, U, source[iCell], iVar, eps);
            }
            val = orig;
            (iCell == 0 ? x.vL : x.vR)[iVar] = (iCell == 0 ? x.bL : x.bR)[iVar];   // restoreValues
          }
        }
      } else if (upL) {
        // computeBoundaryJacobianTerm (FVMCC_ComputeRhsJacob.cxx:240-317)
        double origGhost[Q];
        for (int i = 0; i < Q; ++i) origGhost[i] = U[(size_t)x.R*Q + i];
        for (int iVar = 0; iVar < Q; ++iVar) {
          double& val = U[(size_t)x.L*Q + iVar];
          const double orig = val;
          const double eps = computeEps(iVar, val);
          val += eps;
          setGhostState(f, &U[(size_t)x.L*Q], &U[(size_t)x.R*Q]);
          extrapolate(x, U, o, true);   // cellFlag[L] is already set: the limiter is not recomputed
          if (!convectiveFlux(x, true, o, pertFlux)) { val = orig; return; }
          if (hasDiff) {
            diffusiveFlux(x, U, true, iVar, o, dFlux);
            for (int i = 0; i < Q; ++i) pertFlux[i] -= dFlux[i];
          }
          const double invEps = 1./eps;
          for (int i = 0; i < Q; ++i) fluxDiff[i] = (pertFlux[i] - flux[i])*invEps;
          for (int i = 0; i < Q; ++i) fluxDiff[i] *= cfg.resFactor;
          addColumn(o, x.L, x.L, iVar, fluxDiff);
          if (sourceJacobOnCell[0]) addSourceJacobColumn(o, x.L, U, source[0], iVar, eps);
          val = orig;
          for (int i = 0; i < Q; ++i) U[(size_t)x.R*Q + i] = origGhost[i];
        }
      }
    }
  }

  // ---- pseudo-steady time term (a21): FVMCC_PseudoSteadyTimeRhs::execute (FVMCC_PseudoSteadyTimeRhs.cxx:106-175):
  //   diag = V/dt when dt > 0 (unsteady), updateCoeff/CFL otherwise; with useGlobalDT: V/min_cells(V/updateCoeff*CFL)
  //   (the reference takes the minimum over all local states, :121-127); zeroDiagValue drops the term of an equation;
  //   numerical (:197-262) or analytical (:266-310) transformation matrix dU/dP
  void pseudoSteadyTimeRhs(double* U, const double* pastU, Outputs& o) const {
    const bool doJac = cfg.computeJacobian && o.bsrVal;
    const double dt = cfg.dt;
    double minDt = 0.0;
    if (cfg.timeUseGlobalDT) {
      minDt = mesh.volumes[0]/o.updateCoeff[0]*cfg.cfl;
      for (int c = 1; c < mesh.nCells; ++c) minDt = std::min(mesh.volumes[c]/o.updateCoeff[c]*cfg.cfl, minDt);
    }
    for (int c = 0; c < mesh.nCells; ++c) {
      double diagValue;
      if (!cfg.timeUseGlobalDT) diagValue = (dt > 0.0) ? mesh.volumes[c]/dt : o.updateCoeff[c]/cfg.cfl;
      else diagValue = mesh.volumes[c]/minDt;
      if (!mesh.updatable[c]) continue;
      double tempState[Q], tPast[Q], tPert[Q], fluxDiff[Q];
      P::toCons(cfg, &U[(size_t)c*Q], tempState);
      P::toCons(cfg, &pastU[(size_t)c*Q], tPast);
      for (int iEq = 0; iEq < Q; ++iEq) {
        const double dU = tempState[iEq] - tPast[iEq];
        o.rhs[(size_t)c*Q + iEq] -= (!cfg.timeZeroDiag[iEq]) ? dU*diagValue*cfg.resFactor : 0.0;
      }
      if (!doJac) continue;
      if (!cfg.timeAnalyticalMatrix) {
        for (int iVar = 0; iVar < Q; ++iVar) {
          double& val = U[(size_t)c*Q + iVar];
          const double orig = val;
          const double eps = computeEps(iVar, val);
          val += eps;
          P::toCons(cfg, &U[(size_t)c*Q], tPert);
          const double invEps = 1./eps;
          for (int i = 0; i < Q; ++i) fluxDiff[i] = (tPert[i] - tempState[i])*invEps;
          for (int i = 0; i < Q; ++i) fluxDiff[i] *= (!cfg.timeZeroDiag[i]) ? diagValue*cfg.resFactor : 0.0;
          addColumn(o, c, c, iVar, fluxDiff);
          val = orig;
        }
      } else {
        // computeAnalyticalTransMatrix: value = matrix(iVar,jVar)*diag*resFactor added where !zeroDiagValue[jVar] (sic: the
        // flag is tested on the COLUMN index there)
        double M[Q*Q];
        P::updateToSolutionMatrix(cfg, &U[(size_t)c*Q], M);
        for (int jVar = 0; jVar < Q; ++jVar) {
          if (cfg.timeZeroDiag[jVar]) continue;
          for (int iVar = 0; iVar < Q; ++iVar) fluxDiff[iVar] = M[iVar*Q + jVar]*diagValue*cfg.resFactor;
          addColumn(o, c, c, jVar, fluxDiff);
        }
      }
    }
  }

  const Config& cfg;
  const MeshView& mesh;
  const BC* bc;
  std::vector<double> l[6];
  std::vector<double> weights;
  std::vector<uint8_t> cellFlag;
  bool freezeCoeff = false;
};

} // namespace cfo
