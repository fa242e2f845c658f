// cfb200 — a15, second DerivativeComputer: "Corrected2D" / "Corrected3D" inside NSFlux (NSFlux.ci:59-156).
//   CorrectedDerivative2D.cxx:91-150 (inner), :155-257 (boundary), :262-331 (control volume), :335-349 (average state),
//   :352-430 (boundary stencil), :498-504 (I - n n);  CorrectedDerivative3D.cxx:54-86, :88-176, :190-211
// The face gradient of every gradient variable is the mean of the two cell gradients (the least-squares uX/uY/uZ
// sockets: frozen during the finite-difference perturbations) whose normal part is replaced by the two-point difference
//   grad = (vR - vL)/|xR - xL| * (e.n) n + (I - n n) 0.5 (gradL + gradR).
// On a boundary face gradR is a local least-squares gradient at the ghost centre over the host-built stencil
// Dev::bndSt (left cell, the other ghosts of its stencil, the stencil cells that share a node with the face).
// Included by cfb_face.cuh inside namespace cfb.
#pragma once

template <int D>
struct CorrectedWork {
  static constexpr int Q = D + 2;
  static constexpr int NMAX = (D == 2) ? 2 : 4;
  typedef Phys<D> P;
  double n[3], area;
  double dr, eN[3], oe[3][3], volume, mu, lambda;
  double xR[3];
  bool gg;                // "CorrectedGG2D/3D" (CorrectedDerivativeGG2D.cxx, CorrectedDerivativeGG3D.cxx): the cell gradients
                          // are Green-Gauss sums over the nodal states, the boundary least squares runs over the states
                          // across the L cell's faces that touch the face (unweighted unless useWeights), and the
                          // "average state" reads the shifted control-volume entries of Dev::ggAv (see the oracle)
  bool ggWeights;
  bool stateDiff;         // "StateDiff" (StateDiffDerivative.cxx:56-131): grad = dv * (xR - xL)/|xR - xL|, no tangential part,
                          // average state = mean of the nodal states of the first two face nodes, no boundary stencil
  double avState[Q];
  double gradL[Q][3];     // the left cell's gradient
  double tang[Q][3];      // (I - n n) 0.5 (gradL + gradR)
  double grad[Q][3];

  // computeGradientsGG (CorrectedDerivativeGG2D.cxx:437-476): g[var][dim] = (1/V) sum over the cell's faces (local order) and
  // their nodes (cell-local table order) of nstate/nbNodesInFace * outward area normal
  static CFB_HD void ggGrad(const Dev& d, int cell, double (*g)[3]) {
#pragma unroll
    for (int v = 0; v < Q; ++v) { g[v][0] = 0.0; g[v][1] = 0.0; g[v][2] = 0.0; }
    int p = d.ggPtr[cell];
    for (int k = d.cellFacePtr[cell]; k < d.cellFacePtr[cell+1]; ++k) {
      const int cf = d.cellFaces[k];
      const int nbN = d.faceNodePtr[cf+1] - d.faceNodePtr[cf];
      const bool flip = (d.faceL[cf] != cell);               // isOutward[faceID] != state->getLocalID()
      double fnr[3] = {0., 0., 0.};
#pragma unroll
      for (int a = 0; a < D; ++a) { const double x = d.normals[(size_t)a*d.nFaces + cf]; fnr[a] = flip ? x*(-1.) : x; }
      const double inv = ddiv(1., (double)nbN);
      for (int j = 0; j < nbN; ++j) {
        const int node = d.ggNode[p++];
#pragma unroll
        for (int v = 0; v < Q; ++v) {
          const double vOvN = d.nstates[(size_t)v*d.nNodes + node]*inv;
#pragma unroll
          for (int a = 0; a < D; ++a) g[v][a] += vOvN*fnr[a];
        }
      }
    }
    const Recip rv = recipOf(d.volumes[cell]);
#pragma unroll
    for (int v = 0; v < Q; ++v)
#pragma unroll
      for (int a = 0; a < D; ++a) g[v][a] = divBy(g[v][a], rv);
  }

  // computeControlVolume + computeAverageValues + the transport coefficients of the (never perturbed) nodal average
  CFB_HD void setup(const Dev& d, const Cfg& c, int f, int L, int R, bool isB, const double* unitNormal, double faceArea) {
    area = faceArea;
    stateDiff = (c.derivativeStrategy == CFB200_DV_STATEDIFF);
    gg = (c.derivativeStrategy == CFB200_DV_CORRECTED_GG);
    ggWeights = (c.correctedGGUseWeights != 0);
    double xL[3] = {0., 0., 0.};
#pragma unroll
    for (int k = 0; k < 3; ++k) { n[k] = unitNormal[k]; xR[k] = 0.; }
#pragma unroll
    for (int k = 0; k < D; ++k) { xL[k] = d.centres[(size_t)k*d.nStates + L]; xR[k] = d.centres[(size_t)k*d.nStates + R]; }
    const int b = d.faceNodePtr[f];
    const int nbN = d.faceNodePtr[f+1] - b;
    int node[NMAX];
#pragma unroll
    for (int j = 0; j < NMAX; ++j) node[j] = d.faceNodes[b + ((j < nbN) ? j : nbN - 1)];
    double nv[Q][NMAX];
#pragma unroll
    for (int j = 0; j < NMAX; ++j)
#pragma unroll
      for (int v = 0; v < Q; ++v) nv[v][j] = d.nstates[(size_t)v*d.nNodes + node[j]];
#pragma unroll
    for (int v = 0; v < Q; ++v) {
      const size_t iL = (size_t)v*d.nCells + L;
      if (stateDiff || gg) { gradL[v][0] = 0.; gradL[v][1] = 0.; gradL[v][2] = 0.; continue; }   // no LSQ cell gradients read
      gradL[v][0] = d.uX[iL]; gradL[v][1] = d.uY[iL]; gradL[v][2] = (D == 3) ? d.uZ[iL] : 0.;
    }
    const double volL = d.volumes[L];
    const double volR = isB ? 0. : d.volumes[R];
    double dist = 0.;
#pragma unroll
    for (int k = 0; k < D; ++k) { const double diff = xL[k] - xR[k]; dist += diff*diff; }
    dr = dsqrt(dist);
    const Recip rd = recipOf(dr);
    double eDotN = 0.0;
#pragma unroll
    for (int k = 0; k < D; ++k) eDotN += divBy(xR[k] - xL[k], rd)*n[k];
#pragma unroll
    for (int k = 0; k < 3; ++k) eN[k] = (k < D) ? eDotN*n[k] : 0.;
    if (stateDiff) {                                  // _vectorLR, StateDiffDerivative.cxx:89-94
#pragma unroll
      for (int k = 0; k < D; ++k) eN[k] = divBy(xR[k] - xL[k], rd);
    }
#pragma unroll
    for (int a = 0; a < 3; ++a)
#pragma unroll
      for (int k = 0; k < 3; ++k) oe[a][k] = (a == k) ? 1.0 - n[a]*n[a] : -(n[a]*n[k]);
    volume = isB ? volL : 0.5*(volL + volR);
    const double invNbNodes = ddiv(1., (double)nbN);
#pragma unroll
    for (int v = 0; v < Q; ++v) {
      double a = 0.0;
#pragma unroll
      for (int j = 0; j < NMAX; ++j) if (j < nbN) a += invNbNodes*nv[v][j];
      avState[v] = stateDiff ? 0.5*(nv[v][0] + nv[v][1]) : a;       // StateDiffDerivative.cxx:121-128
    }
    if (gg) {
      ggGrad(d, L, gradL);
      // computeAverageValues (CorrectedDerivativeGG2D.cxx:265-285): sum of the listed entries, then * 1/nbNodesInFace
      int code[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) code[j] = d.ggAv[(size_t)j*d.nFaces + f];
#pragma unroll
      for (int v = 0; v < Q; ++v) {
        double a = 0.0;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          if (code[j] == INT32_MIN) continue;
          a += (code[j] >= 0) ? d.nstates[(size_t)v*d.nNodes + code[j]] : d.U[(size_t)v*d.nStates + (~code[j])];
        }
        avState[v] = a*invNbNodes;
      }
    }
    mu = P::dynViscosity(c, avState)*c.artDiffCoeff;
    lambda = P::thermConductivity(c, mu);
  }

  // correctGradient (CorrectedDerivative2D.hh:127-130): the matrix-vector product accumulates from 0 (MatExprT.hh:356-363)
  CFB_HD void tangOf(const double* gl, double* t) const {
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      double res = 0.0;
#pragma unroll
      for (int k = 0; k < D; ++k) res += oe[a][k]*gl[k];
      t[a] = (a < D) ? res : 0.;
    }
  }
  CFB_HD void gradOf(double dv, const double* t, double* g) const {
#pragma unroll
    for (int a = 0; a < 3; ++a) g[a] = (a < D) ? (stateDiff ? dv*eN[a] : dv*eN[a] + t[a]) : 0.;
  }
  CFB_HD void diffFlux(const Cfg& c, const double (*g)[3], double* dflux) const {
    P::nsFlux(c, avState, g, n, mu, lambda, dflux);
#pragma unroll
    for (int v = 0; v < Q; ++v) dflux[v] *= area;
  }
  CFB_HD double diffUpdateCoeff(const Cfg& c) const {      // NSFlux.ci:127-152
    const double nu = ddiv(mu, P::density(c, avState));
    return ddiv(nu*area*area, volume);
  }

  // ---- interior faces -----------------------------------------------------------------------------------------
  CFB_HD void innerTang(const Dev& d, int R) {
    if (stateDiff) {
#pragma unroll
      for (int v = 0; v < Q; ++v) { tang[v][0] = 0.; tang[v][1] = 0.; tang[v][2] = 0.; }
      return;
    }
    if (gg) {
      double gR[Q][3];
      ggGrad(d, R, gR);
#pragma unroll
      for (int v = 0; v < Q; ++v) {
        double gl[3];
#pragma unroll
        for (int a = 0; a < 3; ++a) gl[a] = (a < D) ? 0.5*(gradL[v][a] + gR[v][a]) : 0.;
        tangOf(gl, tang[v]);
      }
      return;
    }
#pragma unroll
    for (int v = 0; v < Q; ++v) {
      const size_t iR = (size_t)v*d.nCells + R;
      double gl[3];
      gl[0] = 0.5*(gradL[v][0] + d.uX[iR]);
      gl[1] = 0.5*(gradL[v][1] + d.uY[iR]);
      gl[2] = (D == 3) ? 0.5*(gradL[v][2] + d.uZ[iR]) : 0.;
      tangOf(gl, tang[v]);
    }
  }
  // base diffusive flux + update coefficient and, when jac, the 2q perturbed diffusive fluxes staged exactly like
  // FaceWork::innerDiffusive does (column iVar of block(R,L) for side L, of block(L,R) for side R)
  CFB_HD void innerDiffusive(const Dev& d, const Cfg& c, int R, const double* sL, const double* sR, bool jac,
                             double* dfluxBase, double& ucD, double* stg, int stride) {
    innerTang(d, R);
    const Recip rd = recipOf(dr);
#pragma unroll
    for (int v = 0; v < Q; ++v) gradOf(divBy(sR[v] - sL[v], rd), tang[v], grad[v]);
    diffFlux(c, grad, dfluxBase);
    ucD = diffUpdateCoeff(c);
    if (!jac) return;
#pragma unroll 1
    for (int it = 0; it < 2*Q; ++it) {
      const int side = (it >= Q) ? 1 : 0;
      const int iVar = it - side*Q;
      const double oL = pick<Q>(sL, iVar), oR = pick<Q>(sR, iVar);
      const double orig = side ? oR : oL;
      const double pv = orig + fdEps(c, iVar, orig);
      const double dv = side ? divBy(pv - oL, rd) : divBy(oR - pv, rd);
      double t[3], gr[3];
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        double r = tang[0][k];
#pragma unroll
        for (int v = 1; v < Q; ++v) if (iVar == v) r = tang[v][k];
        t[k] = r;
      }
      gradOf(dv, t, gr);
      double g2[Q][3];
#pragma unroll
      for (int v = 0; v < Q; ++v)
#pragma unroll
        for (int k = 0; k < 3; ++k) g2[v][k] = (v == iVar) ? gr[k] : grad[v][k];
      double dflux[Q];
      diffFlux(c, g2, dflux);
      double* blk = stg + (size_t)(side*Q*Q + iVar)*stride;
#pragma unroll
      for (int v = 0; v < Q; ++v) blk[(size_t)v*Q*stride] = dflux[v];
    }
  }

  // ---- boundary faces: computeBoundaryGradients (2D.cxx:155-257, 3D.cxx:88-176) ---------------------------------
  // g = boundary face index; sLs / sRs = the (possibly perturbed) left cell state and the ghost state rebuilt from it.
  // Stencil member 0 is the left cell itself; the other members are read from the state array.
  CFB_HD void boundaryGrads(const Dev& d, int g, const double* sLs, const double* sRs) {
    if (stateDiff) {                                  // the same two-point difference with the ghost state
      const Recip rd0 = recipOf(dr);
      const double t0[3] = {0., 0., 0.};
#pragma unroll
      for (int v = 0; v < Q; ++v) gradOf(divBy(sRs[v] - sLs[v], rd0), t0, grad[v]);
      return;
    }
    double l11 = 0.0, l12 = 0.0, l13 = 0.0, l22 = 0.0, l23 = 0.0, l33 = 0.0;
    double lf1[Q], lf2[Q], lf3[Q];
#pragma unroll
    for (int v = 0; v < Q; ++v) { lf1[v] = 0.0; lf2[v] = 0.0; lf3[v] = 0.0; }
    const int b = d.bndStPtr[g], e = d.bndStPtr[g+1];
    for (int in = b; in < e; ++in) {
      const int s = d.bndSt[in];
      double xs[3] = {0., 0., 0.}, sv[Q];
#pragma unroll
      for (int k = 0; k < D; ++k) xs[k] = d.centres[(size_t)k*d.nStates + s];
#pragma unroll
      for (int v = 0; v < Q; ++v) sv[v] = d.U[(size_t)v*d.nStates + s];
      if (in == b && !gg) {                 // Corrected: member 0 is the (possibly perturbed) left cell itself
#pragma unroll
        for (int v = 0; v < Q; ++v) sv[v] = sLs[v];
      }
      double dist = 0.;
#pragma unroll
      for (int k = 0; k < D; ++k) { const double diff = xR[k] - xs[k]; dist += diff*diff; }
      const double weight = (gg && !ggWeights) ? 1.0 : ddiv(1.0, dsqrt(dist));
      const double dx = weight*(xs[0] - xR[0]);
      const double dy = weight*(xs[1] - xR[1]);
      const double dz = (D == 3) ? weight*(xs[2] - xR[2]) : 0.;
      l11 += dx*dx; l12 += dx*dy; l22 += dy*dy;
      if (D == 3) { l13 += dx*dz; l23 += dy*dz; l33 += dz*dz; }
#pragma unroll
      for (int v = 0; v < Q; ++v) {
        const double du = weight*(sv[v] - sRs[v]);
        lf1[v] += dx*du; lf2[v] += dy*du;
        if (D == 3) lf3[v] += dz*du;
      }
    }
    double invDet, a11 = 0., a12 = 0., a13 = 0., a22 = 0., a23 = 0., a33 = 0.;
    if (D == 2) {
      invDet = ddiv(1., l11*l22 - l12*l12);
    } else {
      const double det = (l11*l22*l33 - l11*l23*l23 - l12*l12*l33 + l12*l13*l23 + l13*l12*l23 - l13*l13*l22);
      invDet = ddiv(1., det);
      a11 = l22*l33 - l23*l23; a22 = l11*l33 - l13*l13; a33 = l11*l22 - l12*l12;
      a12 = -(l12*l33 - l13*l23); a13 = l12*l23 - l13*l22; a23 = -(l11*l23 - l13*l12);
    }
    const Recip rd = recipOf(dr);
#pragma unroll
    for (int v = 0; v < Q; ++v) {
      double uR[3];
      if (D == 2) {
        uR[0] = (l22*lf1[v] - l12*lf2[v])*invDet;
        uR[1] = (l11*lf2[v] - l12*lf1[v])*invDet;
        uR[2] = 0.;
      } else {
        uR[0] = (a11*lf1[v] + a12*lf2[v] + a13*lf3[v])*invDet;
        uR[1] = (a12*lf1[v] + a22*lf2[v] + a23*lf3[v])*invDet;
        uR[2] = (a13*lf1[v] + a23*lf2[v] + a33*lf3[v])*invDet;
      }
      double gl[3], t[3];
#pragma unroll
      for (int k = 0; k < 3; ++k) gl[k] = (k < D) ? 0.5*(gradL[v][k] + uR[k]) : 0.;
      tangOf(gl, t);
      gradOf(divBy(sRs[v] - sLs[v], rd), t, grad[v]);
    }
  }
};
