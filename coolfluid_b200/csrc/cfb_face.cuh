// cfb200 — face kernels: reconstruction, Roe + Navier-Stokes face flux, update coefficients and the
// finite-difference block Jacobian, one thread per face (included by cfb_kernels.cuh).
//
// Register discipline: every small array is indexed with compile-time constants only (loops over
// variables are unrolled, the perturbed variable is handled with select chains, the perturbed side is a
// template parameter), so nothing is demoted to local memory; the loop over the perturbed variable itself
// stays rolled to keep the code size of ~1 flux evaluation per side.
#pragma once
// (included from inside namespace cfb in cfb_kernels.cuh)

template <int N> CFB_HD double pick(const double* a, int i) {
  double r = a[0];
#pragma unroll
  for (int k = 1; k < N; ++k) if (i == k) r = a[k];
  return r;
}

// Per-face working set shared by the flux and the Jacobian part.
// "RoeSA" (RoeSAFlux.cxx:66-197): Sanders' H-correction.  etaSA = half the largest jump between the eigenvalues
// (un, .., un + a, un - a) of the reconstructed L (R) state and those of the CELL (or ghost) state across every face of the
// L (R) cell that shares a node with the face (all its faces with "oldStencil"), each taken with that face's own stored
// normal.  Dev::saFace lists those faces per face and side (host-built at set_mesh; entry 0 is the face itself, whose
// "state across" is the other cell of the face: the one entry a finite-difference perturbation of that cell changes,
// because the reference perturbs the State object in place, FVMCC_ComputeRhsJacob.cxx:321-365).  Boundary faces use
// the left cell only (RoeSAFlux.cxx:92-94).  The triples live in local memory (dynamic index).
template <int D>
struct SaWork {
  static constexpr int Q = D + 2;
  static constexpr int W = CFB_SA_WIDTH;
  typedef Phys<D> P;
  double ev[2][W][3];
  int cnt[2];
  static CFB_HD void triple(const Cfg& c, const double* u, const double* nrm, double* t) {
    PData pd;
    P::physicalData(c, u, pd);                       // updateVarSet->computePhysicalData(*rs, _dataRightState)
    const double un = P::vn(pd, nrm);                // Euler{2,3}DVarSet::computeEigenValues
    t[0] = un; t[1] = un + pd.A; t[2] = un - pd.A;
  }
  CFB_HD void load(const Dev& d, const Cfg& c, int f, int L, int R, bool isB) {
    cnt[0] = 0; cnt[1] = 0;
    const int nSides = isB ? 1 : 2;
    for (int side = 0; side < nSides; ++side) {
      const int cell = side ? R : L;
      for (int k = 0; k < W; ++k) {
        const int nf = d.saFace[(size_t)(side*W + k)*d.nFaces + f];
        if (nf < 0) break;
        const int l = d.faceL[nf];
        const int rs = (l == cell) ? d.faceR[nf] : l;
        const double invArea = ddiv(1., d.areas[nf]);
        double nrm[3] = {0., 0., 0.}, u[Q];
#pragma unroll
        for (int i = 0; i < D; ++i) nrm[i] = d.normals[(size_t)i*d.nFaces + nf]*invArea;
#pragma unroll
        for (int v = 0; v < Q; ++v) u[v] = d.U[(size_t)v*d.nStates + rs];
        triple(c, u, nrm, ev[side][k]);
        cnt[side] = k + 1;
      }
    }
  }
  // the state across the face itself as seen from the cell of `side` (a perturbed cell / ghost state)
  CFB_HD void setAcross(const Cfg& c, int side, const double* u, const double* n) { triple(c, u, n, ev[side][0]); }
  CFB_HD double eta(double unL, double aL, double unR, double aR) const {
    double m = 0.0;
    for (int side = 0; side < 2; ++side) {
      const double un = side ? unR : unL, a = side ? aR : aL;
      const double e1 = un + a, e2 = un - a;
      for (int k = 0; k < cnt[side]; ++k) {
        m = dmax(m, fabs(ev[side][k][0] - un));
        m = dmax(m, fabs(ev[side][k][1] - e1));
        m = dmax(m, fabs(ev[side][k][2] - e2));
      }
    }
    return m*0.5;
  }
};

template <int D>
struct FaceWork {
  static constexpr int Q = D + 2;
  static constexpr int NMAX = (D == 2) ? 2 : 4;
  typedef Phys<D> P;
  int f, L, R, nbN; bool isB;
  double n[3], area;
  double gL[Q], gR[Q];      // limiter * gradientFactor * grad . (xq - x): the frozen reconstruction increments
  // diamond control volume (DiamondVolume{2,3}DDerivative)
  double cvn[8][3];         // 2D: 4 edge normals; 3D quad: 2x4 pyramid normals; 3D tri: 6 tetra normals
  double volume, mu, lambda;
  double val[Q][6];         // control-volume states per variable: face nodes in [0..3], L cell in [4], R state in [5]
  double avState[Q];
  double grad[Q][3];

  // a8: LeastSquareP1PolyRec2D.cxx:158-217, 3D.cxx:204-268.  Boundary faces use the INNER cell's gradient
  // and limiter for the ghost side (reference behaviour).  Also loads the unit normal (FVMCC_ComputeRHS.cxx:518-534).
  CFB_HD void load(const Dev& d, const Cfg& c, int face) {
    f = face; L = d.faceL[f]; R = d.faceR[f]; isB = (R >= d.nCells);
    area = d.areas[f];
    double xq[3] = {0., 0., 0.}, xL[3] = {0., 0., 0.}, xR[3] = {0., 0., 0.}, nr[3] = {0., 0., 0.};
#pragma unroll
    for (int k = 0; k < D; ++k) {
      nr[k] = d.normals[(size_t)k*d.nFaces + f];
      xL[k] = d.centres[(size_t)k*d.nStates + L];
      xR[k] = d.centres[(size_t)k*d.nStates + R];
    }
    if (c.polyRec == CFB200_PR_CONSTANT) {
#pragma unroll
      for (int v = 0; v < Q; ++v) { gL[v] = 0.; gR[v] = 0.; }
    } else {
#pragma unroll
      for (int k = 0; k < D; ++k) xq[k] = d.faceCentres[(size_t)k*d.nFaces + f];
      const int gradR = isB ? L : R;
      const int32_t* zg = isB ? d.bcs[d.faceBC[f]].zeroGrad : nullptr;
      const bool stop = (c.stopLimiting != 0);
      // every gather of the face is issued before the first use (one memory round trip instead of one per variable:
      // with ~8 resident warps per SM the latency of a dependent chain of gathers is not hidden by other warps)
      double gxL[Q], gyL[Q], gzL[Q], gxR[Q], gyR[Q], gzR[Q], lmL[Q], lmR[Q];
#pragma unroll
      for (int v = 0; v < Q; ++v) {
        const size_t iL = (size_t)v*d.nCells + L, iR = (size_t)v*d.nCells + gradR;
        gxL[v] = d.uX[iL]; gyL[v] = d.uY[iL]; gxR[v] = d.uX[iR]; gyR[v] = d.uY[iR];
        if (D == 3) { gzL[v] = d.uZ[iL]; gzR[v] = d.uZ[iR]; } else { gzL[v] = 0.; gzR[v] = 0.; }
        lmL[v] = stop ? 1. : d.lim[iL];
        lmR[v] = stop ? 1. : d.lim[iR];
      }
#pragma unroll
      for (int v = 0; v < Q; ++v) {
        const bool z = zg && zg[v];
        double gcS = 0.0, gcN = 0.0;
        if (!z) {
          if (D == 2) {
            gcS = c.gradientFactor*(gxL[v]*(xq[0] - xL[0]) + gyL[v]*(xq[1] - xL[1]));
            gcN = c.gradientFactor*(gxR[v]*(xq[0] - xR[0]) + gyR[v]*(xq[1] - xR[1]));
          } else {
            gcS = c.gradientFactor*(gxL[v]*(xq[0] - xL[0]) + gyL[v]*(xq[1] - xL[1]) + gzL[v]*(xq[2] - xL[2]));
            gcN = c.gradientFactor*(gxR[v]*(xq[0] - xR[0]) + gyR[v]*(xq[1] - xR[1]) + gzR[v]*(xq[2] - xR[2]));
          }
        }
        gL[v] = lmL[v]*gcS;
        gR[v] = lmR[v]*gcN;
      }
    }
    const double invArea = ddiv(1., area);
    n[2] = 0.;
#pragma unroll
    for (int k = 0; k < D; ++k) n[k] = nr[k]*invArea;
  }

  CFB_HD void reconstruct(const Cfg& c, const double* sL, const double* sR, double* vL, double* vR) const {
    if (c.polyRec == CFB200_PR_CONSTANT) {
#pragma unroll
      for (int v = 0; v < Q; ++v) { vL[v] = sL[v]; vR[v] = sR[v]; }
    } else {
#pragma unroll
      for (int v = 0; v < Q; ++v) { vL[v] = sL[v] + gL[v]; vR[v] = sR[v] + gR[v]; }
    }
  }

  static CFB_HD void tetN(const double* a, const double* b, const double* c2, const double* dd, double fct, double* out) {
    double v1[3], v2[3], v3[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) { v1[k] = a[k] - b[k]; v2[k] = c2[k] - dd[k]; }
    cross3(v1, v2, v3);
#pragma unroll
    for (int k = 0; k < 3; ++k) out[k] = fct*v3[k];
  }

  // a15: control volume of the face, DiamondVolume2DDerivative.cxx:91-137, DiamondVolume3DDerivative.cxx:336-411,
  // and the nodal part of the control-volume states.
  // Raw loads of the diamond (face nodes: coordinates + nodal states; the two cell centres).  Separate from the
  // arithmetic so that the interior-face kernel can issue them together with the other gathers of its prologue.
  struct DiamondRaw { double xL[3], xR[3], xn[NMAX][3]; };
  CFB_HD void fetchDiamond(const Dev& d, DiamondRaw& w) {
#pragma unroll
    for (int k = 0; k < 3; ++k) { w.xL[k] = 0.; w.xR[k] = 0.; }
#pragma unroll
    for (int k = 0; k < D; ++k) { w.xL[k] = d.centres[(size_t)k*d.nStates + L]; w.xR[k] = d.centres[(size_t)k*d.nStates + R]; }
    const int b = d.faceNodePtr[f];
    nbN = d.faceNodePtr[f+1] - b;
    int node[NMAX];
#pragma unroll
    for (int j = 0; j < NMAX; ++j) node[j] = d.faceNodes[b + ((j < nbN) ? j : nbN - 1)];
#pragma unroll
    for (int j = 0; j < NMAX; ++j) {
      w.xn[j][2] = 0.;
#pragma unroll
      for (int k = 0; k < D; ++k) w.xn[j][k] = d.nodes[(size_t)k*d.nNodes + node[j]];
#pragma unroll
      for (int v = 0; v < Q; ++v) val[v][j] = d.nstates[(size_t)v*d.nNodes + node[j]];
    }
  }
  CFB_HD void loadDiamond(const Dev& d) { DiamondRaw w; fetchDiamond(d, w); buildDiamond(d, w); }
  CFB_HD void buildDiamond(const Dev& d, const DiamondRaw& w) {
    const double* xL = w.xL; const double* xR = w.xR;
    const double (*xn)[3] = w.xn;
    if constexpr (D == 2) {
      const double* node0 = xn[0]; const double* node1 = xL; const double* node2 = xn[1]; const double* node3 = xR;
      cvn[0][0] = -node0[1] + node1[1]; cvn[0][1] = -node1[0] + node0[0];
      cvn[1][0] = -node1[1] + node2[1]; cvn[1][1] = -node2[0] + node1[0];
      cvn[2][0] = -node2[1] + node3[1]; cvn[2][1] = -node3[0] + node2[0];
      cvn[3][0] = -node3[1] + node0[1]; cvn[3][1] = -node0[0] + node3[0];
      const double det = (node2[0] - node0[0])*(node3[1] - node1[1]) - (node3[0] - node1[0])*(node2[1] - node0[1]);
      if (det < 0.0) {
#pragma unroll
        for (int k = 0; k < 4; ++k) { cvn[k][0] *= -1.; cvn[k][1] *= -1.; }
      }
      volume = 0.5*fabs(det);
#pragma unroll
      for (int v = 0; v < Q; ++v) avState[v] = 0.5*(val[v][0] + val[v][1]);
    } else if constexpr (D == 3) { if (nbN == 3) {
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
#pragma unroll
      for (int v = 0; v < Q; ++v) avState[v] = (1./3.)*(val[v][0] + val[v][1] + val[v][2]);
    } else {
      // compute2PyramidNormals, DiamondVolume3DDerivative.cxx:336-383
      double factor = (!d.cvSign[f]) ? 1.0 : -1.0;
      triNormal(xn[0], xn[1], xL, factor, cvn[0]); triNormal(xn[1], xn[2], xL, factor, cvn[1]);
      triNormal(xn[2], xn[3], xL, factor, cvn[2]); triNormal(xn[0], xL, xn[3], factor, cvn[3]);
      factor *= -1.;
      triNormal(xn[0], xn[1], xR, factor, cvn[4]); triNormal(xn[1], xn[2], xR, factor, cvn[5]);
      triNormal(xn[2], xn[3], xR, factor, cvn[6]); triNormal(xn[0], xR, xn[3], factor, cvn[7]);
      volume = d.cvol[f];
#pragma unroll
      for (int v = 0; v < Q; ++v) avState[v] = 0.25*(val[v][0] + val[v][1] + val[v][2] + val[v][3]);
    } }
  }

  // Green-Gauss gradient of one variable over the diamond from its 6 control-volume values a[] (nodes 0..3, L, R):
  // DiamondVolume2DDerivative.cxx:139-176, DiamondVolume3DDerivative.cxx:190-298
  CFB_HD void gradRow(const double* a, double* g) const {
    if constexpr (D == 2) {
      const double invVolume = ddiv(0.5, volume);
      const double v01 = a[0] + a[4], v12 = a[4] + a[1], v23 = a[1] + a[5], v30 = a[5] + a[0];
#pragma unroll
      for (int ix = 0; ix < 2; ++ix)
        g[ix] = invVolume*(cvn[0][ix]*v01 + cvn[1][ix]*v12 + cvn[2][ix]*v23 + cvn[3][ix]*v30);
      g[2] = 0.;
    } else if constexpr (D == 3) { if (nbN == 3) {
      const double invVolume = ddiv(1./3., volume);
      const double v031 = a[0] + a[4] + a[1], v132 = a[1] + a[4] + a[2], v023 = a[0] + a[2] + a[4];
      const double v041 = a[0] + a[5] + a[1], v142 = a[1] + a[5] + a[2], v024 = a[0] + a[2] + a[5];
#pragma unroll
      for (int ix = 0; ix < 3; ++ix)
        g[ix] = invVolume*(cvn[0][ix]*v031 + cvn[1][ix]*v132 + cvn[2][ix]*v023 +
                           cvn[3][ix]*v041 + cvn[4][ix]*v142 + cvn[5][ix]*v024);
    } else {
      const double invVolume = ddiv(1./3., volume);
      const double v014 = a[0] + a[1] + a[4], v124 = a[1] + a[2] + a[4], v234 = a[2] + a[3] + a[4], v043 = a[0] + a[4] + a[3];
      const double v015 = a[0] + a[1] + a[5], v125 = a[1] + a[2] + a[5], v235 = a[2] + a[3] + a[5], v053 = a[0] + a[5] + a[3];
#pragma unroll
      for (int ix = 0; ix < 3; ++ix)
        g[ix] = invVolume*(cvn[0][ix]*v014 + cvn[1][ix]*v124 + cvn[2][ix]*v234 + cvn[3][ix]*v043 +
                           cvn[4][ix]*v015 + cvn[5][ix]*v125 + cvn[6][ix]*v235 + cvn[7][ix]*v053);
    } }
  }
  CFB_HD void setCellStates(const double* sL, const double* sR) {
#pragma unroll
    for (int v = 0; v < Q; ++v) { val[v][4] = sL[v]; val[v][5] = sR[v]; }
  }
  CFB_HD void allGrads() {
#pragma unroll
    for (int v = 0; v < Q; ++v) gradRow(val[v], grad[v]);
  }
  // a14: NSFlux.ci:59-156.  mu/lambda depend on the NODAL average state only, which the perturbation loop never
  // touches (and which FreezeDiffCoeff would freeze anyway), so they are evaluated once per face.
  CFB_HD void transport(const Cfg& c) {
    mu = P::dynViscosity(c, avState)*c.artDiffCoeff;
    lambda = P::thermConductivity(c, mu);
  }
  CFB_HD void diffFlux(const Cfg& c, const double (*g)[3], double* dflux) const {
    P::nsFlux(c, avState, g, n, mu, lambda, dflux);
#pragma unroll
    for (int v = 0; v < Q; ++v) dflux[v] *= area;
  }

  // Base (unperturbed) evaluation: a11/a14/a17.  Returns the face-integrated flux (convective - diffusive) and the
  // update coefficients (RoeFlux.cxx:165-185, NSFlux.ci:127-152).
  CFB_HD void baseFlux(const Dev& d, const Cfg& c, const double* sL, const double* sR, const typename P::Side& a,
                       const typename P::Side& b, bool hasDiff, double* flux, double& ucL, double& ucR, double& ucD,
                       double etaSA = 0.) {
    if (!P::numFlux(c, a, b, n, flux, etaSA)) {
      *d.status = CFB200_ERR_BAD_VALUE;
#pragma unroll
      for (int v = 0; v < Q; ++v) flux[v] = 0.;
    }
    const double faceArea = area/1;                     // one quadrature point per face
    ucL = dmax(P::maxEigenValue(a.pd, n), 0.)*faceArea;
    double tn[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) tn[k] = -1.0*n[k];
    ucR = dmax(P::maxEigenValue(b.pd, tn), 0.)*faceArea;
#pragma unroll
    for (int v = 0; v < Q; ++v) flux[v] *= area;        // FVMCC_FluxSplitter.hh:79-83
    ucD = 0.;
    if (hasDiff) {
      loadDiamond(d);          // only now: the diamond working set must not be live across the Roe evaluation
      setCellStates(sL, sR);
      allGrads();
      transport(c);
      double dflux[Q];
      diffFlux(c, grad, dflux);
#pragma unroll
      for (int v = 0; v < Q; ++v) flux[v] -= dflux[v];  // FVMCC_ComputeRHS.cxx:260-276
      const double nu = ddiv(mu, P::density(c, avState));
      ucD = ddiv(nu*area*area, volume);
    }
  }

  // ---- interior faces: ONE rolled loop body serves the base evaluation and all 2q perturbations (code size) -------
  // Diffusive part first: base diffusive flux + ucD, then (JAC) the 2q perturbed diffusive fluxes staged into the
  // two output blocks (column iVar of blkRL for side L, of blkLR for side R).
  CFB_HD void innerDiffusive(const Dev& d, const Cfg& c, const double* sL, const double* sR, bool jac,
                             const DiamondRaw& raw, double* dfluxBase, double& ucD, double* stg, int stride) {
    buildDiamond(d, raw);
    setCellStates(sL, sR);
    allGrads();
    transport(c);
    diffFlux(c, grad, dfluxBase);
    const double nu = ddiv(mu, P::density(c, avState));      // NSFlux.ci:127-152
    ucD = ddiv(nu*area*area, volume);
    if (!jac) return;
#pragma unroll 1
    for (int it = 0; it < 2*Q; ++it) {
      const int side = (it >= Q) ? 1 : 0;
      const int iVar = it - side*Q;
      double orig = sL[0];
#pragma unroll
      for (int v = 0; v < Q; ++v) { if (side == 0 && iVar == v) orig = sL[v]; if (side == 1 && iVar == v) orig = sR[v]; }
      const double pv = orig + fdEps(c, iVar, orig);
      double a[6], gr[3];
#pragma unroll
      for (int j = 0; j < 6; ++j) {
        double r = val[0][j];
#pragma unroll
        for (int v = 1; v < Q; ++v) if (iVar == v) r = val[v][j];
        a[j] = r;
      }
      if (side == 0) a[4] = pv; else a[5] = pv;  // the control volume holds the CELL state, not the reconstructed one
      gradRow(a, gr);
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

  // Convective part: it = -1 is the base evaluation (-> flux, faceFlux), it = 0..2q-1 perturbs variable it%q of side
  // it/q (FVMCC_ComputeRhsJacob.cxx:126-236,321-365; frozen gradient/limiter, LeastSquareP1PolyRec2D.cxx:221-262):
  //   pf = F_c*area - staged F_d ;  column <- sgn * resFactor * (pf - flux)/eps  (row-major q x q, BlockAccumulator.hh:289-296)
  // FS = 0: Roe (fixed side parked as s | z | f);  FS = 1: AUSM family (fixed side parked as rho p H a v[3]);
  // FS = 2: Lax-Friedrichs / HLL (fixed side parked as f | s | un a);  FS = 3: Roe with the entropy fix (s | z | f | un a)
  template <int FS>
  CFB_HD void innerConvective(const Dev& d, const Cfg& c, const double* sL, const double* sR, const double* vL,
                              const double* vR, bool hasDiff, const double* dfluxBase, bool jac,
                              double* stg, int stride) const {
    const double rf = c.resFactor;
    // loop-carried state parked in the staging area (shared memory) instead of registers: the unperturbed other side
    // (s z f) and the base flux; without a staging area (RHS-only kernel: one evaluation) they live in locals
    constexpr int NFX = (FS == 3 || FS == 4) ? 3*Q + 2 : 3*Q;       // doubles of the parked fixed side
    SaWork<D> sa;                                        // FS == 4 ("RoeSA") only
    if constexpr (FS == 4) sa.load(d, c, f, L, R, false);
    double local[NFX + Q];
    double* fx = jac ? stg + (size_t)2*Q*Q*stride : local;
    const int fs = jac ? stride : 1;
    double* fluxS = fx + (size_t)NFX*fs;
    const int last = jac ? 2*Q : 0;
#pragma unroll 1
    for (int it = -1; it < last; ++it) {
      const int side = (it >= Q) ? 1 : 0;
      const int iVar = it - side*Q;                // -1 for the base evaluation
      if (it == -1 || it == Q) {
        double vF[Q];
#pragma unroll
        for (int v = 0; v < Q; ++v) vF[v] = side ? vL[v] : vR[v];
        if constexpr (FS == 0 || FS == 3 || FS == 4) {
          typename P::SideLite fixed;
          P::sideLite(c, vF, n, fixed);              // the unperturbed other side
#pragma unroll
          for (int v = 0; v < Q; ++v) { fx[(size_t)v*fs] = fixed.s[v]; fx[(size_t)(Q + v)*fs] = fixed.z[v]; fx[(size_t)(2*Q + v)*fs] = fixed.f[v]; }
          if constexpr (FS == 3 || FS == 4) { fx[(size_t)(3*Q)*fs] = fixed.un; fx[(size_t)(3*Q + 1)*fs] = fixed.A; }
          if constexpr (FS == 4) { if (it == Q) sa.setAcross(c, 1, sL, n); }   // the L cell is back to its own value
        } else if constexpr (FS == 2) {
          typename P::CSide fixed;
          P::cside(c, vF, n, fixed);
#pragma unroll
          for (int v = 0; v < Q; ++v) { fx[(size_t)v*fs] = fixed.f[v]; fx[(size_t)(Q + v)*fs] = fixed.s[v]; }
          fx[(size_t)(2*Q)*fs] = fixed.un; fx[(size_t)(2*Q + 1)*fs] = fixed.A;
        } else {
          typename P::ASide fixed;
          P::aside(c, vF, fixed);
          fx[0] = fixed.rho; fx[(size_t)fs] = fixed.p; fx[(size_t)2*fs] = fixed.H; fx[(size_t)3*fs] = fixed.A;
#pragma unroll
          for (int k = 0; k < 3; ++k) fx[(size_t)(4 + k)*fs] = fixed.v[k];
        }
      }
      double orig = 0., gp = 0.;
#pragma unroll
      for (int v = 0; v < Q; ++v) {
        if (side == 0 && iVar == v) { orig = sL[v]; gp = gL[v]; }
        if (side == 1 && iVar == v) { orig = sR[v]; gp = gR[v]; }
      }
      const double eps = (iVar >= 0) ? fdEps(c, iVar, orig) : 1.0;
      const double pv = orig + eps;
      const double pvr = (c.polyRec == CFB200_PR_CONSTANT) ? pv : pv + gp;
      double vp[Q];
#pragma unroll
      for (int v = 0; v < Q; ++v) { const double base = side ? vR[v] : vL[v]; vp[v] = (v == iVar) ? pvr : base; }
      double pf[Q];
      if constexpr (FS == 0 || FS == 3 || FS == 4) {
        typename P::SideLite pert;
        P::sideLite(c, vp, n, pert);
        double etaSA = 0.;
        if constexpr (FS == 4) {
          if (iVar >= 0) {
            // the perturbed CELL state is what the other cell sees across this face
            double sp[Q];
#pragma unroll
            for (int v = 0; v < Q; ++v) { const double base = side ? sR[v] : sL[v]; sp[v] = (v == iVar) ? pv : base; }
            sa.setAcross(c, 1 - side, sp, n);
          }
          const double fun = fx[(size_t)(3*Q)*fs], fa = fx[(size_t)(3*Q + 1)*fs];
          etaSA = side ? sa.eta(fun, fa, pert.un, pert.A) : sa.eta(pert.un, pert.A, fun, fa);
          if (!(fabs(etaSA) > 0.)) *d.status = CFB200_ERR_BAD_VALUE;      // assert(std::abs(etaSA) > 0.), RoeSAFlux.cxx:97
        }
        if (!P::template roeOriented<(FS == 3) ? 1 : ((FS == 4) ? 2 : 0)>(c, pert, fx, fs, side == 0, n, pf, etaSA)) {
          *d.status = CFB200_ERR_BAD_VALUE;
#pragma unroll
          for (int v = 0; v < Q; ++v) pf[v] = 0.;
        }
      } else if constexpr (FS == 2) {
        typename P::CSide pert, l, r;
        P::cside(c, vp, n, pert);
        const bool pl = (side == 0);
#pragma unroll
        for (int v = 0; v < Q; ++v) {
          const double xf = fx[(size_t)v*fs], xs = fx[(size_t)(Q + v)*fs];
          l.f[v] = pl ? pert.f[v] : xf; r.f[v] = pl ? xf : pert.f[v];
          l.s[v] = pl ? pert.s[v] : xs; r.s[v] = pl ? xs : pert.s[v];
        }
        { const double x = fx[(size_t)(2*Q)*fs];     l.un = pl ? pert.un : x; r.un = pl ? x : pert.un; }
        { const double x = fx[(size_t)(2*Q + 1)*fs]; l.A = pl ? pert.A : x;   r.A = pl ? x : pert.A; }
        P::centred(c, l, r, pf);
      } else {
        typename P::ASide pert, l, r;
        P::aside(c, vp, pert);
        const bool pl = (side == 0);
        { const double x = fx[0];              l.rho = pl ? pert.rho : x; r.rho = pl ? x : pert.rho; }
        { const double x = fx[(size_t)fs];     l.p = pl ? pert.p : x;     r.p = pl ? x : pert.p; }
        { const double x = fx[(size_t)2*fs];   l.H = pl ? pert.H : x;     r.H = pl ? x : pert.H; }
        { const double x = fx[(size_t)3*fs];   l.A = pl ? pert.A : x;     r.A = pl ? x : pert.A; }
#pragma unroll
        for (int k = 0; k < 3; ++k) { const double x = fx[(size_t)(4 + k)*fs]; l.v[k] = pl ? pert.v[k] : x; r.v[k] = pl ? x : pert.v[k]; }
        P::ausm(c, l, r, n, pf);
      }
      if (iVar < 0) {
#pragma unroll
        for (int v = 0; v < Q; ++v) {
          pf[v] *= area;                            // FVMCC_FluxSplitter.hh:79-83
          if (hasDiff) pf[v] -= dfluxBase[v];       // FVMCC_ComputeRHS.cxx:260-276
          fluxS[(size_t)v*fs] = pf[v];
          d.faceFlux[(size_t)v*d.nFaces + f] = pf[v];
        }
      } else {
        double* blk = stg + (size_t)(side*Q*Q + iVar)*stride;   // column iVar of block(R,L) (side 0) / block(L,R) (side 1)
        const double invEps = ddiv(1., eps);
#pragma unroll
        for (int v = 0; v < Q; ++v) {
          pf[v] *= area;
          if (hasDiff) pf[v] -= blk[(size_t)v*Q*stride];
          double fd = (pf[v] - fluxS[(size_t)v*fs])*invEps;
          fd *= rf;                                 // what row L receives (upFactor[0])
          blk[(size_t)v*Q*stride] = side ? fd : fd*(-1.0);  // block(R,L) is what row R receives: upFactor[1] = -1 times that
        }
      }
    }
  }

  // ---- Roe FD loop, software-pipelined (FS = 0 | 3 with the Jacobian) --------------------------------------------------
  // One evaluation is a long dependent chain (state -> physical data -> Roe vector -> Roe average: seven divide / square
  // root sequences) followed by a wide part (eigenvector matrices, the O(q^3) product).  With two resident warps per
  // scheduler the chain cannot be hidden by other warps, so the loop body is arranged as
  //     front(it + 1)  ||  back(it)
  // in ONE basic block: the chain of the next perturbation has no dependence on the product of the current one and the
  // instruction scheduler interleaves the two.  Same operations, same order inside each evaluation: bit-identical.
  // What lives across iterations is parked in the thread's shared-memory staging area (entry k at stg[k*stride]):
  //   [0, 2QQ) blocks | side L: s z f (un a) | side R: s z f (un a) | base flux (diffusive base flux before it = -1) | gL | gR
  // both unperturbed sides are parked once, the perturbed one is rebuilt per iteration.
  // NB = output blocks staged at a time: 2 (both, k_staged) or 1 (one per phase, k_phased: the staging area then fits three
  // resident blocks per SM)
  static constexpr int RP_NSIDE(bool efix) { return 3*Q + (efix ? 2 : 0); }
  static constexpr int RP_STAGE(bool efix, int nb = 2) { return nb*Q*Q + 2*RP_NSIDE(efix) + 3*Q; }
  struct RoeHand { PData lin; double dU[Q], sumF[Q], eps, lc; bool ok; };   // front -> back
  // the cell value perturbed at iteration `it` (the base evaluation reads any): loaded two iterations ahead by the callers
  CFB_HD double roeOrig(const Dev& d, int it) const {
    const int side = (it >= Q) ? 1 : 0;
    const int iVar = it - side*Q;
    const int iv = (iVar < 0) ? 0 : iVar;
    return d.U[(size_t)iv*d.nStates + (side ? R : L)];
  }
  template <bool EFIX, int UV, int NB = 2>
  CFB_HD void roeFront(const Dev& d, const Cfg& c, int it, double orig, const double* vL, const double* vR, const double* stg,
                       int stride, RoeHand& h) const {
    constexpr int NS = RP_NSIDE(EFIX), OFF_SIDES = NB*Q*Q, OFF_G = NB*Q*Q + 2*NS + Q;
    const int side = (it >= Q) ? 1 : 0;
    const int iVar = it - side*Q;                  // -1: the base evaluation
    const int iv = (iVar < 0) ? 0 : iVar;
    const double gp = stg[(size_t)(OFF_G + side*Q + iv)*stride];
    const double eps = fdEps(c, iv, orig);
    const double pv = orig + eps;
    const double pvr = (c.polyRec == CFB200_PR_CONSTANT) ? pv : pv + gp;
    double vp[Q];
#pragma unroll
    for (int v = 0; v < Q; ++v) vp[v] = dsel(v == iVar, pvr, dsel(side != 0, vR[v], vL[v]));
    typename P::SideLite p;
    P::template sideLiteT<UV>(c, vp, n, p);
    const double* fx = stg + (size_t)(OFF_SIDES + (1 - side)*NS)*stride;     // the unperturbed other side
    {
      double fz[Q];
#pragma unroll
      for (int k = 0; k < Q; ++k) fz[k] = fx[(size_t)(Q + k)*stride];
      h.ok = P::linearize(c, p.z, fz, h.lin);
    }
#pragma unroll
    for (int k = 0; k < Q; ++k) {
      const double fsk = fx[(size_t)k*stride];                 // dU = sR - sL: the operands are selected, not the result
      h.dU[k] = dsel(side == 0, fsk, p.s[k]) - dsel(side == 0, p.s[k], fsk);
    }
#pragma unroll
    for (int i = 0; i < Q; ++i) h.sumF[i] = p.f[i] + fx[(size_t)(2*Q + i)*stride];
    h.lc = 0.;
    if constexpr (EFIX) h.lc = P::lambdaCorrOf(p.un, p.A, fx[(size_t)(3*Q)*stride], fx[(size_t)(3*Q + 1)*stride]);
    h.eps = eps;
  }
  template <bool EFIX, int NB = 2>
  CFB_HD void roeBack(const Dev& d, const Cfg& c, int it, const RoeHand& h, bool hasDiff, double* stg, int stride) const {
    constexpr int NS = RP_NSIDE(EFIX), OFF_FLUX = NB*Q*Q + 2*NS;
    double Lm[Q*Q];
    typename P::RCoef rc;
    if constexpr (EFIX) P::leftScaled(c, h.lin, n, Lm, rc, c.entropyFixID, h.lc); else P::leftScaled(c, h.lin, n, Lm, rc);
    const double kred = c.diffRedCoeff;
    double pf[Q];
#pragma unroll
    for (int i = 0; i < Q; ++i) {
      double Ri[Q];
      P::rightRow(rc, i, Ri);
      double res = 0.;
#pragma unroll
      for (int k = 0; k < Q; ++k) {
        // m = 0 + R(i,0) L(0,k) + ... in the reference's accumulation; the leading `0 +` is dropped: it only turns a -0
        // product into +0, a difference that survives to m solely when every term is a zero, and then
        // res += (k*m)*dU adds a zero of either sign to a sum that starts from +0: the same res, bit for bit
        double m = Ri[0]*Lm[k];
#pragma unroll
        for (int mm = 1; mm < Q; ++mm) m += Ri[mm]*Lm[mm*Q+k];
        res += (kred*m)*h.dU[k];
      }
      pf[i] = 0.5*(h.sumF[i] - res);
    }
    if (!h.ok) *d.status = CFB200_ERR_BAD_VALUE;
#pragma unroll
    for (int v = 0; v < Q; ++v) pf[v] = dsel(h.ok, pf[v], 0.);
    double* fluxS = stg + (size_t)OFF_FLUX*stride;
    const int side = (it >= Q) ? 1 : 0;
    const int iVar = it - side*Q;
    if (it < 0) {
#pragma unroll
      for (int v = 0; v < Q; ++v) {
        double x = pf[v]*area;                                   // FVMCC_FluxSplitter.hh:79-83
        const double dfl = fluxS[(size_t)v*stride];              // base diffusive flux, parked here by the caller
        x = hasDiff ? x - dfl : x;                               // FVMCC_ComputeRHS.cxx:260-276
        fluxS[(size_t)v*stride] = x;
        d.faceFlux[(size_t)v*d.nFaces + f] = x;
      }
    } else {
      double* blk = stg + (size_t)((NB == 2 ? side*Q*Q : 0) + iVar)*stride;   // column iVar of block(R,L) (side 0) / block(L,R) (side 1)
      const double invEps = ddiv(1., h.eps);
      const double rf = c.resFactor;
#pragma unroll
      for (int v = 0; v < Q; ++v) {
        double x = pf[v]*area;
        const double dfl = blk[(size_t)v*Q*stride];              // perturbed diffusive flux staged by innerDiffusive
        x = hasDiff ? x - dfl : x;
        double fd = (x - fluxS[(size_t)v*stride])*invEps;
        fd *= rf;                                                // what row L receives (upFactor[0])
        blk[(size_t)v*Q*stride] = side ? fd : fd*(-1.0);         // block(R,L) is what row R receives: upFactor[1] = -1 times that
      }
    }
  }
  template <bool EFIX, int UV>
  CFB_HD void innerConvectiveRoePipelined(const Dev& d, const Cfg& c, const double* vL, const double* vR, bool hasDiff,
                                          const double* dfluxBase, double* stg, int stride) const {
    constexpr int NS = RP_NSIDE(EFIX), OFF_SIDES = 2*Q*Q, OFF_FLUX = 2*Q*Q + 2*NS, OFF_G = OFF_FLUX + Q;
#pragma unroll
    for (int v = 0; v < Q; ++v) {
      stg[(size_t)(OFF_FLUX + v)*stride] = dfluxBase[v];
      stg[(size_t)(OFF_G + v)*stride] = gL[v];
      stg[(size_t)(OFF_G + Q + v)*stride] = gR[v];
    }
#pragma unroll 1
    for (int sd = 0; sd < 2; ++sd) {
      double vF[Q];
#pragma unroll
      for (int v = 0; v < Q; ++v) vF[v] = sd ? vR[v] : vL[v];
      typename P::SideLite fixed;
      P::template sideLiteT<UV>(c, vF, n, fixed);
      double* fx = stg + (size_t)(OFF_SIDES + sd*NS)*stride;
#pragma unroll
      for (int v = 0; v < Q; ++v) { fx[(size_t)v*stride] = fixed.s[v]; fx[(size_t)(Q + v)*stride] = fixed.z[v]; fx[(size_t)(2*Q + v)*stride] = fixed.f[v]; }
      if constexpr (EFIX) { fx[(size_t)(3*Q)*stride] = fixed.un; fx[(size_t)(3*Q + 1)*stride] = fixed.A; }
    }
    RoeHand cur;
    roeFront<EFIX, UV>(d, c, -1, roeOrig(d, -1), vL, vR, stg, stride, cur);
    double origNext = roeOrig(d, 0);
#pragma unroll 1
    for (int it = -1; it < 2*Q; ++it) {
      RoeHand nxt;
      const int itn = (it + 1 < 2*Q) ? it + 1 : it;      // the last iteration has no successor: its front half is recomputed and dropped
      const double origAfter = roeOrig(d, (it + 2 < 2*Q) ? it + 2 : 2*Q - 1);   // consumed by the NEXT iteration's front half
      roeFront<EFIX, UV>(d, c, itn, origNext, vL, vR, stg, stride, nxt);
      roeBack<EFIX>(d, c, it, cur, hasDiff, stg, stride);
      cur = nxt;
      origNext = origAfter;
    }
  }

  // ---- the same loop, one side per phase (k_phased): staging = one block | both sides | flux | gL gR ---------------------
  // parks what both phases read: the base diffusive flux slots are written by diffusiveSide(0)
  template <bool EFIX, int UV>
  CFB_HD void phasedPark(const Cfg& c, const double* vL, const double* vR, double* stg, int stride) const {
    constexpr int NS = RP_NSIDE(EFIX), OFF_SIDES = Q*Q, OFF_FLUX = Q*Q + 2*NS, OFF_G = OFF_FLUX + Q;
#pragma unroll
    for (int v = 0; v < Q; ++v) {
      stg[(size_t)(OFF_FLUX + v)*stride] = 0.;
      stg[(size_t)(OFF_G + v)*stride] = gL[v];
      stg[(size_t)(OFF_G + Q + v)*stride] = gR[v];
    }
#pragma unroll 1
    for (int sd = 0; sd < 2; ++sd) {
      double vF[Q];
#pragma unroll
      for (int v = 0; v < Q; ++v) vF[v] = dsel(sd != 0, vR[v], vL[v]);
      typename P::SideLite fixed;
      P::template sideLiteT<UV>(c, vF, n, fixed);
      double* fx = stg + (size_t)(OFF_SIDES + sd*NS)*stride;
#pragma unroll
      for (int v = 0; v < Q; ++v) { fx[(size_t)v*stride] = fixed.s[v]; fx[(size_t)(Q + v)*stride] = fixed.z[v]; fx[(size_t)(2*Q + v)*stride] = fixed.f[v]; }
      if constexpr (EFIX) { fx[(size_t)(3*Q)*stride] = fixed.un; fx[(size_t)(3*Q + 1)*stride] = fixed.A; }
    }
  }
  // diffusive part of one phase: the diamond is rebuilt from the mesh (it cannot stay in registers across the convective
  // loop), phase 0 also forms the base diffusive flux (-> flux slots) and the diffusive update coefficient; then the Q
  // perturbed diffusive fluxes of side `side` are staged as the columns of the block
  CFB_HD void diffusiveSide(const Dev& d, const Cfg& c, int side, double* stg, int stride, int offFlux) {
    {
      double sL[Q], sR[Q];
#pragma unroll
      for (int v = 0; v < Q; ++v) { sL[v] = d.U[(size_t)v*d.nStates + L]; sR[v] = d.U[(size_t)v*d.nStates + R]; }
      loadDiamond(d);
      setCellStates(sL, sR);
    }
    allGrads();
    transport(c);
    if (side == 0) {
      double dfluxBase[Q];
      diffFlux(c, grad, dfluxBase);
#pragma unroll
      for (int v = 0; v < Q; ++v) stg[(size_t)(offFlux + v)*stride] = dfluxBase[v];
      const double nu = ddiv(mu, P::density(c, avState));      // NSFlux.ci:127-152
      d.ucD[f] = ddiv(nu*area*area, volume);
    }
    // The control-volume values of the perturbed variable (face nodes, L cell, R cell) are read back from memory per
    // perturbation (they hit L1 / L2) instead of keeping all 6 x Q of them in registers next to the 24 normals and 15
    // gradients; the loads of the next perturbation are issued before the arithmetic of the current one.
    int node[NMAX];
    {
      const int b0 = d.faceNodePtr[f];
#pragma unroll
      for (int j = 0; j < NMAX; ++j) node[j] = d.faceNodes[b0 + ((j < nbN) ? j : nbN - 1)];
    }
    const int pc = side ? R : L, oc = side ? L : R;
    double an[6];
#pragma unroll
    for (int j = 0; j < NMAX; ++j) an[j] = d.nstates[node[j]];
    if (NMAX < 4) { an[2] = 0.; an[3] = 0.; }
    an[4] = d.U[pc]; an[5] = d.U[oc];
#pragma unroll 1
    for (int iVar = 0; iVar < Q; ++iVar) {
      double a[6];
#pragma unroll
      for (int j = 0; j < 6; ++j) a[j] = an[j];
      {
        const int nv = (iVar + 1 < Q) ? iVar + 1 : iVar;
#pragma unroll
        for (int j = 0; j < NMAX; ++j) an[j] = d.nstates[(size_t)nv*d.nNodes + node[j]];
        an[4] = d.U[(size_t)nv*d.nStates + pc]; an[5] = d.U[(size_t)nv*d.nStates + oc];
      }
      const double orig = a[4], other = a[5];
      const double pv = orig + fdEps(c, iVar, orig);
      a[4] = dsel(side == 0, pv, other);         // the control volume holds the CELL state, not the reconstructed one
      a[5] = dsel(side == 0, other, pv);
      double gr[3];
      gradRow(a, gr);
      double g2[Q][3];
#pragma unroll
      for (int v = 0; v < Q; ++v)
#pragma unroll
        for (int k = 0; k < 3; ++k) g2[v][k] = dsel(v == iVar, gr[k], grad[v][k]);
      double dflux[Q];
      diffFlux(c, g2, dflux);
      double* blk = stg + (size_t)iVar*stride;
#pragma unroll
      for (int v = 0; v < Q; ++v) blk[(size_t)v*Q*stride] = dflux[v];
    }
  }
  // convective part of one phase, software-pipelined like innerConvectiveRoePipelined: phase 0 = base evaluation + side L,
  // phase 1 = side R
  template <bool EFIX, int UV>
  CFB_HD void convectiveSide(const Dev& d, const Cfg& c, int side, bool hasDiff, double* stg, int stride) const {
    // the reconstructed face states are re-formed (cell state + parked increment: the additions of reconstruct()) rather
    // than kept in registers across the diffusive part
    constexpr int OFF_G = Q*Q + 2*RP_NSIDE(EFIX) + Q;
    double vL[Q], vR[Q];
    const bool constant = (c.polyRec == CFB200_PR_CONSTANT);
#pragma unroll
    for (int v = 0; v < Q; ++v) {
      const double sl = d.U[(size_t)v*d.nStates + L], sr = d.U[(size_t)v*d.nStates + R];
      vL[v] = dsel(constant, sl, sl + stg[(size_t)(OFF_G + v)*stride]);
      vR[v] = dsel(constant, sr, sr + stg[(size_t)(OFF_G + Q + v)*stride]);
    }
    const int first = side ? Q : -1, last = side ? 2*Q - 1 : Q - 1;
    RoeHand cur;
    roeFront<EFIX, UV, 1>(d, c, first, roeOrig(d, first), vL, vR, stg, stride, cur);
    double origNext = roeOrig(d, (first < last) ? first + 1 : last);
#pragma unroll 1
    for (int it = first; it <= last; ++it) {
      RoeHand nxt;
      const int itn = (it < last) ? it + 1 : it;       // the last iteration of a phase has no successor: recomputed and dropped
      const double origAfter = roeOrig(d, (it + 2 < last) ? it + 2 : last);
      roeFront<EFIX, UV, 1>(d, c, itn, origNext, vL, vR, stg, stride, nxt);
      roeBack<EFIX, 1>(d, c, it, cur, hasDiff, stg, stride);
      cur = nxt;
      origNext = origAfter;
    }
  }

  // a18 (boundary-kernel helpers, kept per side): FVMCC_ComputeRhsJacob.cxx:126-236,321-365.
  // Done in two passes over the perturbed variables so that the diamond-volume working set (24 normals, 30 control
  // volume values, 15 gradients) and the Roe working set are never live in registers at the same time:
  //   diffPass : column iVar of blk <- perturbed diffusive flux (staging, same thread reads it back)
  //   convPass : perturbed convective flux, pf = F_c*area - staged F_d, then
  //              blk column <- sgn * resFactor * (pf - flux)/eps   (row-major q x q, BlockAccumulator.hh:289-296)
  template <int SIDE>
  CFB_HD void diffPass(const Cfg& c, const double* sP, double* blk) const {
#pragma unroll 1
    for (int iVar = 0; iVar < Q; ++iVar) {
      const double orig = pick<Q>(sP, iVar);
      const double pv = orig + fdEps(c, iVar, orig);
      double a[6], gr[3];
#pragma unroll
      for (int j = 0; j < 6; ++j) {
        double r = val[0][j];
#pragma unroll
        for (int v = 1; v < Q; ++v) if (iVar == v) r = val[v][j];
        a[j] = r;
      }
      a[4 + SIDE] = pv;                        // the control volume holds the CELL state, not the reconstructed one
      gradRow(a, gr);
      double g2[Q][3];
#pragma unroll
      for (int v = 0; v < Q; ++v)
#pragma unroll
        for (int k = 0; k < 3; ++k) g2[v][k] = (v == iVar) ? gr[k] : grad[v][k];
      double dflux[Q];
      diffFlux(c, g2, dflux);
#pragma unroll
      for (int v = 0; v < Q; ++v) blk[v*Q + iVar] = dflux[v];
    }
  }

  template <int SIDE>
  CFB_HD void convPass(const Dev& d, const Cfg& c, const double* sP, const double* vBase,
                       const typename P::Side& fixed, const double* flux, bool hasDiff, bool negate, double* blk) const {
    const double rf = c.resFactor;
#pragma unroll 1
    for (int iVar = 0; iVar < Q; ++iVar) {
      const double orig = pick<Q>(sP, iVar);
      const double eps = fdEps(c, iVar, orig);
      const double pv = orig + eps;
      const double pvr = (c.polyRec == CFB200_PR_CONSTANT) ? pv : pv + pick<Q>(SIDE == 0 ? gL : gR, iVar);
      double vp[Q];
#pragma unroll
      for (int v = 0; v < Q; ++v) vp[v] = (v == iVar) ? pvr : vBase[v];
      typename P::Side pert;
      P::side(c, vp, n, pert);
      double pf[Q];
      const bool ok = (SIDE == 0) ? P::numFlux(c, pert, fixed, n, pf) : P::numFlux(c, fixed, pert, n, pf);
      if (!ok) {
        *d.status = CFB200_ERR_BAD_VALUE;
#pragma unroll
        for (int v = 0; v < Q; ++v) pf[v] = 0.;
      }
      const double invEps = ddiv(1., eps);
#pragma unroll
      for (int v = 0; v < Q; ++v) {
        pf[v] *= area;
        if (hasDiff) pf[v] -= blk[v*Q + iVar];
        double fd = (pf[v] - flux[v])*invEps;
        fd *= rf;                                // what row L receives (upFactor[0])
        blk[v*Q + iVar] = negate ? fd*(-1.0) : fd;   // row R receives upFactor[1] = -1 times that
      }
    }
  }
};

#include "cfb_corrected.cuh"

// ------------------------------------------------------------------------------------------------
// Interior faces: flux + update coefficients into face-owned slots and, when JAC, the two off-diagonal blocks
//   block(R,L) = -resFactor * dF/dU_L        block(L,R) = +resFactor * dF/dU_R
// assembled in the thread's shared-memory staging area (which also parks the perturbed diffusive fluxes between the
// two passes) and then written by the warp, in whole rows, straight into the BSR array (each has exactly one writer).  The diagonal contributions are the negatives of
// these and are gathered by DiagGather.  One-sided updatable faces (partition overlap) use the same values
// (FVMCC_ComputeRhsJacob.cxx:184-236: row idx receives (-1)^idx * resFactor * dF/dU); the blocks that land in
// non-updatable rows are staging for DiagGather and are zeroed afterwards by ZeroRows.
// UV >= 0 (Roe families with the Jacobian): the update-variable set as a compile-time constant and the software-pipelined
// finite-difference loop (FaceWork::innerConvectiveRoePipelined).
template <int D, bool JAC, int FS = 0, int DV = 0, int UV = -1>
struct InnerFaceK {   // idx over interior faces; FS = 0 Roe, 1 AUSM+ / AUSM+up, 2 LaxFried / HLL; DV = 0 DiamondVolume, 1 Corrected
  static constexpr int Q = D + 2;
  static constexpr int BLK = Q*Q;
  static constexpr int NOUT = 2*Q*Q;    // per-thread staging (shared memory, entry k at stg[k*stride]): block(R,L) | block(L,R)
  static constexpr bool PIPE = JAC && (FS == 0 || FS == 3) && UV >= 0;
  // RHS-only launches (one evaluation per face, k_run): three resident 128-thread blocks per SM (168 registers), as for
  // the analytic kernel (6.10 ms at 220^3 against 6.3 with two blocks and 6.99 with four); the Jacobian variants go
  // through launchStaged, which does not read this
  static constexpr int MINB = 3;
  // ... | fixed side s z f (un a) | base flux (loop-carried state of the FD loop); PIPE: FaceWork::RP_STAGE
  static constexpr int STAGE = PIPE ? FaceWork<D>::RP_STAGE(FS == 3) : NOUT + 4*Q + ((FS == 3 || FS == 4) ? 2 : 0);
  Dev d; Cfg c;
  CFB_HD void operator()(int f) const { run(f, nullptr, 0); }
  CFB_HD double* out() const { return d.bsrVal; }
  CFB_HD void slots(int f, int& s0, int& s1) const { s0 = d.slotRL[f]; s1 = d.slotLR[f]; }
  CFB_HD void run(int f, double* stg, int stride) const {
    typedef Phys<D> P;
    const bool hasDiff = (c.diffusive == CFB200_DIFF_NS);
    const int L = d.faceL[f], R = d.faceR[f];
    if (!(d.updatable[L] || d.updatable[R])) {             // FVMCC_ComputeRHS.cxx:212-213
#pragma unroll
      for (int v = 0; v < Q; ++v) d.faceFlux[(size_t)v*d.nFaces + f] = 0.;
      d.ucL[f] = 0.; d.ucR[f] = 0.; d.ucD[f] = 0.;
      if (JAC) for (int k = 0; k < NOUT; ++k) stg[(size_t)k*stride] = 0.;
      return;
    }
    // prologue: every gather of the face (states, diamond nodes, gradients, limiters) is issued before the first
    // arithmetic that depends on one of them
    FaceWork<D> w;
    w.f = f; w.L = L; w.R = R;
    double sL[Q], sR[Q], vL[Q], vR[Q];
#pragma unroll
    for (int v = 0; v < Q; ++v) { sL[v] = d.U[(size_t)v*d.nStates + L]; sR[v] = d.U[(size_t)v*d.nStates + R]; }
    typename FaceWork<D>::DiamondRaw raw;
    if (DV == 0 && hasDiff) w.fetchDiamond(d, raw);        // consumed by innerDiffusive
    w.load(d, c, f);
    w.reconstruct(c, sL, sR, vL, vR);
    {
      // update coefficients, RoeFlux.cxx:165-185 (one quadrature point per face)
      PData pL, pR;
      P::physicalData(c, vL, pL);
      P::physicalData(c, vR, pR);
      const double faceArea = w.area/1;
      d.ucL[f] = dmax(P::maxEigenValue(pL, w.n), 0.)*faceArea;
      double tn[3];
#pragma unroll
      for (int k = 0; k < 3; ++k) tn[k] = -1.0*w.n[k];
      d.ucR[f] = dmax(P::maxEigenValue(pR, tn), 0.)*faceArea;
    }
    double dfluxBase[Q], ucD = 0.;
#pragma unroll
    for (int v = 0; v < Q; ++v) dfluxBase[v] = 0.;
    if (hasDiff) {
      if constexpr (DV == 1) {
        CorrectedWork<D> cw;
        cw.setup(d, c, f, L, R, false, w.n, w.area);
        cw.innerDiffusive(d, c, R, sL, sR, JAC, dfluxBase, ucD, stg, stride);
      } else {
        w.innerDiffusive(d, c, sL, sR, JAC, raw, dfluxBase, ucD, stg, stride);
      }
    }
    d.ucD[f] = ucD;
    if constexpr (PIPE) w.template innerConvectiveRoePipelined<FS == 3, UV>(d, c, vL, vR, hasDiff, dfluxBase, stg, stride);
    else w.template innerConvective<FS>(d, c, sL, sR, vL, vR, hasDiff, dfluxBase, JAC, stg, stride);
  }
};

// The Roe finite-difference kernel in two phases (k_phased): phase 0 = base evaluation + the perturbations of the left
// state -> block(R,L), written out by the warp; phase 1 = the perturbations of the right state -> block(L,R).  Only one
// block is staged at a time, so the per-thread staging area is 71 doubles instead of 95 and THREE 128-thread blocks are
// resident per SM (168 registers per thread): the FP64 pipe needs about four independent instruction streams per
// scheduler to stay busy (tools/microbench/fp64_lat.cu), two resident warps with the software-pipelined loop give ~2.
// Price: the diamond control volume of the diffusive flux is rebuilt in phase 1 (~5 % more instructions).
template <int D, int FS, int UV>
struct InnerFacePhased {   // idx over interior faces; FS = 0 Roe | 3 Roe with the entropy fix; DiamondVolume derivatives
  static constexpr int Q = D + 2;
  static constexpr int BLK = Q*Q;
  static constexpr bool EFIX = (FS == 3);
  static constexpr int OFF_FLUX = BLK + 2*FaceWork<D>::RP_NSIDE(EFIX);
  static constexpr int SLOTROW = FaceWork<D>::RP_STAGE(EFIX, 1);      // one more row: the destination slot of the phase
  static constexpr int STAGE = SLOTROW + 1;
  Dev d; Cfg c;
  struct State { FaceWork<D> w; };
  CFB_HD double* out() const { return d.bsrVal; }
  CFB_HD int slot(int f, int ph) const { return ph == 0 ? d.slotRL[f] : d.slotLR[f]; }
  // false: the face has no updatable side (FVMCC_ComputeRHS.cxx:212-213): zero flux, zero blocks
  CFB_HD bool begin(int f, State& st, double* stg, int stride) const {
    typedef Phys<D> P;
    const int L = d.faceL[f], R = d.faceR[f];
    if (!(d.updatable[L] || d.updatable[R])) {
#pragma unroll
      for (int v = 0; v < Q; ++v) d.faceFlux[(size_t)v*d.nFaces + f] = 0.;
      d.ucL[f] = 0.; d.ucR[f] = 0.; d.ucD[f] = 0.;
      for (int k = 0; k < BLK; ++k) stg[(size_t)k*stride] = 0.;
      return false;
    }
    FaceWork<D>& w = st.w;
    w.f = f; w.L = L; w.R = R;
    double sL[Q], sR[Q], vL[Q], vR[Q];
#pragma unroll
    for (int v = 0; v < Q; ++v) { sL[v] = d.U[(size_t)v*d.nStates + L]; sR[v] = d.U[(size_t)v*d.nStates + R]; }
    w.load(d, c, f);
    w.reconstruct(c, sL, sR, vL, vR);
    {
      PData pL, pR;
      P::physicalData(c, vL, pL);
      P::physicalData(c, vR, pR);
      const double faceArea = w.area/1;
      d.ucL[f] = dmax(P::maxEigenValue(pL, w.n), 0.)*faceArea;          // RoeFlux.cxx:165-185
      double tn[3];
#pragma unroll
      for (int k = 0; k < 3; ++k) tn[k] = -1.0*w.n[k];
      d.ucR[f] = dmax(P::maxEigenValue(pR, tn), 0.)*faceArea;
    }
    if (c.diffusive != CFB200_DIFF_NS) d.ucD[f] = 0.;
    w.template phasedPark<EFIX, UV>(c, vL, vR, stg, stride);
    return true;
  }
  CFB_HD void phase(int ph, State& st, double* stg, int stride) const {
    const bool hasDiff = (c.diffusive == CFB200_DIFF_NS);
    if (hasDiff) st.w.diffusiveSide(d, c, ph, stg, stride, OFF_FLUX);
    st.w.template convectiveSide<EFIX, UV>(d, c, ph, hasDiff, stg, stride);
  }
};

// a19: "NumJacobAnalytic" (FVMCC_ComputeRhsJacobAnalytic.cxx:68-246) for AUSM+up in conservative variables, no
// diffusive term: flux + update coefficients as InnerFaceK, and the two off-diagonal blocks from the analytic flux
// Jacobians of the face states, integrated over the face (FVMCC_FluxSplitter.hh:88-96):
//   block(R,L) = -(resFactor * area * dF/dU_L)       block(L,R) = resFactor * area * dF/dU_R
// (the diagonal contributions are their negatives, gathered by DiagGather like the finite-difference ones; the same values
// serve the one-sided updatable faces).  Boundary faces keep the finite-difference term (FVMCC_FluxSplitter.cxx:117-124).
template <int D, int DV = 0>
struct AnalyticInnerFaceK {   // idx over interior faces; staged like InnerFaceK (the warp writes whole blocks)
  static constexpr int Q = D + 2;
  static constexpr int BLK = Q*Q;
  static constexpr int NOUT = 2*Q*Q;
  static constexpr int STAGE = NOUT + 2;
  // one straight evaluation per face, no finite-difference loop to amortise the prologue gathers over: latency-bound at
  // 8 warps per SM (profile: 12 % occupancy, 42 % FP64 pipe), so three resident blocks (168 registers) instead of two:
  // 14.7 -> 12.2 ms for configs[2] at 220^3, 8.0 -> 7.8 ms for configs[1] at 4096^2; four (128 registers, 900 B of stack)
  // were measured at 17.6 / 9.2 ms
  static constexpr int STAGED_MINB = 3;
  Dev d; Cfg c;
  CFB_HD double* out() const { return d.bsrVal; }
  CFB_HD void slots(int f, int& s0, int& s1) const { s0 = d.slotRL[f]; s1 = d.slotLR[f]; }
  CFB_HD void run(int f, double* stg, int stride) const {
    typedef Phys<D> P;
    const int L = d.faceL[f], R = d.faceR[f];
    if (!(d.updatable[L] || d.updatable[R])) {
#pragma unroll
      for (int v = 0; v < Q; ++v) d.faceFlux[(size_t)v*d.nFaces + f] = 0.;
      d.ucL[f] = 0.; d.ucR[f] = 0.; d.ucD[f] = 0.;
      for (int k = 0; k < NOUT; ++k) stg[(size_t)k*stride] = 0.;
      return;
    }
    const bool hasDiff = (c.diffusive == CFB200_DIFF_NS);
    FaceWork<D> w;
    w.f = f; w.L = L; w.R = R;
    double sL[Q], sR[Q], vL[Q], vR[Q];
#pragma unroll
    for (int v = 0; v < Q; ++v) { sL[v] = d.U[(size_t)v*d.nStates + L]; sR[v] = d.U[(size_t)v*d.nStates + R]; }
    w.load(d, c, f);
    w.reconstruct(c, sL, sR, vL, vR);
    PData pL, pR;
    P::physicalData(c, vL, pL);
    P::physicalData(c, vR, pR);
    double flux[Q];
    {
      const double faceArea = w.area/1;
      d.ucL[f] = dmax(P::maxEigenValue(pL, w.n), 0.)*faceArea;         // AUSMFlux.ci:410-440
      double tn[3];
#pragma unroll
      for (int k = 0; k < 3; ++k) tn[k] = -1.0*w.n[k];
      d.ucR[f] = dmax(P::maxEigenValue(pR, tn), 0.)*faceArea;
      typename P::ASide a, b;
      P::aside(pL, a); P::aside(pR, b);
      P::ausm(c, a, b, w.n, flux);
#pragma unroll
      for (int v = 0; v < Q; ++v) flux[v] *= w.area;
    }
    // NSFluxAndJacob::computeFlux (NSFluxAndJacob.cxx:54-118): the diffusive flux, its update coefficient and
    // d grad / d state of the two cells (DerivativeComputer::getGradientsJacob)
    double gj[2][3] = {{0., 0., 0.}, {0., 0., 0.}}, avS[Q], mu = 0., lambda = 0., ucD = 0.;
#pragma unroll
    for (int v = 0; v < Q; ++v) avS[v] = 0.;
    if (hasDiff) {
      double dflux[Q];
      if constexpr (DV == 1) {
        CorrectedWork<D> cw;
        cw.setup(d, c, f, L, R, false, w.n, w.area);
        cw.innerDiffusive(d, c, R, sL, sR, false, dflux, ucD, stg, stride);
        const double ovDr = ddiv(1., cw.dr);                    // CorrectedDerivative2D.cxx:461-469
#pragma unroll
        for (int k = 0; k < D; ++k) { gj[0][k] = (-ovDr)*cw.eN[k]; gj[1][k] = ovDr*cw.eN[k]; }
#pragma unroll
        for (int v = 0; v < Q; ++v) avS[v] = cw.avState[v];
        mu = cw.mu; lambda = cw.lambda;
      } else {
        w.loadDiamond(d);
        w.setCellStates(sL, sR);
        w.allGrads();
        w.transport(c);
        w.diffFlux(c, w.grad, dflux);
        const double nu = ddiv(w.mu, P::density(c, w.avState));
        ucD = ddiv(nu*w.area*w.area, w.volume);
        if constexpr (D == 2) {                                 // DiamondVolume2DDerivative.cxx:153-159
          const double iv = ddiv(0.5, w.volume);
#pragma unroll
          for (int k = 0; k < 2; ++k) { gj[0][k] = iv*(w.cvn[0][k] + w.cvn[1][k]); gj[1][k] = iv*(w.cvn[3][k] + w.cvn[2][k]); }
        } else {                                                // DiamondVolume3DDerivative.cxx:607-625
          const double iv = ddiv(1./3., w.volume);
          if (w.nbN == 3) {
#pragma unroll
            for (int k = 0; k < 3; ++k) {
              gj[0][k] = iv*(w.cvn[0][k] + w.cvn[1][k] + w.cvn[2][k]);
              gj[1][k] = iv*(w.cvn[3][k] + w.cvn[4][k] + w.cvn[5][k]);
            }
          } else {
#pragma unroll
            for (int k = 0; k < 3; ++k) {
              gj[0][k] = iv*(w.cvn[0][k] + w.cvn[1][k] + w.cvn[2][k] + w.cvn[3][k]);
              gj[1][k] = iv*(w.cvn[4][k] + w.cvn[5][k] + w.cvn[6][k] + w.cvn[7][k]);
            }
          }
        }
#pragma unroll
        for (int v = 0; v < Q; ++v) avS[v] = w.avState[v];
        mu = w.mu; lambda = w.lambda;
      }
#pragma unroll
      for (int v = 0; v < Q; ++v) flux[v] -= dflux[v];          // FVMCC_ComputeRHS.cxx:260-276
    }
    d.ucD[f] = ucD;
#pragma unroll
    for (int v = 0; v < Q; ++v) d.faceFlux[(size_t)v*d.nFaces + f] = flux[v];
    const double rf = c.resFactor;
    typename P::JSide jl, jr;
    typename P::JPvt kp;
    typename P::JPre kc;
    const bool pvt = (c.updateVar == CFB200_UV_PVT);            // AUSMPlusUpFlux.ci:1238-1273
    if (pvt) P::jprePvt(c, pL, pR, w.n, kp);
    else { P::jside(c, vL, jl); P::jside(c, vR, jr); P::jpre(c, jl, jr, w.n, kc); }   // once for both sides
#pragma unroll 1
    for (int side = 0; side < 2; ++side) {
      double J[Q*Q];
      if (pvt) P::ausmUpJacobPvt(c, kp, pL, pR, w.n, side, J);
      else P::ausmUpJacobCons(c, kc, jl, jr, w.n, side, J);
#pragma unroll
      for (int k = 0; k < Q*Q; ++k) J[k] *= w.area;             // FVMCC_FluxSplitter.hh:88-96
      if (hasDiff) {
        // "note the opposite sign of the diffusive jacob compared to convective jacob", FVMCC_ComputeRhsJacobAnalytic.cxx:83-91
        double DJ[Q*Q];
        P::nsFluxJacobian(c, avS, gj[side], w.n, mu, lambda, DJ);
#pragma unroll
        for (int k = 0; k < Q*Q; ++k) J[k] -= DJ[k]*w.area;
      }
      if (side == 0) {
#pragma unroll
        for (int k = 0; k < Q*Q; ++k) stg[(size_t)k*stride] = (J[k]*rf)*(-1.);
      } else {
#pragma unroll
        for (int k = 0; k < Q*Q; ++k) stg[(size_t)(Q*Q + k)*stride] = J[k]*rf;
      }
    }
  }
};

// Boundary faces: flux as above (the ghost is the R state) and, when JAC, FVMCC_ComputeRhsJacob.cxx:240-317: the inner
// state is perturbed, the ghost state is rebuilt by the boundary condition, both sides are re-extrapolated, and
// resFactor * dF/dU_L goes to the boundary block of the face (added to the diagonal by DiagGather).
template <int D, bool JAC, int DV = 0>
struct BoundaryFaceK {   // idx over boundary faces; DV = 0 DiamondVolume, 1 Corrected
  Dev d; Cfg c;
  CFB_HD void operator()(int g) const {
    constexpr int Q = D + 2;
    typedef Phys<D> P;
    const bool hasDiff = (c.diffusive == CFB200_DIFF_NS);
    const int f = d.nInner + g;
    const int L = d.faceL[f], R = d.faceR[f];
    double* blk = &d.bndBlocks[(size_t)g*Q*Q];
    if (!d.updatable[L]) {
#pragma unroll
      for (int v = 0; v < Q; ++v) d.faceFlux[(size_t)v*d.nFaces + f] = 0.;
      d.ucL[f] = 0.; d.ucR[f] = 0.; d.ucD[f] = 0.;
      if (JAC) for (int k = 0; k < Q*Q; ++k) blk[k] = 0.;
      return;
    }
    FaceWork<D> w;
    w.load(d, c, f);
    double sL[Q], sR[Q], vL[Q], vR[Q], flux[Q];
#pragma unroll
    for (int v = 0; v < Q; ++v) { sL[v] = d.U[(size_t)v*d.nStates + L]; sR[v] = d.U[(size_t)v*d.nStates + R]; }
    w.reconstruct(c, sL, sR, vL, vR);
    const bool roeSA = (c.fluxSplitter == CFB200_FS_ROESA);
    SaWork<D> sa;
    if (roeSA) sa.load(d, c, f, L, R, true);
    {
      typename P::Side a, b;
      P::side(c, vL, w.n, a);
      P::side(c, vR, w.n, b);
      double ucL, ucR, ucD, etaSA = 0.;
      if (roeSA) {
        etaSA = sa.eta(P::vn(a.pd, w.n), a.pd.A, 0., 0.);
        if (!(fabs(etaSA) > 0.)) *d.status = CFB200_ERR_BAD_VALUE;
      }
      w.baseFlux(d, c, sL, sR, a, b, hasDiff && DV == 0, flux, ucL, ucR, ucD, etaSA);
      d.ucL[f] = ucL; d.ucR[f] = ucR; d.ucD[f] = ucD;
    }
    CorrectedWork<D> cw;
    if (DV == 1 && hasDiff) {
      cw.setup(d, c, f, L, R, true, w.n, w.area);
      cw.boundaryGrads(d, g, sL, sR);
      double dflux[Q];
      cw.diffFlux(c, cw.grad, dflux);
#pragma unroll
      for (int v = 0; v < Q; ++v) flux[v] -= dflux[v];    // FVMCC_ComputeRHS.cxx:260-276
      d.ucD[f] = cw.diffUpdateCoeff(c);
    }
#pragma unroll
    for (int v = 0; v < Q; ++v) d.faceFlux[(size_t)v*d.nFaces + f] = flux[v];
    if (!JAC) return;
    double nr[D];
#pragma unroll
    for (int k = 0; k < D; ++k) nr[k] = d.normals[(size_t)k*d.nFaces + f];
    const BCond& bc = d.bcs[d.faceBC[f]];
    const double rf = c.resFactor;
#pragma unroll 1
    for (int iVar = 0; iVar < Q; ++iVar) {
      const double orig = pick<Q>(sL, iVar);
      const double eps = fdEps(c, iVar, orig);
      double sLp[Q], sRp[Q];
#pragma unroll
      for (int v = 0; v < Q; ++v) sLp[v] = (v == iVar) ? orig + eps : sL[v];
      ghostState<D>(d, g, c, bc, nr, sLp, sRp);
      w.reconstruct(c, sLp, sRp, vL, vR);
      typename P::Side a, b;
      P::side(c, vL, w.n, a);
      P::side(c, vR, w.n, b);
      double pf[Q], etaSA = 0.;
      if (roeSA) {
        sa.setAcross(c, 0, sRp, w.n);             // the ghost state rebuilt from the perturbed inner state
        etaSA = sa.eta(P::vn(a.pd, w.n), a.pd.A, 0., 0.);
        if (!(fabs(etaSA) > 0.)) *d.status = CFB200_ERR_BAD_VALUE;
      }
      if (!P::numFlux(c, a, b, w.n, pf, etaSA)) {
        *d.status = CFB200_ERR_BAD_VALUE;
#pragma unroll
        for (int v = 0; v < Q; ++v) pf[v] = 0.;
      }
#pragma unroll
      for (int v = 0; v < Q; ++v) pf[v] *= w.area;
      if (DV == 1 && hasDiff) {
        cw.boundaryGrads(d, g, sLp, sRp);
        double dflux[Q];
        cw.diffFlux(c, cw.grad, dflux);
#pragma unroll
        for (int v = 0; v < Q; ++v) pf[v] -= dflux[v];
      } else if (hasDiff) {
        w.setCellStates(sLp, sRp);
        w.allGrads();
        double dflux[Q];
        w.diffFlux(c, w.grad, dflux);
#pragma unroll
        for (int v = 0; v < Q; ++v) pf[v] -= dflux[v];
      }
      const double invEps = ddiv(1., eps);
#pragma unroll
      for (int v = 0; v < Q; ++v) {
        double fd = (pf[v] - flux[v])*invEps;
        fd *= rf;
        blk[v*Q + iVar] = fd;
      }
    }
  }
};
