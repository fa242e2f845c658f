// cfb200 — face kernels of the centred splitters (Lax-Friedrichs, HLL) for physics policies that provide
//   PH::CSide / PH::cside() / PH::centred() / PH::maxEigenValue()   (MhdPhys: cfb_mhd.cuh)
// One thread per face, no diffusive flux.  Same contract as InnerFaceK / BoundaryFaceK (cfb_face.cuh): flux and update
// coefficients into face-owned slots; with JAC the finite-difference blocks block(R,L) = -resFactor dF/dU_L and
// block(L,R) = +resFactor dF/dU_R straight into the BSR array (one writer each), boundary blocks into bndBlocks.
// (included from inside namespace cfb in cfb_kernels.cuh)
#pragma once

// a8: frozen reconstruction increments limiter * gradientFactor * grad . (xq - x) of both sides and the unit normal
// (LeastSquareP1PolyRec3D.cxx:204-268, FVMCC_ComputeRHS.cxx:518-534); boundary faces use the inner cell's gradient
// and limiter for the ghost side.
template <int D, int Q>
CFB_HD void faceRecon(const Dev& d, const Cfg& c, int f, int L, int R, double* n, double& area, double* gL, double* gR) {
  const bool isB = (R >= d.nCells);
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
#pragma unroll
    for (int v = 0; v < Q; ++v) {
      const size_t iL = (size_t)v*d.nCells + L, iR = (size_t)v*d.nCells + gradR;
      const bool z = zg && zg[v];
      double gcS = 0.0, gcN = 0.0;
      if (!z) {
        if (D == 2) {
          gcS = c.gradientFactor*(d.uX[iL]*(xq[0] - xL[0]) + d.uY[iL]*(xq[1] - xL[1]));
          gcN = c.gradientFactor*(d.uX[iR]*(xq[0] - xR[0]) + d.uY[iR]*(xq[1] - xR[1]));
        } else {
          gcS = c.gradientFactor*(d.uX[iL]*(xq[0] - xL[0]) + d.uY[iL]*(xq[1] - xL[1]) + d.uZ[iL]*(xq[2] - xL[2]));
          gcN = c.gradientFactor*(d.uX[iR]*(xq[0] - xR[0]) + d.uY[iR]*(xq[1] - xR[1]) + d.uZ[iR]*(xq[2] - xR[2]));
        }
      }
      gL[v] = (stop ? 1. : d.lim[iL])*gcS;
      gR[v] = (stop ? 1. : d.lim[iR])*gcN;
    }
  }
  const double invArea = ddiv(1., area);
  n[2] = 0.;
#pragma unroll
  for (int k = 0; k < D; ++k) n[k] = nr[k]*invArea;
}

// B0 at the quadrature point of face f (both extrapolated states sit there: FVMCC_PolyRec.cxx:170-185)
template <class PH>
CFB_HD void faceBackgroundB(const Dev& d, const Cfg& c, int f, double* B0) {
  double xq[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) xq[k] = (c.mhdPotentialBType != 0) ? d.faceCentres[(size_t)k*d.nFaces + f] : 1.0;
  PH::backgroundB(c, xq, B0);
}

template <class PH, bool JAC>
struct CentredInnerFaceK {   // idx over interior faces
  static constexpr int D = PH::DIM, Q = PH::Q;
  // JAC: launched in two phases through k_staged2 (cfb_capi.cu): phase 0 evaluates the face and perturbs the L state ->
  // block(R,L); phase 1 perturbs the R state -> block(L,R).  One q x q block is staged in shared memory per thread and
  // written out by the warp in whole rows (a 9 x 9 block is 648 B: staging both at once would leave one resident block
  // per SM).  What phase 1 needs of phase 0 (normal, the R cell state and its reconstruction increment, the unperturbed
  // left side, the base flux) stays in registers across the write-out (State): no second gather, no second base evaluation.
  static constexpr int BLK = Q*Q;
  static constexpr int STAGE = BLK + 1;
  struct State { double n[3], area, b0[3], sR[Q], gR[Q], flux[Q]; typename PH::CSide a; bool active; };
  Dev d; Cfg c;
  CFB_HD double* out() const { return d.bsrVal; }
  CFB_HD int slot(int f, int phase) const { return phase ? d.slotLR[f] : d.slotRL[f]; }
  CFB_HD void operator()(int f) const { State st; run(f, 0, st, nullptr, 0); }        // RHS only
  // the finite-difference columns of one side: `fixed` is the unperturbed other side, sP / gP / vBase the cell state, the
  // reconstruction increment and the reconstructed state of the perturbed side
  // (FVMCC_ComputeRhsJacob.cxx:126-236,321-365 with the frozen gradient / limiter, LeastSquareP1PolyRec3D.cxx:272-318)
  template <int SIDE>
  CFB_HD void fdColumns(const State& st, const typename PH::CSide& fixed, const double* sP, const double* gP,
                        const double* vBase, bool lin, double* stg, int stride) const {
    const double rf = c.resFactor;
#pragma unroll 1
    for (int iVar = 0; iVar < Q; ++iVar) {
      const double orig = pick<Q>(sP, iVar), gp = pick<Q>(gP, iVar);
      const double eps = fdEps(c, iVar, orig);
      const double pv = orig + eps;
      const double pvr = lin ? pv + gp : pv;
      double vp[Q];
#pragma unroll
      for (int v = 0; v < Q; ++v) vp[v] = (v == iVar) ? pvr : vBase[v];
      typename PH::CSide pert;
      PH::cside(c, vp, st.n, pert, st.b0);
      double pf[Q];
      if (SIDE == 0) PH::centred(c, pert, fixed, pf); else PH::centred(c, fixed, pert, pf);
      const double invEps = ddiv(1., eps);
#pragma unroll
      for (int v = 0; v < Q; ++v) {
        double fd = (pf[v]*st.area - st.flux[v])*invEps;
        fd *= rf;                                                    // what row L receives (upFactor[0])
        stg[(size_t)(v*Q + iVar)*stride] = SIDE ? fd : fd*(-1.0);    // row R receives upFactor[1] = -1 times that
      }
    }
  }
  CFB_HD void run(int f, int phase, State& st, double* stg, int stride) const {
    const bool lin = (c.polyRec != CFB200_PR_CONSTANT);
    if (phase == 1) {
      if (!st.active) { for (int k = 0; k < BLK; ++k) stg[(size_t)k*stride] = 0.; return; }
      double vR[Q];
#pragma unroll
      for (int v = 0; v < Q; ++v) vR[v] = lin ? st.sR[v] + st.gR[v] : st.sR[v];
      fdColumns<1>(st, st.a, st.sR, st.gR, vR, lin, stg, stride);
      return;
    }
    const int L = d.faceL[f], R = d.faceR[f];
    st.active = d.updatable[L] || d.updatable[R];
    if (!st.active) {             // FVMCC_ComputeRHS.cxx:212-213
      for (int v = 0; v < Q; ++v) d.faceFlux[(size_t)v*d.nFaces + f] = 0.;
      d.ucL[f] = 0.; d.ucR[f] = 0.; d.ucD[f] = 0.;
      if (JAC) for (int k = 0; k < BLK; ++k) stg[(size_t)k*stride] = 0.;
      return;
    }
    double gL[Q], sL[Q], vL[Q], vR[Q];
#pragma unroll
    for (int v = 0; v < Q; ++v) { sL[v] = d.U[(size_t)v*d.nStates + L]; st.sR[v] = d.U[(size_t)v*d.nStates + R]; }
    faceRecon<D, Q>(d, c, f, L, R, st.n, st.area, gL, st.gR);
    faceBackgroundB<PH>(d, c, f, st.b0);
#pragma unroll
    for (int v = 0; v < Q; ++v) { vL[v] = lin ? sL[v] + gL[v] : sL[v]; vR[v] = lin ? st.sR[v] + st.gR[v] : st.sR[v]; }
    typename PH::CSide b;
    PH::cside(c, vL, st.n, st.a, st.b0);
    PH::cside(c, vR, st.n, b, st.b0);
    PH::centred(c, st.a, b, st.flux);
#pragma unroll
    for (int v = 0; v < Q; ++v) st.flux[v] *= st.area;
    {
      // update coefficients: LaxFriedFlux.cxx:120-188 / HLLFlux.cxx:135-150 (left: _leftEv.max(); right: getMaxEigenValue
      // with the reversed normal)
      const double faceArea = st.area/1;
      d.ucL[f] = dmax(st.a.evVecMax, 0.)*faceArea;
      double tn[3];
#pragma unroll
      for (int k = 0; k < 3; ++k) tn[k] = -1.0*st.n[k];
      typename PH::Data pR;
      PH::physicalData(c, vR, pR);
      pR.b0[0] = st.b0[0]; pR.b0[1] = st.b0[1]; pR.b0[2] = st.b0[2];
      d.ucR[f] = dmax(PH::maxEigenValue(c, pR, tn), 0.)*faceArea;
      d.ucD[f] = 0.;
#pragma unroll
      for (int v = 0; v < Q; ++v) d.faceFlux[(size_t)v*d.nFaces + f] = st.flux[v];
    }
    if (!JAC) return;
    fdColumns<0>(st, b, sL, gL, vL, lin, stg, stride);
  }
};

template <class PH, bool JAC>
struct CentredBoundaryFaceK {   // idx over boundary faces
  static constexpr int D = PH::DIM, Q = PH::Q;
  Dev d; Cfg c;
  CFB_HD void operator()(int g) const {
    const int f = d.nInner + g;
    const int L = d.faceL[f], R = d.faceR[f];
    double* blk = &d.bndBlocks[(size_t)g*Q*Q];
    if (!d.updatable[L]) {
      for (int v = 0; v < Q; ++v) d.faceFlux[(size_t)v*d.nFaces + f] = 0.;
      d.ucL[f] = 0.; d.ucR[f] = 0.; d.ucD[f] = 0.;
      if (JAC) for (int k = 0; k < Q*Q; ++k) blk[k] = 0.;
      return;
    }
    double n[3], area, gL[Q], gR[Q], sL[Q], sR[Q], vL[Q], vR[Q];
#pragma unroll
    for (int v = 0; v < Q; ++v) { sL[v] = d.U[(size_t)v*d.nStates + L]; sR[v] = d.U[(size_t)v*d.nStates + R]; }
    faceRecon<D, Q>(d, c, f, L, R, n, area, gL, gR);
    double b0[3];
    faceBackgroundB<PH>(d, c, f, b0);
    const bool lin = (c.polyRec != CFB200_PR_CONSTANT);
#pragma unroll
    for (int v = 0; v < Q; ++v) { vL[v] = lin ? sL[v] + gL[v] : sL[v]; vR[v] = lin ? sR[v] + gR[v] : sR[v]; }
    double flux[Q];
    {
      typename PH::CSide a, b;
      PH::cside(c, vL, n, a, b0);
      PH::cside(c, vR, n, b, b0);
      PH::centred(c, a, b, flux);
      d.ucL[f] = dmax(a.evVecMax, 0.)*(area/1);
      d.ucR[f] = 0.; d.ucD[f] = 0.;               // the ghost side has no cell (FVMCC_ComputeRHS.cxx:455-514)
    }
#pragma unroll
    for (int v = 0; v < Q; ++v) { flux[v] *= area; d.faceFlux[(size_t)v*d.nFaces + f] = flux[v]; }
    if (!JAC) return;
    // FVMCC_ComputeRhsJacob.cxx:240-317: perturb the inner state, rebuild the ghost state, re-extrapolate both sides
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
      ghostState<D, PH>(d, g, c, bc, nr, sLp, sRp);
#pragma unroll
      for (int v = 0; v < Q; ++v) { vL[v] = lin ? sLp[v] + gL[v] : sLp[v]; vR[v] = lin ? sRp[v] + gR[v] : sRp[v]; }
      typename PH::CSide a, b;
      PH::cside(c, vL, n, a, b0);
      PH::cside(c, vR, n, b, b0);
      double pf[Q];
      PH::centred(c, a, b, pf);
      const double invEps = ddiv(1., eps);
#pragma unroll
      for (int v = 0; v < Q; ++v) {
        double fd = (pf[v]*area - flux[v])*invEps;
        fd *= rf;
        blk[v*Q + iVar] = fd;
      }
    }
  }
};
