// cfb200 — MHD3DProjection (a13): the ideal-MHD + hyperbolic divergence cleaning model of plugins/MHD in conservative
// variables (rho, rho u, rho v, rho w, Bx, By, Bz, rho E, phi; q = 9) and the face kernels of the centred splitters
// (Lax-Friedrichs, HLL) for it.  Included by cfb_kernels.cuh inside namespace cfb.
//   MHD3DProjectionCons.cxx:540-574   computePhysicalData          :1132-1148 computeStateFromPhysicalData
//   MHD3DProjectionVarSet.cxx:42-78   computeFlux   :398-436 getMaxEigenValue   :499-548 computeEigenValues
//   FiniteVolumeMHD/MirrorMHD3DProjection.cxx:56-96
// Background potential field B0: none or "Dipole" (computeMagneticDipole :479-495), evaluated at the face quadrature point;
// it enters the wave speeds only (computeFlux has no B0 term).
#pragma once

struct MhdData { double rho, p, A, V, B, phi, v[3], b[3], b0[3]; };   // b0: background field at the point of the data

struct MhdPhys {
  static constexpr int DIM = 3;
  static constexpr int Q = 9;
  static constexpr int TID = 7;
  static constexpr bool CENTRED_ONLY = true;    // only the centred splitters (cfb_centred_face.cuh), no diffusive flux
  typedef MhdData Data;

  static CFB_HD void physicalData(const Cfg& c, const double* s, MhdData& d) {
    const double rho = s[0];
    const Recip rr = recipOf(rho);
    const double u = divBy(s[1], rr), v = divBy(s[2], rr), w = divBy(s[3], rr);
    const double Bx = s[4], By = s[5], Bz = s[6];
    const double V2 = u*u + v*v + w*w;
    const double B2 = Bx*Bx + By*By + Bz*Bz;
    d.v[0] = u; d.v[1] = v; d.v[2] = w;
    d.b[0] = Bx; d.b[1] = By; d.b[2] = Bz;
    d.B = dsqrt(B2);
    d.rho = rho;
    d.V = dsqrt(V2);
    d.p = (c.gamma - 1.)*(s[7] - 0.5*(rho*V2 + B2));
    d.phi = s[8];
    d.A = dsqrt(divBy(c.gamma*d.p, rr));
    d.b0[0] = 0.0; d.b0[1] = 0.0; d.b0[2] = 0.0;
  }
  // MHD3DProjectionVarSet::computeMagneticDipole (MHD3DProjectionVarSet.cxx:479-495) at x; zero without potentialBType
  static CFB_HD void backgroundB(const Cfg& c, const double* x, double* B0) {
    B0[0] = 0.0; B0[1] = 0.0; B0[2] = 0.0;
    if (c.mhdPotentialBType == 1) {
      const double mX = c.mhdDipole[0], mY = c.mhdDipole[1], mZ = c.mhdDipole[2];
      const double xCoord = x[0], yCoord = x[1], zCoord = x[2];
      const double r = dsqrt(xCoord*xCoord + yCoord*yCoord + zCoord*zCoord);
      const double rCubeInv = ddiv(1.0, r*r*r);
      const double rSqInv = ddiv(1.0, r*r);
      const double mDotr = mX*xCoord + mY*yCoord + mZ*zCoord;
      const double coeff = 3.0*rSqInv*mDotr;
      B0[0] = rCubeInv*(coeff*xCoord - mX);
      B0[1] = rCubeInv*(coeff*yCoord - mY);
      B0[2] = rCubeInv*(coeff*zCoord - mZ);
    }
  }
  static CFB_HD void stateFromPhysicalData(const Cfg& c, const MhdData& d, double* s) {
    const double rho = d.rho;
    s[0] = rho;
    s[1] = rho*d.v[0]; s[2] = rho*d.v[1]; s[3] = rho*d.v[2];
    s[4] = d.b[0]; s[5] = d.b[1]; s[6] = d.b[2];
    s[7] = ddiv(d.p, c.gamma - 1.0) + 0.5*(rho*d.V*d.V + d.B*d.B);
    s[8] = d.phi;
  }
  static CFB_HD void mirrorGhost(const Cfg& c, const double* nr, const double* inner, double* ghost) {
    double nx = nr[0], ny = nr[1], nz = nr[2];
    const double invFaceLength = ddiv(1., dsqrt(nx*nx + ny*ny + nz*nz));
    nx *= invFaceLength; ny *= invFaceLength; nz *= invFaceLength;
    MhdData di, dg;
    physicalData(c, inner, di);
    const double vn = di.v[0]*nx + di.v[1]*ny + di.v[2]*nz;
    const double bn = di.b[0]*nx + di.b[1]*ny + di.b[2]*nz;
    dg.rho = di.rho;
    dg.v[0] = di.v[0] - 2.0*vn*nx; dg.v[1] = di.v[1] - 2.0*vn*ny; dg.v[2] = di.v[2] - 2.0*vn*nz;
    dg.b[0] = di.b[0] - 2.0*bn*nx; dg.b[1] = di.b[1] - 2.0*bn*ny; dg.b[2] = di.b[2] - 2.0*bn*nz;
    dg.V = di.V; dg.A = di.A; dg.p = di.p; dg.B = di.B; dg.phi = di.phi;
    stateFromPhysicalData(c, dg, ghost);
  }
  // SuperOutletMHD3DProjection.cxx:76-239.  coronal (option coronalBC, the default): r^2 Br, Btheta, r Bphi, r^3 rho,
  // r^2 rho Vr, rho Vtheta, r Vphi and r^3 p continuous across the face, phi mirrored, on the state entries as they are;
  // else the inner state rebuilt from its physical data with phi = refPhi.  Every quotient a/b of the reference is one
  // correctly rounded division here too (divBy with the reciprocal of b shared between the quotients over the same b).
  static CFB_HD void superOutletGhost(const Cfg& c, bool coronal, double refPhi, const double* xi, const double* xg,
                                      const double* inner, double* ghost) {
    if (!coronal) {
      MhdData dg;
      physicalData(c, inner, dg);
      dg.phi = refPhi;
      stateFromPhysicalData(c, dg, ghost);
      return;
    }
    const double rI = dsqrt(xi[0]*xi[0] + xi[1]*xi[1] + xi[2]*xi[2]), rhoI = dsqrt(xi[0]*xi[0] + xi[1]*xi[1]);
    const double rG = dsqrt(xg[0]*xg[0] + xg[1]*xg[1] + xg[2]*xg[2]), rhoG = dsqrt(xg[0]*xg[0] + xg[1]*xg[1]);
    const Recip qrI = recipOf(rI), qrhoI = recipOf(rhoI), qrrI = recipOf(rhoI*rI);
    const Recip qrG = recipOf(rG), qrhoG = recipOf(rhoG), qrrG = recipOf(rhoG*rG);
    // spherical components of a Cartesian vector at the inner point
    const double er[3] = {divBy(xi[0], qrI), divBy(xi[1], qrI), divBy(xi[2], qrI)};
    const double et[3] = {divBy(xi[0]*xi[2], qrrI), divBy(xi[1]*xi[2], qrrI), divBy(rhoI, qrI)};
    const double ep[2] = {divBy(-xi[1], qrhoI), divBy(xi[0], qrhoI)};
    const double* V = inner + 1; const double* Bv = inner + 4;
    const double BrI = er[0]*Bv[0] + er[1]*Bv[1] + er[2]*Bv[2];
    const double BtI = et[0]*Bv[0] + et[1]*Bv[1] - et[2]*Bv[2];
    const double BpI = ep[0]*Bv[0] + ep[1]*Bv[1];
    const double VrI = er[0]*V[0] + er[1]*V[1] + er[2]*V[2];
    const double VtI = et[0]*V[0] + et[1]*V[1] - et[2]*V[2];
    const double VpI = ep[0]*V[0] + ep[1]*V[1];
    const double rI2 = rI*rI, rG2 = rG*rG, rG3 = rG2*rG;
    const Recip qrG3 = recipOf(rG3);
    const double ratio = divBy(rI, qrG);
    const double BrG = divBy(rI2, recipOf(rG2))*BrI, BtG = BtI, BpG = ratio*BpI;
    const double rhoGst = divBy(rI2*rI*inner[0], qrG3);
    const Recip qRhoG = recipOf(rhoGst);
    const double VrG = ddiv(rI2*VrI*inner[0], rG2*rhoGst);
    const double VtG = divBy(inner[0]*VtI, qRhoG);
    const double VpG = ratio*VpI;
    // back to Cartesian components at the ghost point
    const double gr[3] = {divBy(xg[0], qrG), divBy(xg[1], qrG), divBy(xg[2], qrG)};
    const double gt[3] = {divBy(xg[0]*xg[2], qrrG), divBy(xg[1]*xg[2], qrrG), divBy(rhoG, qrG)};
    const double gp[2] = {divBy(xg[1], qrhoG), divBy(xg[0], qrhoG)};
    ghost[0] = rhoGst;
    ghost[1] = gr[0]*VrG + gt[0]*VtG - gp[0]*VpG;
    ghost[2] = gr[1]*VrG + gt[1]*VtG + gp[1]*VpG;
    ghost[3] = gr[2]*VrG - gt[2]*VtG;
    ghost[4] = gr[0]*BrG + gt[0]*BtG - gp[0]*BpG;
    ghost[5] = gr[1]*BrG + gt[1]*BtG + gp[1]*BpG;
    ghost[6] = gr[2]*BrG - gt[2]*BtG;
    ghost[7] = divBy(inner[7]*rI*rI*rI, qrG3);
    ghost[8] = -inner[8];
  }
  static CFB_HD void farFieldGhost(const Cfg&, const double*, const double*, const double* inner, double* ghost) {
    for (int i = 0; i < Q; ++i) ghost[i] = inner[i];         // FarFieldEuler is rejected for MHD at set_mesh
  }
  static CFB_HD void physFlux(const Cfg& c, const MhdData& d, const double* n, double* f) {
    const double nx = n[0], ny = n[1], nz = n[2];
    const double rho = d.rho, u = d.v[0], v = d.v[1], w = d.v[2], Bx = d.b[0], By = d.b[1], Bz = d.b[2];
    const double Vn = u*nx + v*ny + w*nz;
    const double Bn = Bx*nx + By*ny + Bz*nz;
    const double B2 = Bx*Bx + By*By + Bz*Bz;
    const double VdotB = u*Bx + v*By + w*Bz;
    const double P = d.p + 0.5*B2;
    const double E = ddiv(d.p, c.gamma - 1.0) + 0.5*(rho*(u*u + v*v + w*w) + B2);
    const double mn = Vn*rho;
    f[0] = mn;
    f[1] = mn*u - Bn*Bx + P*nx;
    f[2] = mn*v - Bn*By + P*ny;
    f[3] = mn*w - Bn*Bz + P*nz;
    f[4] = (v*Bx - By*u)*ny + (w*Bx - Bz*u)*nz + d.phi*nx;
    f[5] = (u*By - Bx*v)*nx + (w*By - Bz*v)*nz + d.phi*ny;
    f[6] = (u*Bz - Bx*w)*nx + (v*Bz - By*w)*ny + d.phi*nz;
    f[7] = Vn*(E + P) - Bn*VdotB;
    f[8] = (c.refSpeed*c.refSpeed)*Bn;
  }
  // The wave speeds the eigenvalue vector is made of (computeEigenValues): Vn, cf, cs, ca.
  struct Waves { double Vn, cf, cs, ca; };
  static CFB_HD void waves(const Cfg& c, const MhdData& d, const double* n, Waves& w) {
    const double Bx = d.b[0] + d.b0[0], By = d.b[1] + d.b0[1], Bz = d.b[2] + d.b0[2];      // B1 + B0
    const Recip rr = recipOf(d.rho);
    const double invRho = divBy(1., rr);
    const double sqB = Bx*Bx + By*By + Bz*Bz;
    w.Vn = d.v[0]*n[0] + d.v[1]*n[1] + d.v[2]*n[2];
    const double Bn = Bx*n[0] + By*n[1] + Bz*n[2];
    const double sqrbar = dsqrt(d.rho);
    const double astar2 = (c.gamma*d.p + sqB)*invRho;
    const double astarb = dsqrt(astar2*astar2 - 4.0*c.gamma*d.p*Bn*Bn*invRho*invRho);
    w.cf = dsqrt(0.5*(astar2 + astarb));
    w.cs = dsqrt(0.5*(astar2 - astarb));
    w.ca = ddiv(fabs(Bn), sqrbar);
  }
  static CFB_HD void eigenValues(const Cfg& c, const Waves& w, double* ev) {
    ev[0] = w.Vn - w.cf; ev[1] = w.Vn - w.ca; ev[2] = w.Vn - w.cs; ev[3] = w.Vn;
    ev[4] = c.refSpeed; ev[5] = -c.refSpeed;
    ev[6] = w.Vn + w.cs; ev[7] = w.Vn + w.ca; ev[8] = w.Vn + w.cf;
  }
  static CFB_HD double maxEigenValue(const Cfg& c, const MhdData& d, const double* n) {
    Waves w;
    waves(c, d, n, w);
    return dmax(c.refSpeed, (w.Vn + w.cf));
  }
  static CFB_HD void toCons(const Cfg&, const double* s, double* r) {
#pragma unroll
    for (int i = 0; i < Q; ++i) r[i] = s[i];
  }

  // What the centred splitters read of one side: flux, solution state and the three folds over the eigenvalue vector
  // that LaxFriedFlux.cxx:96-102 / HLLFlux.cxx:93-107 form (each starting from 0.0), plus RealVector::max() of it
  // (ArrayT.hh:61-66) for the left update coefficient.
  struct CSide { double f[Q], s[Q], evMax0, evMin0, evAbs0, evVecMax; };
  static CFB_HD void cside(const Cfg& c, const double* u, const double* n, CSide& o, const double* b0) {
    MhdData pd;
    physicalData(c, u, pd);
    pd.b0[0] = b0[0]; pd.b0[1] = b0[1]; pd.b0[2] = b0[2];
    physFlux(c, pd, n, o.f);
#pragma unroll
    for (int i = 0; i < Q; ++i) o.s[i] = u[i];
    Waves w;
    waves(c, pd, n, w);
    // The folds over ev = (Vn-cf, Vn-ca, Vn-cs, Vn, c, -c, Vn+cs, Vn+ca, Vn+cf) (eigenValues()) without the entries that can
    // never decide them.  In floating point cf >= cs >= 0 and ca >= 0 (square roots / a quotient of non-negative numbers,
    // 0.5*(a2 + ab) >= 0.5*(a2 - ab) for ab >= 0) and x + y is monotone in y, so Vn-cs, Vn, Vn+cs lie in [Vn-cf, Vn+cf],
    // Vn-ca <= Vn+cf and Vn+ca >= Vn-cf: those entries are never the strict maximum, minimum or largest magnitude, and
    // max / min / abs do not round, so the values equal the 9-entry loops of LaxFriedFlux.cxx:96-102 / HLLFlux.cxx:93-107
    // (which start from 0.0) and RealVector::max() (which starts from ev[0]).  Only the folds the configured splitter
    // reads are formed (warp-uniform branch); a fold nobody reads at a call site is dead code there.
    const double e0 = w.Vn - w.cf, e1 = w.Vn - w.ca, e7 = w.Vn + w.ca, e8 = w.Vn + w.cf;
    const double cp = c.refSpeed, cm = -c.refSpeed;
    o.evMax0 = 0.0; o.evMin0 = 0.0; o.evAbs0 = 0.0;
    if (c.fluxSplitter == CFB200_FS_CENTRED) {
      // no fold is read by the flux; the left update coefficient is getMaxEigenValue (CentredFlux.cxx:75-77), whose value
      // max(refSpeed, Vn + cf) is what evVecMax carries for it
      o.evVecMax = dmax(c.refSpeed, e8);
      return;
    }
    if (c.fluxSplitter == CFB200_FS_LAXFRIED) {
      o.evAbs0 = dmax(dmax(dmax(dmax(dmax(dmax(0.0, fabs(e0)), fabs(e1)), fabs(cp)), fabs(cm)), fabs(e7)), fabs(e8));
    } else {
      o.evMax0 = dmax(dmax(dmax(dmax(0.0, cp), cm), e7), e8);
      o.evMin0 = dmin(dmin(dmin(dmin(0.0, e0), e1), cp), cm);
    }
    o.evVecMax = dmax(dmax(dmax(dmax(e0, cp), cm), e7), e8);
  }
  // LaxFriedFlux.cxx:79-118:  0.5*(fR + fL - a*k*(sR - sL));  HLLFlux.cxx:73-133
  static CFB_HD void centred(const Cfg& c, const CSide& l, const CSide& r, double* result) {
    if (c.fluxSplitter == CFB200_FS_CENTRED) {    // CentredFlux.cxx:62-64
#pragma unroll
      for (int i = 0; i < Q; ++i) result[i] = 0.5*(r.f[i] + l.f[i]);
      return;
    }
    if (c.fluxSplitter == CFB200_FS_LAXFRIED) {
      const double aDiff = dmax(r.evAbs0, l.evAbs0)*c.diffRedCoeff;
#pragma unroll
      for (int i = 0; i < Q; ++i) result[i] = 0.5*((r.f[i] + l.f[i]) - aDiff*(r.s[i] - l.s[i]));
    } else {
      const double amax = dmax(0.0, dmax(r.evMax0, l.evMax0)), amin = dmin(0.0, dmin(r.evMin0, l.evMin0));
      const Recip rd = recipOf(amax - amin);
      const double c1 = divBy(amax + amin, rd);
      double c2;
      if (!c.hllUseAlpha) c2 = divBy(amax*amin, rd);
      else c2 = divBy(((divBy(dmax(fabs(amin), fabs(amax)), rd)*2.0)*amax)*amin, rd);
#pragma unroll
      for (int i = 0; i < Q; ++i)
        result[i] = 0.5*(((l.f[i] + r.f[i]) - c1*(r.f[i] - l.f[i])) + c2*(r.s[i] - l.s[i]));
    }
  }
};
