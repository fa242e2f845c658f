// ORACLE — TEST INFRASTRUCTURE ONLY (see cfo_types.hpp header).
// Leaf arithmetic of the FVMCC path: perfect-gas Euler/Navier-Stokes variable sets,
// Roe linearisation, Roe flux, Navier-Stokes diffusive flux with the DiamondVolume
// derivative.  Every function follows the cited reference lines operation by
// operation (same association order, no FMA: build with -ffp-contract=off).
#pragma once
#include "cfo_types.hpp"

namespace cfo {

struct PData {            // EulerTerm enum (plugins/NavierStokes/EulerTerm.hh:85-86)
  double rho, p, H, E, A, T, V, v[3];
};

// Correctly-rounded x^1.5 and x^3 (see DESIGN.md "pow"): the reference calls glibc pow(),
// which is correctly rounded in all but ~1e-4 of cases; oracle and product both use
// the correctly rounded value so that they agree bit for bit.
static inline double cr_pow15(double x) {
  if (!(x > 0.0)) return (x == 0.0) ? 0.0 : std::numeric_limits<double>::quiet_NaN();
  const double s0 = std::sqrt(x);
  const __float128 X = x, S0 = s0;
  const __float128 S = S0 + (X - S0*S0)/(2*S0);   // Newton step: |err| ~ 2^-105
  return (double)(X*S);
}
static inline double cr_pow3(double x) {
  const __float128 X = x;
  return (double)(X*X*X);
}

// ---------------------------------------------------------------------------
template <int D>
struct Phys {
  static constexpr int Q = D + 2;
  static constexpr int TID = D + 1;
  typedef PData Data;

  // Euler2DCons.cxx:441-465, Euler3DCons.cxx:538-562, Euler2DPuvt.cxx:78-108, Euler3DPvt.ci:76-116
  static void setCoordinates(PData&, const double*) {}      // the Euler physical data hold no coordinates
  static void computePhysicalData(const Config& c, const double* s, PData& d) {
    if (c.updateVar == UV_CONS) {
      const double rho = s[0];
      const double ovRho = 1./rho;
      double V2;
      if (D == 2) {
        const double u = s[1]*ovRho, v = s[2]*ovRho;
        V2 = u*u + v*v;
        d.v[0] = u; d.v[1] = v; d.v[2] = 0.;
      } else {
        const double u = s[1]*ovRho, v = s[2]*ovRho, w = s[3]*ovRho;
        V2 = u*u + v*v + w*w;
        d.v[0] = u; d.v[1] = v; d.v[2] = w;
      }
      const double gammaMinus1 = c.gamma - 1.;
      d.rho = rho;
      d.p = gammaMinus1*(s[TID] - 0.5*rho*V2);
      const double pOvRho = d.p*ovRho;
      d.E = s[TID]*ovRho;
      d.H = d.E + pOvRho;
      d.A = std::sqrt(c.gamma*pOvRho);
      d.T = pOvRho/c.R;
      d.V = std::sqrt(V2);
    } else {
      const double p = s[0] + c.pOffset;                 // EulerTerm.hh:199-202
      const double T = s[TID];
      const double rho = p/(c.R*T);                       // EulerTerm.hh:208-212
      double V2;
      if (D == 2) { V2 = s[1]*s[1] + s[2]*s[2]; d.v[0] = s[1]; d.v[1] = s[2]; d.v[2] = 0.; }
      else { V2 = s[1]*s[1] + s[2]*s[2] + s[3]*s[3]; d.v[0] = s[1]; d.v[1] = s[2]; d.v[2] = s[3]; }
      const double gammaDivGammaMinus1 = c.gamma/(c.gamma - 1.);
      const double pOvRho = p/rho;
      d.rho = rho;
      d.p = s[0];
      d.H = gammaDivGammaMinus1*pOvRho + 0.5*V2;
      d.E = d.H - pOvRho;
      d.A = std::sqrt(c.gamma*pOvRho);
      d.T = T;
      d.V = std::sqrt(V2);
    }
  }

  // Euler2DCons.cxx / Euler3DCons.cxx:564-574 computeStateFromPhysicalData (Cons only; used by mirror BC)
  static void stateFromPhysicalDataCons(const PData& d, double* s) {
    const double rho = d.rho;
    s[0] = rho;
    for (int i = 0; i < D; ++i) s[1+i] = rho*d.v[i];
    s[TID] = rho*d.H - d.p;
  }

  // MirrorEuler2D.cxx:56-94 (+ 3D analogue): nr is the area-weighted face normal
  static void mirrorGhost(const Config& cfg, const double* nr, const double* inner, double* ghost) {
    double n[3] = {0,0,0};
    double s2 = 0.;
    if (D == 2) s2 = nr[0]*nr[0] + nr[1]*nr[1]; else s2 = nr[0]*nr[0] + nr[1]*nr[1] + nr[2]*nr[2];
    const double invFaceLength = 1./std::sqrt(s2);
    for (int i = 0; i < D; ++i) n[i] = nr[i]*invFaceLength;
    PData di, dg;
    computePhysicalData(cfg, inner, di);
    double vn;
    if (D == 2) vn = di.v[0]*n[0] + di.v[1]*n[1]; else vn = di.v[0]*n[0] + di.v[1]*n[1] + di.v[2]*n[2];
    const double gamma = cfg.gamma;
    const double gammaDivGammaMinus1 = gamma/(gamma - 1.0);
    dg.rho = di.rho;
    dg.v[2] = 0.;
    for (int i = 0; i < D; ++i) dg.v[i] = di.v[i] - 2.0*vn*n[i];
    dg.V = di.V;
    dg.p = di.p;
    dg.H = (gammaDivGammaMinus1*dg.p + 0.5*dg.rho*dg.V*dg.V)/dg.rho;
    dg.A = std::sqrt(gamma*dg.p/dg.rho);
    dg.T = di.T;
    stateFromPhysicalDataCons(dg, ghost);
  }

  // FarFieldEuler2D.cxx:127-300 / FarFieldEuler3D.cxx:62-196: the number of imposed quantities follows the normal
  // Mach number of the inner state; inf = (p, u, v, [w,] T) at infinity (Pressure, Uinf, Vinf, Winf, Temperature).
  // The 2D command leaves EulerTerm::T of the ghost data unset in two of its branches, so only its conservative form
  // is a function of its inputs (Euler2DCons::computeStateFromPhysicalData does not read T); the 3D one sets T always.
  static void superOutletGhost(const Config&, bool, double, const double*, const double*, const double*, double*) {
    throw std::runtime_error("SuperOutletMHD3DProjection is a command of the MHD3DProjection model");
  }
  static void farFieldGhost(const Config& cfg, const double* nr, const double* inf, const double* inner, double* ghost) {
    if (D == 2 && cfg.updateVar != UV_CONS) throw std::runtime_error("FarFieldEuler2D: conservative variables only");
    double n[3] = {0., 0., 0.};
    double s2;
    if (D == 2) s2 = nr[0]*nr[0] + nr[1]*nr[1]; else s2 = nr[0]*nr[0] + nr[1]*nr[1] + nr[2]*nr[2];
    const double invFaceLength = 1./std::sqrt(s2);
    for (int i = 0; i < D; ++i) n[i] = nr[i]*invFaceLength;
    PData di, dg;
    computePhysicalData(cfg, inner, di);
    const double gamma = cfg.gamma;
    const double gammaDivGammaMinus1 = gamma/(gamma - 1.0);
    const double R = cfg.R;
    const double pInf = inf[0], TInf = inf[D+1];
    const double rho = di.rho, pInnerState = di.p;
    double vn;
    if (D == 2) vn = di.v[0]*n[0] + di.v[1]*n[1]; else vn = di.v[0]*n[0] + di.v[1]*n[1] + di.v[2]*n[2];
    const double machInner = vn/di.A;
    auto speed = [&](const PData& x) {
      return (D == 2) ? std::sqrt(x.v[0]*x.v[0] + x.v[1]*x.v[1]) : std::sqrt(x.v[0]*x.v[0] + x.v[1]*x.v[1] + x.v[2]*x.v[2]);
    };
    auto finish = [&](bool setT) {
      if (setT) dg.T = dg.p/(dg.rho*R);
      if (cfg.updateVar == UV_CONS) stateFromPhysicalDataCons(dg, ghost);
      else { ghost[0] = dg.p; for (int i = 0; i < D; ++i) ghost[1+i] = dg.v[i]; ghost[TID] = dg.T; }   // Euler3DPvt.ci:118-128
    };
    dg.v[2] = 0.;
    if (machInner >= 1.0) {                                  // supersonic outlet
      for (int i = 0; i < Q; ++i) ghost[i] = inner[i];
    }
    if (machInner <= -1.0) {                                 // supersonic inlet
      const double rhoInf = pInf/(R*TInf);
      dg.rho = 2.0*rhoInf - rho;
      for (int i = 0; i < D; ++i) dg.v[i] = 2.0*inf[1+i] - di.v[i];
      dg.p = 2.0*pInf - pInnerState;
      dg.A = std::sqrt(gamma*dg.p/dg.rho);
      dg.V = speed(dg);
      dg.H = (gammaDivGammaMinus1*dg.p + 0.5*dg.rho*dg.V*dg.V)/dg.rho;
      finish(D == 3);
    }
    if ((machInner < 1.0) && (machInner >= 0.0)) {           // subsonic outlet
      dg.rho = di.rho;
      for (int i = 0; i < D; ++i) dg.v[i] = di.v[i];
      dg.V = di.V;
      dg.p = 2.0*pInf - pInnerState;
      dg.H = (gammaDivGammaMinus1*dg.p + 0.5*dg.rho*dg.V*dg.V)/dg.rho;
      dg.A = std::sqrt(gamma*dg.p/dg.rho);
      finish(D == 3);
    }
    if ((machInner > -1.0) && (machInner < 0.0)) {           // subsonic inlet
      dg.rho = pInnerState/(R*TInf);
      for (int i = 0; i < D; ++i) dg.v[i] = 2.0*inf[1+i] - di.v[i];
      dg.V = speed(dg);
      dg.p = pInnerState;
      dg.H = (gammaDivGammaMinus1*dg.p + 0.5*dg.rho*dg.V*dg.V)/dg.rho;
      dg.A = std::sqrt(gamma*dg.p/dg.rho);
      dg.T = TInf;
      finish(D == 3);
    }
  }

  // FarFieldEulerChar2D.cxx:36-254 / FarFieldEulerChar3D.cxx:40-200 (no StateInterpolator): far field through the Riemann
  // invariants R+ = vn + 2a/(gamma-1) of the inner state and R- = vn_inf - 2a_inf/(gamma-1).  Supersonic branches as in
  // FarFieldEuler; subsonic: outflow when (R+ + R-)/2 > 0 (inner rho and p, normal velocity corrected), inflow otherwise
  // (rho, p from infinity).  Every branch sets T, so all variable sets are served.  2D takes `<= 0` for the inflow
  // branch, 3D `< 0` (an exactly zero mean then leaves the ghost state as it was: restated here as inflow).
  static void farFieldCharGhost(const Config& cfg, const double* nr, const double* inf, const double* inner, double* ghost) {
    double n[3] = {0., 0., 0.};
    double s2;
    if (D == 2) s2 = nr[0]*nr[0] + nr[1]*nr[1]; else s2 = nr[0]*nr[0] + nr[1]*nr[1] + nr[2]*nr[2];
    const double invFaceLength = 1./std::sqrt(s2);
    for (int i = 0; i < D; ++i) n[i] = nr[i]*invFaceLength;
    PData di, dg;
    computePhysicalData(cfg, inner, di);
    const double gamma = cfg.gamma, gammaMinus1 = gamma - 1.0;
    const double gammaDivGammaMinus1 = gamma/(gamma - 1.0);
    const double R = cfg.R;
    const double pInf = inf[0], TInf = inf[D+1];
    const double rho = di.rho, pInnerState = di.p, aInnerState = di.A;
    double vn, vnInf;
    if (D == 2) { vn = di.v[0]*n[0] + di.v[1]*n[1]; vnInf = inf[1]*n[0] + inf[2]*n[1]; }
    else { vn = di.v[0]*n[0] + di.v[1]*n[1] + di.v[2]*n[2]; vnInf = inf[1]*n[0] + inf[2]*n[1] + inf[3]*n[2]; }
    const double machInner = vn/aInnerState;
    const double rhoInf = pInf/(R*TInf);
    const double aInf = std::sqrt(gamma*pInf/rhoInf);
    const double RPlus_n = vn + (2.*aInnerState)/(gammaMinus1);
    const double RMin_n = vnInf - (2.*aInf)/(gammaMinus1);
    const double RMinRmax = (RPlus_n + RMin_n)*0.5;
    auto finish = [&]() {
      if (D == 2) dg.V = std::sqrt(dg.v[0]*dg.v[0] + dg.v[1]*dg.v[1]);
      else dg.V = std::sqrt(dg.v[0]*dg.v[0] + dg.v[1]*dg.v[1] + dg.v[2]*dg.v[2]);
      dg.H = (gammaDivGammaMinus1*dg.p + 0.5*dg.rho*dg.V*dg.V)/dg.rho;
      dg.A = std::sqrt(gamma*dg.p/dg.rho);
      dg.T = dg.p/(dg.rho*R);
      if (cfg.updateVar == UV_CONS) stateFromPhysicalDataCons(dg, ghost);
      else { ghost[0] = dg.p; for (int i = 0; i < D; ++i) ghost[1+i] = dg.v[i]; ghost[TID] = dg.T; }
    };
    dg.v[2] = 0.;
    if (machInner >= 1.0) for (int i = 0; i < Q; ++i) ghost[i] = inner[i];
    if (machInner <= -1.0) {
      dg.rho = 2.0*rhoInf - rho;
      for (int i = 0; i < D; ++i) dg.v[i] = 2.0*inf[1+i] - di.v[i];
      dg.p = 2.0*pInf - pInnerState;
      finish();
    }
    if (std::fabs(machInner) < 1.0) {
      if (RMinRmax > 0.) {
        const double vnInner = vn;
        dg.rho = di.rho;
        for (int i = 0; i < D; ++i) { const double bound = di.v[i] + (RMinRmax - vnInner)*n[i]; dg.v[i] = 2.0*bound - di.v[i]; }
        dg.p = di.p;
        finish();
      } else {
        dg.rho = pInf/(R*TInf);
        for (int i = 0; i < D; ++i) { const double bound = inf[1+i] + (RMinRmax - vnInf)*n[i]; dg.v[i] = 2.0*bound - di.v[i]; }
        dg.p = pInf;
        finish();
      }
    }
  }

  // Euler2DVarSet.cxx:83-117, Euler3DVarSet.cxx:78-105
  static void convFlux(const PData& d, const double* n, double* f) {
    double un;
    if (D == 2) un = d.v[0]*n[0] + d.v[1]*n[1];
    else        un = d.v[0]*n[0] + d.v[1]*n[1] + d.v[2]*n[2];
    const double rhoVn = d.rho*un;
    f[0] = rhoVn;
    for (int i = 0; i < D; ++i) f[1+i] = d.p*n[i] + d.v[i]*rhoVn;
    f[TID] = rhoVn*d.H;
  }

  // Euler2DVarSet.cxx:39-45, Euler3DVarSet.cxx:41-46
  static double maxEigenValue(const PData& d, const double* n) {
    double un;
    if (D == 2) un = d.v[0]*n[0] + d.v[1]*n[1];
    else        un = d.v[0]*n[0] + d.v[1]*n[1] + d.v[2]*n[2];
    return un + d.A;
  }

  static double maxEigenValue(const Config&, const PData& d, const double* n) { return maxEigenValue(d, n); }
  // RealVector::max() of computeEigenValues (ArrayT.hh:61-66)
  static double evMaxOfVector(const Config&, const PData& d, const double* n) {
    double ev[Q];
    eigenValues(d, n, ev);
    double emax = ev[0];
    for (int i = 1; i < Q; ++i) emax = std::max(emax, ev[i]);
    return emax;
  }

  // update -> solution(Cons): identity or EulerPvtToCons.cxx:49-70
  static void toCons(const Config& c, const double* s, double* r) {
    if (c.updateVar == UV_CONS) { for (int i = 0; i < Q; ++i) r[i] = s[i]; return; }
    const double p = s[0] + c.pOffset;
    const double T = s[TID];
    const double rho = p/(c.R*T);
    r[0] = rho;
    double V2 = 0.;
    for (int i = 0; i < D; ++i) { const double u = s[1+i]; r[1+i] = rho*u; V2 += u*u; }
    r[TID] = p/(c.gamma - 1.) + 0.5*rho*V2;
  }

  // dU/dP of the update -> solution transformation, row-major Q x Q: Euler2DPuvtToConsInPuvt.cxx:42-67,
  // Euler3DPvtToConsInPvt.cxx:42-76; identity for conservative update variables (IdentityVarSetMatrixTransformer)
  static void updateToSolutionMatrix(const Config& c, const double* state, double* M) {
    for (int i = 0; i < Q*Q; ++i) M[i] = 0.;
    if (c.updateVar == UV_CONS) { for (int i = 0; i < Q; ++i) M[i*Q + i] = 1.; return; }
    const double R = c.R;
    const double p = state[0] + c.pOffset;
    const double T = state[TID];
    const double rho = p/(R*T);
    const double rhoOvT = rho/T;
    const double invRT = 1./(R*T);
    double V2;
    if (D == 2) V2 = state[1]*state[1] + state[2]*state[2];
    else V2 = state[1]*state[1] + state[2]*state[2] + state[3]*state[3];
    const double gammaMinus1 = c.gamma - 1.;
    M[0*Q + 0] = invRT;
    M[0*Q + TID] = -rhoOvT;
    for (int i = 0; i < D; ++i) {
      const double u = state[1+i];
      M[(1+i)*Q + 0] = u*invRT;
      M[(1+i)*Q + 1+i] = rho;
      M[(1+i)*Q + TID] = -rhoOvT*u;
      M[TID*Q + 1+i] = rho*u;
    }
    M[TID*Q + 0] = 1./gammaMinus1 + 0.5*V2*invRT;
    M[TID*Q + TID] = -0.5*rhoOvT*V2;
  }

  // Euler2DConsToRoe.cxx:41-52, Euler3DConsToRoe.cxx:41-58
  static void consToRoe(const Config& c, const double* s, double* r) {
    const double gamma = c.gamma;
    const double gammaMinus1 = gamma - 1.;
    if (D == 2) {
      const double sqrho = std::sqrt(s[0]);
      r[0] = sqrho;
      r[1] = s[1]/sqrho;
      r[2] = s[2]/sqrho;
      const double sqrhoU2 = r[1]*r[1] + r[2]*r[2];
      r[3] = (gamma*s[3] - 0.5*gammaMinus1*sqrhoU2)/sqrho;
    } else {
      const double rho = s[0], rhoU = s[1], rhoV = s[2], rhoW = s[3], rhoE = s[4];
      r[0] = std::sqrt(rho);
      r[1] = rhoU/r[0];
      r[2] = rhoV/r[0];
      r[3] = rhoW/r[0];
      r[4] = (gamma*rhoE - 0.5*gammaMinus1*(rhoU*rhoU + rhoV*rhoV + rhoW*rhoW)/rho)/r[0];
    }
  }

  // Euler2DLinearRoe.cxx:44-101, Euler3DLinearRoe.cxx:43-92.  Returns false if a2 < 0
  // (the reference throws BadValueException there).
  static bool linearize(const Config& c, const double* zL, const double* zR, PData& d) {
    double sumZ[Q];
    for (int i = 0; i < Q; ++i) { sumZ[i] = 0.; sumZ[i] += zL[i]; sumZ[i] += zR[i]; }
    const double gamma = c.gamma;
    if (D == 2) {
      const double invNbStates = 1./2.;
      double avZ[Q];
      for (int i = 0; i < Q; ++i) avZ[i] = sumZ[i]*invNbStates;
      const double ovAvZ0 = 1./avZ[0];
      d.rho = avZ[0]*avZ[0];
      d.v[0] = avZ[1]*ovAvZ0;
      d.v[1] = avZ[2]*ovAvZ0;
      d.v[2] = 0.;
      const double speed2 = (d.v[0]*d.v[0] + d.v[1]*d.v[1]);
      d.V = std::sqrt(speed2);
      d.H = avZ[3]*ovAvZ0;
      const double a2 = (gamma-1.)*(d.H - 0.5*speed2);
      if (a2 < 0.0) return false;
      d.A = std::sqrt(a2);
      d.T = a2/(gamma*c.R);
      const double RT = c.R*d.T;
      d.p = d.rho*RT;
      d.E = d.H - RT;
    } else {
      double avZ0 = sumZ[0]/2.;
      d.rho = avZ0*avZ0;
      d.v[0] = sumZ[1]/sumZ[0];
      d.v[1] = sumZ[2]/sumZ[0];
      d.v[2] = sumZ[3]/sumZ[0];
      d.H = sumZ[4]/sumZ[0];
      d.V = std::sqrt(d.v[0]*d.v[0] + d.v[1]*d.v[1] + d.v[2]*d.v[2]);
      const double speed2 = d.V*d.V;
      const double a2 = (gamma - 1.)*(d.H - 0.5*speed2);
      if (a2 < 0.0) return false;
      d.A = std::sqrt(a2);
      d.T = a2/(gamma*c.R);
      d.p = d.rho*c.R*d.T;
      d.E = d.H - d.p/d.rho;
    }
    return true;
  }

  // Euler2DCons.cxx:156-228, Euler3DCons.cxx:150-239 (row-major Q x Q)
  static void eigen(const Config& c, const PData& lin, const double* nrm, double* R, double* L, double* ev) {
    const double gamma = c.gamma;
    const double gammaMinus1 = gamma - 1.;
    if (D == 2) {
      const double nx = nrm[0], ny = nrm[1];
      const double avRho = lin.rho, avU = lin.v[0], avV = lin.v[1], avH = lin.H, avA = lin.A;
      const double ovAvA = 1./avA;
      const double ovAvA2 = ovAvA*ovAvA;
      const double um = avU*nx + avV*ny;
      const double ra = 0.5*avRho*ovAvA;
      const double coeffM2 = 0.5*gammaMinus1*(avU*avU + avV*avV)*ovAvA2;
      const double ovAvRho = 1./avRho;
      const double uDivA = gammaMinus1*avU*ovAvA;
      const double vDivA = gammaMinus1*avV*ovAvA;
      const double ovAvRhoA = ovAvRho*ovAvA;
#define RE(i,j) R[(i)*4+(j)]
#define LE(i,j) L[(i)*4+(j)]
      RE(0,0) = 1.; RE(0,1) = 0.; RE(0,2) = ra; RE(0,3) = ra;
      RE(1,0) = avU; RE(1,1) = avRho*ny; RE(1,2) = ra*(avU + avA*nx); RE(1,3) = ra*(avU - avA*nx);
      RE(2,0) = avV; RE(2,1) = -avRho*nx; RE(2,2) = ra*(avV + avA*ny); RE(2,3) = ra*(avV - avA*ny);
      RE(3,0) = 0.5*(avU*avU + avV*avV); RE(3,1) = avRho*(avU*ny - avV*nx);
      RE(3,2) = ra*(avH + avA*um); RE(3,3) = ra*(avH - avA*um);
      LE(0,0) = 1.- coeffM2; LE(0,1) = uDivA*ovAvA; LE(0,2) = vDivA*ovAvA; LE(0,3) = -gammaMinus1*ovAvA2;
      LE(1,0) = ovAvRho*(avV*nx - avU*ny); LE(1,1) = ovAvRho*ny; LE(1,2) = -ovAvRho*nx; LE(1,3) = 0.0;
      LE(2,0) = avA*ovAvRho*(coeffM2 - um*ovAvA); LE(2,1) = ovAvRho*(nx - uDivA);
      LE(2,2) = ovAvRho*(ny - vDivA); LE(2,3) = gammaMinus1*ovAvRhoA;
      LE(3,0) = avA*ovAvRho*(coeffM2 + um*ovAvA); LE(3,1) = -ovAvRho*(nx + uDivA);
      LE(3,2) = -ovAvRho*(ny + vDivA); LE(3,3) = gammaMinus1*ovAvRhoA;
#undef RE
#undef LE
      ev[0] = um; ev[1] = um; ev[2] = um + avA; ev[3] = um - avA;
    } else {
      const double nx = nrm[0], ny = nrm[1], nz = nrm[2];
      const double avRho = lin.rho, avU = lin.v[0], avV = lin.v[1], avW = lin.v[2];
      const double avH = lin.H, avA = lin.A;
      const double avK = 0.5*lin.V*lin.V;
      const double um = avU*nx + avV*ny + avW*nz;
      const double ra = 0.5*avRho/avA;
      const double avA2 = avA*avA;
      const double invAvRho = 1./avRho;
      const double k1 = gammaMinus1*avK/avA2;
      const double k2 = 1.0 - k1;
      const double k3 = -gammaMinus1/avA2;
      const double uDivA = gammaMinus1*avU/avA;
      const double vDivA = gammaMinus1*avV/avA;
      const double wDivA = gammaMinus1*avW/avA;
      const double uDivA2 = uDivA/avA;
      const double vDivA2 = vDivA/avA;
      const double wDivA2 = wDivA/avA;
      const double rhoA = avRho*avA;
#define RE(i,j) R[(i)*5+(j)]
#define LE(i,j) L[(i)*5+(j)]
      RE(0,0) = nx; RE(0,1) = ny; RE(0,2) = nz; RE(0,3) = ra; RE(0,4) = ra;
      RE(1,0) = avU*nx; RE(1,1) = avU*ny - avRho*nz; RE(1,2) = avU*nz + avRho*ny;
      RE(1,3) = ra*(avU + avA*nx); RE(1,4) = ra*(avU - avA*nx);
      RE(2,0) = avV*nx + avRho*nz; RE(2,1) = avV*ny; RE(2,2) = avV*nz - avRho*nx;
      RE(2,3) = ra*(avV + avA*ny); RE(2,4) = ra*(avV - avA*ny);
      RE(3,0) = avW*nx - avRho*ny; RE(3,1) = avW*ny + avRho*nx; RE(3,2) = avW*nz;
      RE(3,3) = ra*(avW + avA*nz); RE(3,4) = ra*(avW - avA*nz);
      RE(4,0) = avK*nx + avRho*(avV*nz - avW*ny);
      RE(4,1) = avK*ny + avRho*(avW*nx - avU*nz);
      RE(4,2) = avK*nz + avRho*(avU*ny - avV*nx);
      RE(4,3) = ra*(avH + avA*um); RE(4,4) = ra*(avH - avA*um);
      LE(0,0) = nx*k2 - invAvRho*(avV*nz - avW*ny); LE(0,1) = uDivA2*nx;
      LE(0,2) = vDivA2*nx + nz*invAvRho; LE(0,3) = wDivA2*nx - ny*invAvRho; LE(0,4) = k3*nx;
      LE(1,0) = ny*k2 - invAvRho*(avW*nx - avU*nz); LE(1,1) = uDivA2*ny - nz*invAvRho;
      LE(1,2) = vDivA2*ny; LE(1,3) = wDivA2*ny + nx*invAvRho; LE(1,4) = k3*ny;
      LE(2,0) = nz*k2 - invAvRho*(avU*ny - avV*nx); LE(2,1) = uDivA2*nz + ny*invAvRho;
      LE(2,2) = vDivA2*nz - nx*invAvRho; LE(2,3) = wDivA2*nz; LE(2,4) = k3*nz;
      LE(3,0) = avA*invAvRho*(k1 - um/avA); LE(3,1) = invAvRho*(nx - uDivA);
      LE(3,2) = invAvRho*(ny - vDivA); LE(3,3) = invAvRho*(nz - wDivA); LE(3,4) = gammaMinus1/rhoA;
      LE(4,0) = avA*invAvRho*(k1 + um/avA); LE(4,1) = invAvRho*(-nx - uDivA);
      LE(4,2) = invAvRho*(-ny - vDivA); LE(4,3) = invAvRho*(-nz - wDivA); LE(4,4) = gammaMinus1/rhoA;
#undef RE
#undef LE
      ev[0] = um; ev[1] = um; ev[2] = um; ev[3] = um + avA; ev[4] = um - avA;
    }
  }

  // RoeFlux.cxx:120-163: flux per unit area.  uL/uR are the reconstructed states in
  // UPDATE variables, pL/pR their physical data.  Returns false on a2 < 0.
  //   result = 0.5*(_sumFlux - k*(_rightEv*(_absEvalues*_leftEv))*(stateR - stateL))
  // is evaluated lazily by MatExprT.hh:255-390 / ExprT.hh as
  //   result_i = 0.5*(sumFlux_i - sum_k [ (k * sum_m R(i,m)*(|ev_m|*L(m,k))) * (sR_k - sL_k) ])
  // (pinned against the reference headers by oracle/ref_mathtools.cpp).
  // etaSA: RoeSA only (RoeSAFlux.cxx:66-120), the H-correction the solver gathered from the neighbouring faces
  static bool roeFlux(const Config& c, const double* uL, const double* uR,
                      const PData& pL, const PData& pR, const double* n, double* result, double etaSA = 0.) {
    double sL[Q], sR[Q], zL[Q], zR[Q];
    toCons(c, uL, sL); toCons(c, uR, sR);
    consToRoe(c, sL, zL); consToRoe(c, sR, zR);
    PData lin;
    if (!linearize(c, zL, zR, lin)) return false;
    double Rm[Q*Q], Lm[Q*Q], ev[Q], aev[Q];
    eigen(c, lin, n, Rm, Lm, ev);
    for (int i = 0; i < Q; ++i) aev[i] = std::abs(ev[i]);     // RoeFlux.cxx:200-203
    if (c.fluxSplitter == FS_ROESA) {
      // RoeSAFlux::setAbsEigenValues, RoeSAFlux.cxx:100-119
      for (int i = 0; i < Q; ++i) {
        const double absEv = std::abs(ev[i]);
        switch (c.entropyFixID) {
        case 1: aev[i] = (absEv >= 2.*etaSA) ? absEv : 0.25*absEv*absEv/etaSA + etaSA; break;
        case 2: aev[i] = std::max(absEv, etaSA); break;
        case 3: aev[i] = absEv + etaSA; break;
        default: aev[i] = absEv + etaSA; break;
        }
      }
    } else if (c.entropyFixID != 0) {
      // RoeEntropyFixFlux::setAbsEigenValues (RoeEntropyFixFlux.cxx:57-90) + computeLambdaCorr (RoeEntropyFixFlux.hh:49-56)
      double evL[Q], evR[Q];
      eigenValues(pL, n, evL);
      eigenValues(pR, n, evR);
      double lambdaCorr = 0.0;
      for (int i = 0; i < Q; ++i) lambdaCorr = std::max(lambdaCorr, std::abs(evR[i] - evL[i]));
      lambdaCorr *= 0.5;
      if (c.entropyFixID == 1) {
        for (int i = 0; i < Q; ++i) aev[i] = std::max(std::abs(ev[i]), lambdaCorr);
      } else if (c.entropyFixID == 2) {
        const double twoLambdaCorr = 2.0*lambdaCorr;
        for (int i = 0; i < Q; ++i) {
          const double absLambda = std::abs(ev[i]);
          aev[i] = (!(std::abs(ev[i]) < twoLambdaCorr)) ? absLambda : (absLambda*absLambda/(4.0*lambdaCorr) + lambdaCorr);
        }
      }
    }
    double sumFlux[Q], f2[Q];
    convFlux(pL, n, sumFlux);
    convFlux(pR, n, f2);
    for (int i = 0; i < Q; ++i) sumFlux[i] += f2[i];
    const double kred = c.diffRedCoeff;
    for (int i = 0; i < Q; ++i) {
      double res = 0.;
      for (int k = 0; k < Q; ++k) {
        double m = 0.;
        for (int mm = 0; mm < Q; ++mm) m += Rm[i*Q+mm]*(aev[mm]*Lm[mm*Q+k]);
        res += (kred*m)*(sR[k] - sL[k]);
      }
      result[i] = 0.5*(sumFlux[i] - res);
    }
    return true;
  }

  // ---- AUSM family ---------------------------------------------------------
  // AUSMFlux.hh:147-174 (pow(x,2.) is folded to x*x by the compiler)
  static double mach1Plus(double m) { return 0.5*(m + std::abs(m)); }
  static double mach1Min(double m) { return 0.5*(m - std::abs(m)); }
  static double mach2Plus(double m) { return 0.25*((m + 1.0)*(m + 1.0)); }
  static double mach2Min(double m) { return -0.25*((m - 1.0)*(m - 1.0)); }

  // AUSMFlux.ci:547-581 computeInterfaceSoundSpeed -> computeSoundSpeed1/2/3 (:585-646)
  static double ausmSoundSpeed(const Config& c, const PData& L, const PData& R, double unL, double unR) {
    switch (c.choiceA12) {
      case 2: return 0.5*(L.A + R.A);
      case 3: return std::sqrt(L.A*R.A);
      default: {
        const double gammaL = c.gamma, gammaMinus1L = gammaL - 1.0;
        const double gammaR = c.gamma, gammaMinus1R = gammaR - 1.0;
        const double aCritL = std::sqrt((2.0*gammaMinus1L/(gammaL + 1.0))*L.H);
        const double aCritR = std::sqrt((2.0*gammaMinus1R/(gammaR + 1.0))*R.H);
        const double acL = (aCritL*aCritL)/std::max(aCritL, unL);
        const double acR = (aCritR*aCritR)/std::max(aCritR, -unR);
        return std::min(acL, acR);
      }
    }
  }

  // AUSMFlux.ci:123-222 compute (single Euler set, no mesh speed, no incompressible correction:
  // AUSMFlux.hh:97 switches it off) with computeMassFlux / computePressureFlux of
  // AUSMPlusFlux.ci:47-105 (FS_AUSMPLUS) or AUSMPlusUpFlux.ci:50-135 (FS_AUSMPLUSUP), and
  // computeMassFluxImpl AUSMFlux.ci:497-543.  Flux per unit area.
  static bool ausmFlux(const Config& c, const PData& L, const PData& R, const double* n, double* result) {
    double unL = 0.0, unR = 0.0;
    for (int i = 0; i < D; ++i) { unL += L.v[i]*n[i]; unR += R.v[i]*n[i]; }
    const double a12 = ausmSoundSpeed(c, L, R, unL, unR);
    const double mL = unL/a12, mR = unR/a12;
    const double rhoL = L.rho, rhoR = R.rho, pL = L.p, pR = R.p;
    double mflux12, p12;
    if (c.fluxSplitter == FS_AUSMPLUS) {
      const double beta = c.ausmBeta, alpha = c.ausmAlpha;
      const double sqL = (mL*mL - 1.0)*(mL*mL - 1.0), sqR = (mR*mR - 1.0)*(mR*mR - 1.0);
      const double M4Plus = (std::abs(mL) >= 1.0) ? mach1Plus(mL) : mach2Plus(mL) + beta*sqL;
      const double M4Minus = (std::abs(mR) >= 1.0) ? mach1Min(mR) : mach2Min(mR) - beta*sqR;
      const double m12 = M4Plus + M4Minus;
      mflux12 = (m12 > 0.0) ? a12*m12*rhoL : a12*m12*rhoR;
      const double P5Plus = (std::abs(mL) >= 1.0) ? 0.5*(1 + ((mL < 0.0) ? -1.0 : 1.0)) :
        mach2Plus(mL)*(2.0 - mL) + alpha*mL*sqL;
      const double P5Minus = (std::abs(mR) >= 1.0) ? 0.5*(1 - ((mR < 0.0) ? -1.0 : 1.0)) :
        -mach2Min(mR)*(2.0 + mR) - alpha*mR*sqR;
      p12 = P5Plus*pL + P5Minus*pR;
    } else {
      const double beta = c.ausmBeta;
      const double mBarSq = (unL*unL + unR*unR)/(2.0*a12*a12);
      const double mInf = c.machInf;
      const double mZero = std::sqrt(std::min(1.0, std::max(mBarSq, mInf*mInf)));
      const double fa = mZero*(2.0 - mZero);
      const double M4Plus = (std::abs(mL) >= 1.0) ? mach1Plus(mL) : mach2Plus(mL)*(1.0 - 16.*beta*mach2Min(mL));
      const double M4Minus = (std::abs(mR) >= 1.0) ? mach1Min(mR) : mach2Min(mR)*(1.0 + 16.*beta*mach2Plus(mR));
      const double rhoa2 = 0.5*(rhoL + rhoR)*a12*a12;
      const double mP = (-c.ausmKp/fa)*std::max(1.0 - c.ausmSigma*mBarSq, 0.0)*(pR - pL)/rhoa2;
      const double m12 = M4Plus + M4Minus + mP;
      mflux12 = (m12 > 0.0) ? a12*m12*rhoL : a12*m12*rhoR;
      const double alpha = (3.0/16.0)*(-4.0 + 5.0*fa*fa);
      const double P5Plus = (std::abs(mL) >= 1.0) ? mach1Plus(mL)/mL :
        mach2Plus(mL)*((2.0 - mL) - 16.*alpha*mL*mach2Min(mL));
      const double P5Minus = (std::abs(mR) >= 1.0) ? mach1Min(mR)/mR :
        mach2Min(mR)*((-2.0 - mR) + 16.*alpha*mR*mach2Plus(mR));
      const double pU = -c.ausmKu*P5Plus*P5Minus*(rhoL + rhoR)*fa*a12*(unR - unL);
      p12 = P5Plus*pL + P5Minus*pR + pU;
    }
    const PData& up = (mflux12 > 0.0) ? L : R;
    result[0] = mflux12;
    for (int i = 0; i < D; ++i) result[1+i] = mflux12*up.v[i] + p12*n[i];
    result[TID] = mflux12*up.H;
    return true;
  }

  // ---- a19: analytic Jacobian of the AUSM+up flux in conservative variables ----------------------------
  // MathFunctions.hh:58-61
  static double sigmoidF(double x, double y, double K) { return 0.5*(1.0 - (1.0 + K)*(x - y)/(-std::abs(x - y) - K)); }

  // ---- a19: analytic Jacobian of the AUSM+up flux in Puvt / Pvt variables -------------------------------------------
  // AUSMPlusUpFlux<UPDATEVAR>::ComputedFluxdPvt (AUSMPlusUpFlux.ci:160-668), column by column: p (:320-391), u v w
  // (:394-597, one form), T (:599-664).  Its quirks are kept: d rho / d p and d rho / d T use the dimensional constants of
  // air (Mmass, Rgas :183-184) whatever the model's R; the T column leaves the d rho12 term out of d mP (:633-635) and sums
  // d mdot12 in another order than the p column (:641 vs :363); the p column has no mdot12 dH term in the energy row (:390).
  static void ausmPlusUpJacobPvt(const Config& c, const PData& L, const PData& R, const double* n, int side, double* J) {
    const bool LEFT = (side == 0);
    const double rhoL = L.rho, rhoR = R.rho, HL = L.H, HR = R.H, TL = L.T, TR = R.T;
    const double gammaL = c.gamma, gammaR = c.gamma, pL = L.p, pR = R.p;
    const double UL = L.v[0], UR = R.v[0], VL = L.v[1], VR = R.v[1], WL = L.v[2], WR = R.v[2];
    const double Mmass = 28.965338E-3, Rgas = 8.3145;
    const double beta = c.ausmBeta, Ksig = c.coeffSigmoid, sigma = c.ausmSigma;
    double qnL = 0.0, qnR = 0.0;
    for (int i = 0; i < D; ++i) { qnL += L.v[i]*n[i]; qnR += R.v[i]*n[i]; }
    // computeSoundSpeedForJacobian (AUSMFlux.ci:795-822)
    const double aCritL = std::sqrt(2.0*(gammaL - 1.0)/(gammaL + 1.0)*HL);
    const double aCritR = std::sqrt(2.0*(gammaR - 1.0)/(gammaR + 1.0)*HR);
    const double aCritLsig = sigmoidF(aCritL, qnL, Ksig);
    const double OneaCritLsig = (1.0 - aCritLsig);
    const double aCritRsig = sigmoidF(aCritR, -qnR, Ksig);
    const double OneaCritRsig = -(1.0 - aCritRsig);
    const double asigL = aCritLsig*aCritL + (1.0 - aCritLsig)*qnL;
    const double asigR = aCritRsig*aCritR - (1.0 - aCritRsig)*qnR;
    const double ahatL = aCritL*aCritL/std::max(aCritL, qnL);
    const double ahatR = aCritR*aCritR/std::max(aCritR, -qnR);
    const double a12 = (1.0 - sigmoidF(ahatL, ahatR, Ksig))*aCritL*aCritL/asigL + sigmoidF(ahatL, ahatR, Ksig)*aCritR*aCritR/asigR;

    const double rho12 = 0.5*(rhoL + rhoR);
    const double ML = qnL/a12, MR = qnR/a12;
    const double mInf = c.machInf;
    const double MBarSq = 0.5*(ML*ML + MR*MR);
    const double MinfSq = mInf*mInf;
    const double sigMbarsqMinfsq = sigmoidF(MBarSq, MinfSq, Ksig);
    const double sig2 = sigMbarsqMinfsq*MBarSq + (1.0 - sigMbarsqMinfsq)*MinfSq;
    const double MZeroSig = sigmoidF(1.0, sig2, Ksig);
    const double mZero = std::sqrt((1 - MZeroSig)*1.0 + MZeroSig*sig2);
    const double fa = mZero*(2.0 - mZero);
    const double sigmaMBarSq = 1.0 - sigma*MBarSq;
    const double mPSig = sigmoidF(sigmaMBarSq, 0.0, Ksig);
    const double dMpCon = -(c.ausmKp)*mPSig;
    const double M4Plus = (std::abs(ML) >= 1.0) ? mach1Plus(ML) : mach2Plus(ML)*(1.0 - 16.*beta*mach2Min(ML));
    const double M4Minus = (std::abs(MR) >= 1.0) ? mach1Min(MR) : mach2Min(MR)*(1.0 + 16.*beta*mach2Plus(MR));
    const double mP = -(c.ausmKp)*mPSig*sigmaMBarSq*(pR - pL)/(fa*rho12*a12*a12);
    const double M12 = M4Plus + M4Minus + mP;
    const double sigrho = sigmoidF(M12, 0.0, Ksig);
    const double Onesigrho = (1.0 - sigrho);
    const double rhostar = sigrho*rhoL + Onesigrho*rhoR;
    const double mdot12 = M12*a12*rhostar;
    const double sigm = sigmoidF(mdot12, 0.0, Ksig);
    const double Onesigm = (1.0 - sigm);
    const double alpha = (3.0/16.0)*(-4.0 + 5.0*fa*fa);
    const double P5Plus = (std::abs(ML) >= 1.0) ? mach1Plus(ML)/ML : mach2Plus(ML)*((2.0 - ML) - 16.*alpha*ML*mach2Min(ML));
    const double P5Minus = (std::abs(MR) >= 1.0) ? mach1Min(MR)/MR : mach2Min(MR)*((-2.0 - MR) + 16.*alpha*MR*mach2Plus(MR));
    // global constants (:283-317)
    const double rhoa2 = fa*rho12*a12*a12;
    const double prho2a2 = -(pR - pL)/(fa*rho12*rho12*a12*a12);
    const double prhoa3 = -2.0*(pR - pL)/(fa*rho12*a12*a12*a12);
    const double prhoa2fa = -(pR - pL)/(fa*fa*rho12*a12*a12);
    const double daCritLconst = std::sqrt((gammaL - 1.0)/(2.0*(gammaL + 1.0)*HL));
    const double daCritRconst = std::sqrt((gammaR - 1.0)/(2.0*(gammaR + 1.0)*HR));
    const double dahatCon1L = 2.0*aCritL/asigL, dahatCon1R = 2.0*aCritR/asigR;
    const double dahatCon2L = -aCritL*aCritL/(asigL*asigL), dahatCon2R = -aCritR*aCritR/(asigR*asigR);
    const double dahatSig = sigmoidF(ahatL, ahatR, Ksig);
    const double OnedahatSig = (1.0 - dahatSig);
    const double dMCon = 1.0/a12;
    const double dMConL = -qnL/(a12*a12), dMConR = -qnR/(a12*a12);
    const double dmZeroCon = 0.5/mZero*sigMbarsqMinfsq*MZeroSig*sigmaMBarSq;
    const double dfaCon = 2.0*(1.0 - mZero);
    const double dM4Lsign = (ML < 0.0) ? -1.0 : 1.0, dM4Rsign = (MR < 0.0) ? -1.0 : 1.0;
    const double OnedM4Lsign = 0.5*(1. + dM4Lsign), OnedM4Rsign = 0.5*(1. - dM4Rsign);
    const double dM4Lsuper = (0.5*(ML + 1.) + 4.*beta*ML*(ML*ML - 1.));
    const double dM4Rsuper = -(0.5*(MR - 1.) + 4.*beta*MR*(MR*MR - 1.));
    const double a12rhostar = a12*rhostar, M12a12 = M12*a12, M12rhostar = M12*rhostar;
    const double dalphaCon = 15.0/8.0*fa;
    const double dP5Lconst1 = 0.5*(10.0*alpha*ML*ML*ML*ML - (1.5 + 12.0*alpha)*ML*ML + 1.5 + 2.0*alpha);
    const double dP5Rconst1 = -0.5*(10.0*alpha*MR*MR*MR*MR - (1.5 + 12.0*alpha)*MR*MR + 1.5 + 2.0*alpha);
    const double dP5Lconst2 = ML*(ML*ML - 1)*(ML*ML - 1);
    const double dP5Rconst2 = -MR*(MR*MR - 1)*(MR*MR - 1);
    const double pudP5Plus = P5Minus*rho12*fa*a12*(qnR - qnL);
    const double pudP5Minus = P5Plus*rho12*fa*a12*(qnR - qnL);
    const double pudrho12 = P5Plus*P5Minus*fa*a12*(qnR - qnL);
    const double pudfa = P5Plus*P5Minus*rho12*a12*(qnR - qnL);
    const double puda12 = P5Plus*P5Minus*rho12*fa*(qnR - qnL);
    const double pudqn = P5Plus*P5Minus*rho12*fa;
    const double dpuCon = 0.5*c.ausmKu;
    const double Usig = sigm*UL + Onesigm*UR, Vsig = sigm*VL + Onesigm*VR, Wsig = sigm*WL + Onesigm*WR;
    const double Hsig = sigm*HL + Onesigm*HR;
    const bool supL = std::abs(ML) >= 1.0, supR = std::abs(MR) >= 1.0;
    for (int k = 0; k < Q*Q; ++k) J[k] = 0.;
    const double vs[3] = {Usig, Vsig, Wsig};

    // kind 0: p column, 1: a velocity column, 2: T column
    struct Col { double dmdot, dp12; };
    auto column = [&](int kind, double dqnL, double dqnR, double dhL, double dhR, double dpL, double dpR,
                      double drhoLd, double drhoRd) -> Col {
      const double drho12d = 0.5*(drhoLd + drhoRd);
      const double drhostard = sigrho*drhoLd + Onesigrho*drhoRd;
      const double daCritLd = daCritLconst*dhL, daCritRd = daCritRconst*dhR;
      const double dasigLd = (kind == 1) ? aCritLsig*daCritLd + OneaCritLsig*dqnL : aCritLsig*daCritLd;
      const double dasigRd = (kind == 1) ? aCritRsig*daCritRd + OneaCritRsig*dqnR : aCritRsig*daCritRd;
      const double dahatLd = (dahatCon1L*daCritLd + dahatCon2L*dasigLd);
      const double dahatRd = (dahatCon1R*daCritRd + dahatCon2R*dasigRd);
      const double da12d = OnedahatSig*dahatLd + dahatSig*dahatRd;
      const double dMLd = (kind == 1) ? dMCon*dqnL + dMConL*da12d : dMConL*da12d;
      const double dMRd = (kind == 1) ? dMCon*dqnR + dMConR*da12d : dMConR*da12d;
      const double dMBarSqd = ML*dMLd + MR*dMRd;
      const double dmZerod = dmZeroCon*dMBarSqd;
      const double dfad = dfaCon*dmZerod;
      const double dM4Plusd = supL ? OnedM4Lsign*dMLd : dM4Lsuper*dMLd;
      const double dM4Minusd = supR ? OnedM4Rsign*dMRd : dM4Rsuper*dMRd;
      const double dMpdda = prhoa3*da12d;
      const double dMpddfa = prhoa2fa*dfad;
      double dMpd;
      if (kind == 0) {
        const double dMpddp = (dpR - dpL)/rhoa2;
        const double dMpddrho = prho2a2*drho12d;
        dMpd = dMpCon*(dMpddp + dMpddrho + dMpdda + dMpddfa);
      } else dMpd = dMpCon*(dMpdda + dMpddfa);
      const double dM12d = dM4Plusd + dM4Minusd + dMpd;
      double dmdotd;
      if (kind == 0) dmdotd = dM12d*a12rhostar + M12a12*drhostard + M12rhostar*da12d;
      else if (kind == 1) dmdotd = dM12d*a12rhostar + M12rhostar*da12d;
      else dmdotd = dM12d*a12rhostar + M12rhostar*da12d + M12a12*drhostard;
      const double dalphad = dalphaCon*dfad;
      const double dP5Plusd = supL ? 0. : dP5Lconst1*dMLd + dP5Lconst2*dalphad;
      const double dP5Minusd = supR ? 0. : dP5Rconst1*dMRd + dP5Rconst2*dalphad;
      const double t1 = pudP5Plus*dP5Plusd, t2 = pudP5Minus*dP5Minusd, tfa = pudfa*dfad, tda = puda12*da12d;
      const double trho = (kind == 1) ? 0.0 : pudrho12*drho12d;
      const double tqn = (kind == 1) ? pudqn*da12d*(dqnR - dqnL) : 0.0;
      const double dpud = dpuCon*(t1 + t2 + trho + tfa + tda + tqn);
      Col o;
      o.dmdot = dmdotd;
      if (kind == 0) o.dp12 = dP5Plusd*pL + P5Plus*dpL + dP5Minusd*pR + P5Minus*dpR + dpud;
      else o.dp12 = dP5Plusd*pL + dP5Minusd*pR + dpud;
      return o;
    };
    // dFdp
    {
      const double drhoLd = LEFT ? Mmass/(Rgas*TL) : 0., drhoRd = LEFT ? 0. : Mmass/(Rgas*TR);
      const double dhL = LEFT ? -gammaL*pL/((gammaL - 1)*rhoL*rhoL) : 0.;
      const double dhR = LEFT ? 0. : -gammaR*pR/((gammaR - 1)*rhoR*rhoR);
      const Col k = column(0, 0., 0., dhL, dhR, LEFT ? 1. : 0., LEFT ? 0. : 1., drhoLd, drhoRd);
      J[0*Q + 0] = k.dmdot;
      for (int r = 0; r < D; ++r) J[(1+r)*Q + 0] = k.dmdot*vs[r] + k.dp12*n[r];
      J[(Q-1)*Q + 0] = k.dmdot*Hsig;
    }
    // dFdu, dFdv, dFdw
    for (int m = 0; m < D; ++m) {
      const double dqnLd = LEFT ? n[m] : 0., dqnRd = LEFT ? 0. : n[m];
      const double dhL = LEFT ? L.v[m] : 0., dhR = LEFT ? 0. : R.v[m];
      const double dvelL = LEFT ? 1. : 0., dvelR = LEFT ? 0. : 1.;
      const Col k = column(1, dqnLd, dqnRd, dhL, dhR, 0., 0., 0., 0.);
      const int col = 1 + m;
      J[0*Q + col] = k.dmdot;
      for (int r = 0; r < D; ++r) {
        if (r == m) J[(1+r)*Q + col] = k.dmdot*vs[r] + mdot12*(sigm*dvelL + Onesigm*dvelR) + k.dp12*n[r];
        else J[(1+r)*Q + col] = k.dmdot*vs[r] + k.dp12*n[r];
      }
      J[(Q-1)*Q + col] = k.dmdot*Hsig + mdot12*(sigm*dhL + Onesigm*dhR);
    }
    // dFdT
    {
      const double drhoLd = LEFT ? -Mmass*pL/(Rgas*TL*TL) : 0., drhoRd = LEFT ? 0. : -Mmass*pR/(Rgas*TR*TR);
      const double dhL = LEFT ? gammaL/(rhoL*(gammaL - 1)) : 0., dhR = LEFT ? 0. : gammaR/(rhoR*(gammaR - 1));
      const Col k = column(2, 0., 0., dhL, dhR, 0., 0., drhoLd, drhoRd);
      const int col = Q - 1;
      J[0*Q + col] = k.dmdot;
      for (int r = 0; r < D; ++r) J[(1+r)*Q + col] = k.dmdot*vs[r] + k.dp12*n[r];
      J[(Q-1)*Q + col] = k.dmdot*Hsig + mdot12*(sigm*dhL + Onesigm*dhR);
    }
  }


  // AUSMPlusUpFlux<UPDATEVAR>::ComputedFluxdCons (AUSMPlusUpFlux.ci:675-1203) with
  // AUSMFlux::computeSoundSpeedForJacobian (AUSMFlux.ci:795-822).  side 0 = LEFT, 1 = RIGHT;
  // J is row-major Q x Q: J[i*Q + j] = d F_i / d U_j (per unit area).  The columns follow the reference one by
  // one, including the terms it leaves out where a derivative vanishes and its row-(w, rho w) quirk.
  static void ausmPlusUpJacobCons(const Config& c, const PData& L, const PData& R, const double* n, int side, double* J) {
    const bool LEFT = (side == 0);
    const double rhoL = L.rho, rhoR = R.rho, HL = L.H, HR = R.H, EL = L.E, ER = R.E;
    const double gammaL = c.gamma, gammaR = c.gamma, pL = L.p, pR = R.p;
    const double UL = L.v[0], UR = R.v[0], VL = L.v[1], VR = R.v[1], WL = L.v[2], WR = R.v[2];
    const double beta = c.ausmBeta, Ksig = c.coeffSigmoid, sigma = c.ausmSigma;
    double qnL = 0.0, qnR = 0.0;
    for (int i = 0; i < D; ++i) { qnL += L.v[i]*n[i]; qnR += R.v[i]*n[i]; }
    // computeSoundSpeedForJacobian
    const double aCritL = std::sqrt(2.0*(gammaL - 1.0)/(gammaL + 1.0)*HL);
    const double aCritR = std::sqrt(2.0*(gammaR - 1.0)/(gammaR + 1.0)*HR);
    const double aCritLsig = sigmoidF(aCritL, qnL, Ksig);
    const double OneaCritLsig = (1.0 - aCritLsig);
    const double aCritRsig = sigmoidF(aCritR, -qnR, Ksig);
    const double OneaCritRsig = -(1.0 - aCritRsig);
    const double asigL = aCritLsig*aCritL + (1.0 - aCritLsig)*qnL;
    const double asigR = aCritRsig*aCritR - (1.0 - aCritRsig)*qnR;
    const double ahatL = aCritL*aCritL/std::max(aCritL, qnL);
    const double ahatR = aCritR*aCritR/std::max(aCritR, -qnR);
    const double a12 = (1.0 - sigmoidF(ahatL, ahatR, Ksig))*aCritL*aCritL/asigL + sigmoidF(ahatL, ahatR, Ksig)*aCritR*aCritR/asigR;

    const double rho12 = 0.5*(rhoL + rhoR);
    const double ML = qnL/a12, MR = qnR/a12;
    const double mInf = c.machInf;
    const double MBarSq = 0.5*(ML*ML + MR*MR);
    const double MinfSq = mInf*mInf;
    const double sigMbarsqMinfsq = sigmoidF(MBarSq, MinfSq, Ksig);
    const double sig2 = sigMbarsqMinfsq*MBarSq + (1.0 - sigMbarsqMinfsq)*MinfSq;
    const double MZeroSig = sigmoidF(1.0, sig2, Ksig);
    const double mZero = std::sqrt((1 - MZeroSig)*1.0 + MZeroSig*sig2);
    const double fa = mZero*(2.0 - mZero);
    const double sigmaMBarSq = 1.0 - sigma*MBarSq;
    const double mPSig = sigmoidF(sigmaMBarSq, 0.0, Ksig);
    const double dMpCon = -(c.ausmKp)*mPSig;
    const double M4Plus = (std::abs(ML) >= 1.0) ? mach1Plus(ML) : mach2Plus(ML)*(1.0 - 16.*beta*mach2Min(ML));
    const double M4Minus = (std::abs(MR) >= 1.0) ? mach1Min(MR) : mach2Min(MR)*(1.0 + 16.*beta*mach2Plus(MR));
    const double mP = -(c.ausmKp)*mPSig*(pR - pL)/(rho12*a12*a12);
    const double M12 = M4Plus + M4Minus + mP;
    const double sigrho = sigmoidF(M12, 0.0, Ksig);
    const double Onesigrho = (1.0 - sigrho);
    const double rhostar = sigrho*rhoL + Onesigrho*rhoR;
    const double mdot12 = M12*a12*rhostar;
    const double sigm = sigmoidF(mdot12, 0.0, Ksig);
    const double Onesigm = (1.0 - sigm);
    const double alpha = (3.0/16.0)*(-4.0 + 5.0*fa*fa);
    const double P5Plus = (std::abs(ML) >= 1.0) ? mach1Plus(ML)/ML : mach2Plus(ML)*((2.0 - ML) - 16.*alpha*ML*mach2Min(ML));
    const double P5Minus = (std::abs(MR) >= 1.0) ? mach1Min(MR)/MR : mach2Min(MR)*((-2.0 - MR) + 16.*alpha*MR*mach2Plus(MR));
    // global constants
    const double rhoa2 = rho12*a12*a12;
    const double prho2a2 = (pR - pL)/(rho12*rho12*a12*a12);
    const double prhoa3 = -2.0*(pR - pL)/(rho12*a12*a12*a12);
    const double daCritLconst = std::sqrt((gammaL - 1.0)/(2.0*(gammaL + 1.0)*HL));
    const double daCritRconst = std::sqrt((gammaR - 1.0)/(2.0*(gammaR + 1.0)*HR));
    const double dahatCon1L = 2.0*aCritL/asigL, dahatCon1R = 2.0*aCritR/asigR;
    const double dahatCon2L = -aCritL*aCritL/(asigL*asigL), dahatCon2R = -aCritR*aCritR/(asigR*asigR);
    const double dahatSig = sigmoidF(ahatL, ahatR, Ksig);
    const double OnedahatSig = (1.0 - dahatSig);
    const double dMCon = 1.0/a12;
    const double dMConL = -qnL/(a12*a12), dMConR = -qnR/(a12*a12);
    const double dmZeroCon = 0.5/mZero*sigMbarsqMinfsq*MZeroSig;
    const double dfaCon = 2.0*(1.0 - mZero);
    const double dM4Lsign = (ML < 0.0) ? -1.0 : 1.0, dM4Rsign = (MR < 0.0) ? -1.0 : 1.0;
    const double OnedM4Lsign = 0.5*(1. + dM4Lsign), OnedM4Rsign = 0.5*(1. - dM4Rsign);
    const double dM4Lsuper = (0.5*(ML + 1.) + 4.*beta*ML*(ML*ML - 1.));
    const double dM4Rsuper = -(0.5*(MR - 1.) + 4.*beta*MR*(MR*MR - 1.));
    const double a12rhostar = a12*rhostar, M12a12 = M12*a12, M12rhostar = M12*rhostar;
    const double dalphaCon = 15.0/8.0*fa;
    const double dP5Lconst1 = 0.5*(10.0*alpha*ML*ML*ML*ML - (1.5 + 12.0*alpha)*ML*ML + 1.5 + 2.0*alpha);
    const double dP5Rconst1 = -0.5*(10.0*alpha*MR*MR*MR*MR - (1.5 + 12.0*alpha)*MR*MR + 1.5 + 2.0*alpha);
    const double dP5Lconst2 = ML*(ML*ML - 1)*(ML*ML - 1);
    const double dP5Rconst2 = -MR*(MR*MR - 1)*(MR*MR - 1);
    const double pudP5Plus = P5Minus*rho12*fa*a12*(qnR - qnL);
    const double pudP5Minus = P5Plus*rho12*fa*a12*(qnR - qnL);
    const double pudrho12 = P5Plus*P5Minus*fa*a12*(qnR - qnL);
    const double pudfa = P5Plus*P5Minus*rho12*a12*(qnR - qnL);
    const double puda12 = P5Plus*P5Minus*rho12*fa*(qnR - qnL);
    const double pudqn = P5Plus*P5Minus*rho12*fa;
    const double dpuCon = 0.5*c.ausmKu;
    const double Usig = sigm*UL + Onesigm*UR, Vsig = sigm*VL + Onesigm*VR, Wsig = sigm*WL + Onesigm*WR;
    const double Hsig = sigm*HL + Onesigm*HR;
    const bool supL = std::abs(ML) >= 1.0, supR = std::abs(MR) >= 1.0;
    for (int k = 0; k < Q*Q; ++k) J[k] = 0.;

    // One column: the chain shared by all of them; `kind` says which optional terms the reference writes.
    //   kind 0: rho (drho12, drhostar, qn in asig and pu)   1: momentum (qn in asig and pu)   2: rhoE (no qn terms)
    struct Col { double dmdot, dp12; };
    auto column = [&](int kind, double dqnL, double dqnR, double dhL, double dhR, double dpL, double dpR,
                      double drhoLd, double drhoRd) -> Col {
      const double daCritLd = daCritLconst*dhL, daCritRd = daCritRconst*dhR;
      const double dasigLd = (kind == 2) ? aCritLsig*daCritLd : aCritLsig*daCritLd + OneaCritLsig*dqnL;
      const double dasigRd = (kind == 2) ? aCritRsig*daCritRd : aCritRsig*daCritRd + OneaCritRsig*dqnR;
      const double dahatLd = (dahatCon1L*daCritLd + dahatCon2L*dasigLd);
      const double dahatRd = (dahatCon1R*daCritRd + dahatCon2R*dasigRd);
      const double da12d = OnedahatSig*dahatLd + dahatSig*dahatRd;
      const double dMLd = (kind == 2) ? dMConL*da12d : dMCon*dqnL + dMConL*da12d;
      const double dMRd = (kind == 2) ? dMConR*da12d : dMCon*dqnR + dMConR*da12d;
      const double dMBarSqd = ML*dMLd + MR*dMRd;
      const double dmZerod = dmZeroCon*dMBarSqd;
      const double dfad = dfaCon*dmZerod;
      const double dM4Plusd = supL ? OnedM4Lsign*dMLd : dM4Lsuper*dMLd;
      const double dM4Minusd = supR ? OnedM4Rsign*dMRd : dM4Rsuper*dMRd;
      const double dMpddp = (dpR - dpL)/rhoa2;
      const double dMpdda = prhoa3*da12d;
      double dMpd, dmdotd;
      const double drho12d = 0.5*(drhoLd + drhoRd);
      if (kind == 0) {
        const double dMpddrho = prho2a2*drho12d;
        dMpd = dMpCon*(dMpddp + dMpddrho + dMpdda);
      } else dMpd = dMpCon*(dMpddp + dMpdda);
      const double dM12d = dM4Plusd + dM4Minusd + dMpd;
      if (kind == 0) {
        const double drhostard = sigrho*drhoLd + Onesigrho*drhoRd;
        dmdotd = dM12d*a12rhostar + M12rhostar*da12d + M12a12*drhostard;
      } else dmdotd = dM12d*a12rhostar + M12rhostar*da12d;
      const double dalphad = dalphaCon*dfad;
      const double dP5Plusd = supL ? 0. : dP5Lconst1*dMLd + dP5Lconst2*dalphad;
      const double dP5Minusd = supR ? 0. : dP5Rconst1*dMRd + dP5Rconst2*dalphad;
      const double t1 = pudP5Plus*dP5Plusd, t2 = pudP5Minus*dP5Minusd, tfa = pudfa*dfad, tda = puda12*da12d;
      double dpud;
      if (kind == 0) {
        const double trho = pudrho12*drho12d;
        const double tqn = pudqn*da12d*(dqnR - dqnL);
        dpud = dpuCon*(t1 + t2 + trho + tfa + tda + tqn);
      } else if (kind == 1) {
        const double tqn = pudqn*da12d*(dqnR - dqnL);
        dpud = dpuCon*(t1 + t2 + tfa + tda + tqn);
      } else dpud = dpuCon*(t1 + t2 + tfa + tda);
      Col o;
      o.dmdot = dmdotd;
      o.dp12 = dP5Plusd*pL + P5Plus*dpL + dP5Minusd*pR + P5Minus*dpR + dpud;
      return o;
    };

    // dFdrho
    {
      const double dqnLd = LEFT ? -qnL/rhoL : 0., dqnRd = LEFT ? 0. : -qnR/rhoR;
      double VLsq, VRsq;
      if (D == 3) { VLsq = UL*UL + VL*VL + WL*WL; VRsq = UR*UR + VR*VR + WR*WR; }
      else { VLsq = UL*UL + VL*VL; VRsq = UR*UR + VR*VR; }
      const double dhL = LEFT ? -gammaL*EL/rhoL + (gammaL - 1.)*VLsq/rhoL : 0.;
      const double dhR = LEFT ? 0. : -gammaR*ER/rhoR + (gammaR - 1.)*VRsq/rhoR;
      const double dpL = LEFT ? 0.5*(gammaL - 1.)*VLsq : 0., dpR = LEFT ? 0. : 0.5*(gammaR - 1.)*VRsq;
      const double duL = LEFT ? -UL/rhoL : 0., duR = LEFT ? 0. : -UR/rhoR;
      const double dvL = LEFT ? -VL/rhoL : 0., dvR = LEFT ? 0. : -VR/rhoR;
      double dwL = 0., dwR = 0.;
      if (D == 3) { dwL = LEFT ? -WL/rhoL : 0.; dwR = LEFT ? 0. : -WR/rhoR; }
      const Col k = column(0, dqnLd, dqnRd, dhL, dhR, dpL, dpR, LEFT ? 1. : 0., LEFT ? 0. : 1.);
      J[0*Q + 0] = k.dmdot;
      J[1*Q + 0] = k.dmdot*Usig + mdot12*(sigm*duL + Onesigm*duR) + k.dp12*n[0];
      J[2*Q + 0] = k.dmdot*Vsig + mdot12*(sigm*dvL + Onesigm*dvR) + k.dp12*n[1];
      if (D == 3) J[3*Q + 0] = k.dmdot*Wsig + mdot12*(sigm*dwL + Onesigm*dwR) + k.dp12*n[2];
      J[(Q-1)*Q + 0] = k.dmdot*Hsig + mdot12*(sigm*dhL + Onesigm*dhR);
    }
    // dFdrhoU, dFdrhoV, dFdrhoW
    double dmdotV = 0.;
    for (int m = 0; m < D; ++m) {
      const double vLm = L.v[m], vRm = R.v[m];
      const double dqnLd = LEFT ? n[m]/rhoL : 0., dqnRd = LEFT ? 0. : n[m]/rhoR;
      const double dhL = LEFT ? -(gammaL - 1.)*vLm/rhoL : 0., dhR = LEFT ? 0 : -(gammaR - 1.)*vRm/rhoR;
      const double dpL = LEFT ? -(gammaL - 1.)*vLm : 0., dpR = LEFT ? 0. : -(gammaR - 1.)*vRm;
      const double dvelL = LEFT ? 1/rhoL : 0., dvelR = LEFT ? 0. : 1/rhoR;
      const Col k = column(1, dqnLd, dqnRd, dhL, dhR, dpL, dpR, 0., 0.);
      if (m == 1) dmdotV = k.dmdot;
      const int col = 1 + m;
      const double vs[3] = {Usig, Vsig, Wsig};
      J[0*Q + col] = k.dmdot;
      for (int r = 0; r < D; ++r) {
        if (r == m) {
          // the reference's (3,3) entry multiplies Wsigmoid by the rho V column's dmdot12 (AUSMPlusUpFlux.ci:1141)
          const double dm = (m == 2) ? dmdotV : k.dmdot;
          J[(1+r)*Q + col] = dm*vs[r] + mdot12*(sigm*dvelL + Onesigm*dvelR) + k.dp12*n[r];
        } else {
          J[(1+r)*Q + col] = k.dmdot*vs[r] + k.dp12*n[r];
        }
      }
      J[(Q-1)*Q + col] = k.dmdot*Hsig + mdot12*(sigm*dhL + Onesigm*dhR);
    }
    // dFdrhoE
    {
      const double dhL = LEFT ? gammaL/rhoL : 0., dhR = LEFT ? 0. : gammaR/rhoR;
      const double dpL = LEFT ? gammaL - 1. : 0., dpR = LEFT ? 0. : gammaR - 1.;
      const Col k = column(2, 0., 0., dhL, dhR, dpL, dpR, 0., 0.);
      const int col = Q - 1;
      J[0*Q + col] = k.dmdot;
      J[1*Q + col] = k.dmdot*Usig + k.dp12*n[0];
      J[2*Q + col] = k.dmdot*Vsig + k.dp12*n[1];
      if (D == 3) J[3*Q + col] = k.dmdot*Wsig + k.dp12*n[2];
      J[(Q-1)*Q + col] = k.dmdot*Hsig + mdot12*(sigm*dhL + Onesigm*dhR);
    }
  }

  // ---- Lax-Friedrichs and HLL (a13), Euler variable sets ---------------------
  // Euler2DVarSet.cxx:47-66 / Euler3DVarSet.cxx:64-84 computeEigenValues
  static void eigenValues(const PData& d, const double* n, double* ev) {
    double un;
    if (D == 2) un = d.v[0]*n[0] + d.v[1]*n[1];
    else        un = d.v[0]*n[0] + d.v[1]*n[1] + d.v[2]*n[2];
    for (int i = 0; i < D; ++i) ev[i] = un;
    ev[D] = un + d.A;
    ev[D+1] = un - d.A;
  }
  // LaxFriedFlux.cxx:79-118 and HLLFlux.cxx:73-133, flux per unit area; uL/uR are the reconstructed UPDATE states
  static bool centredFlux(const Config& c, const double* uL, const double* uR, const PData& pL, const PData& pR,
                          const double* n, double* result) {
    double fL[Q], fR[Q], evL[Q], evR[Q], sL[Q], sR[Q];
    convFlux(pR, n, fR); convFlux(pL, n, fL);
    eigenValues(pR, n, evR); eigenValues(pL, n, evL);
    toCons(c, uL, sL); toCons(c, uR, sR);
    if (c.fluxSplitter == FS_CENTRED) {          // CentredFlux.cxx:62-64: _sumFlux = fR; _sumFlux += fL; result = 0.5*_sumFlux
      for (int i = 0; i < Q; ++i) { const double sumFlux = fR[i] + fL[i]; result[i] = 0.5*sumFlux; }
      return true;
    }
    if (c.fluxSplitter == FS_LAXFRIED) {
      double aR = 0.0, aL = 0.0;
      for (int i = 0; i < Q; ++i) { aR = std::max(aR, std::abs(evR[i])); aL = std::max(aL, std::abs(evL[i])); }
      const double a = std::max(aR, aL);
      const double aDiff = a*c.diffRedCoeff;
      for (int i = 0; i < Q; ++i) {
        const double sumFlux = fR[i] + fL[i];
        result[i] = 0.5*(sumFlux - aDiff*(sR[i] - sL[i]));
      }
    } else {
      double aR = 0.0, aL = 0.0;
      for (int i = 0; i < Q; ++i) { aR = std::max(aR, evR[i]); aL = std::max(aL, evL[i]); }
      const double amax = std::max(0.0, std::max(aR, aL));
      aR = 0.0; aL = 0.0;
      for (int i = 0; i < Q; ++i) { aR = std::min(aR, evR[i]); aL = std::min(aL, evL[i]); }
      const double amin = std::min(0.0, std::min(aR, aL));
      const double c1 = (amax + amin)/(amax - amin);
      double c2;
      if (!c.hllUseAlpha) c2 = amax*amin/(amax - amin);
      else {
        const double alpha = std::max(std::abs(amin), std::abs(amax))/(amax - amin);
        c2 = alpha*2.0*amax*amin/(amax - amin);
      }
      for (int i = 0; i < Q; ++i)
        result[i] = 0.5*(fL[i] + fR[i] - c1*(fR[i] - fL[i]) + c2*(sR[i] - sL[i]));
    }
    return true;
  }

  // ---- Navier-Stokes -------------------------------------------------------
  // SimplerSutherlandDynViscosity.cxx:104-107, SutherlandDynViscosity.cxx:107-110
  static double dynViscosityDim(const Config& c, double /*pdim*/, double Tdim) {
    switch (c.viscLaw) {
      case VISC_SIMPLER_SUTHERLAND: return c.viscRef*cr_pow15(Tdim)/(c.viscTRef + Tdim);
      case VISC_SUTHERLAND: return c.viscRef*(c.viscTRef + c.viscSuthConst)/(Tdim + c.viscSuthConst)*cr_pow15(Tdim/c.viscTRef);
      default: return c.fixedVisc;
    }
  }
  // NavierStokes2DPuvt / NavierStokes3DPvt::getDynViscosity (NavierStokes3DPvt.cxx:103-109)
  static double dynViscosity(const Config& c, const double* avState) {
    const double Tdim = c.tempRef*avState[TID];
    const double pdim = c.pRef*(avState[0] + c.pOffset);
    return dynViscosityDim(c, pdim, Tdim)/c.muRef;
  }
  // NavierStokes3DPvt.hh:92-99
  static double thermConductivity(const Config& c, double mu) {
    if (c.isAdimensional) return mu;
    return mu*c.cpOverPrandtl;
  }
  // NavierStokes3DPvt.cxx:113-117
  static double nsDensity(const Config& c, const double* avState) {
    const double p = avState[0] + c.pOffset;
    return p/(c.R*avState[TID]);
  }

  // NavierStokesVarSet.cxx:70-129 + NavierStokes{2,3}DVarSet::getFlux (2D: NavierStokes2DVarSet.cxx:93-115,
  // 3D: NavierStokes3DVarSet.cxx:25-45).  grad[v][ix]; returns flux per unit area in f[0..Q).
  static void nsFluxUnit(const Config& c, const double* avState, const double (*grad)[3],
                         const double* n, double mu, double lambda, double* f) {
    const double twoThird = 2./3.;
    double divTerm = 0.0;
    if (D == 3) divTerm = grad[3][2];
    const double* gradU = grad[1];
    const double* gradV = grad[2];
    const double twoThirdDivV = twoThird*(gradU[0] + gradV[1] + divTerm);
    const double coeffTauMu = c.coeffTau*mu;
    double tau[3][3];
    tau[0][0] = coeffTauMu*(2.*gradU[0] - twoThirdDivV);
    tau[0][1] = tau[1][0] = coeffTauMu*(gradU[1] + gradV[0]);
    tau[1][1] = coeffTauMu*(2.*gradV[1] - twoThirdDivV);
    if (D == 3) {
      const double* gradW = grad[3];
      tau[0][2] = tau[2][0] = coeffTauMu*(gradU[2] + gradW[0]);
      tau[1][2] = tau[2][1] = coeffTauMu*(gradV[2] + gradW[1]);
      tau[2][2] = coeffTauMu*(2.*gradW[2] - twoThirdDivV);
    }
    double ip = 0.;
    for (int i = 0; i < D; ++i) ip += grad[TID][i]*n[i];
    const double qFlux = -c.coeffQ*lambda*ip;
    f[0] = 0.;
    for (int i = 0; i < D; ++i) {
      double r = 0.;
      for (int k = 0; k < D; ++k) r += tau[i][k]*n[k];
      f[1+i] = r;
    }
    if (D == 3) {
      double r = 0.;
      for (int i = 0; i < D; ++i) r += f[1+i]*avState[1+i];
      f[TID] = r - qFlux;
    } else {
      f[TID] = (tau[0][0]*avState[1] + tau[0][1]*avState[2])*n[0] +
               (tau[0][1]*avState[1] + tau[1][1]*avState[2])*n[1] - qFlux;
    }
  }

  // a19: analytic Jacobian of the diffusive flux with respect to the state of one side, given d grad / d state of that
  // side (gj, from DerivativeComputer::getGradientsJacob): NavierStokes2DPuvt::computeFluxJacobian
  // (NavierStokes2DPuvt.cxx:126-161) / NavierStokes3DPvt::computeFluxJacobian (NavierStokes3DPvt.cxx:129-193).
  // Only the velocity / temperature entries are written, the rest of J stays zero (the matrices are sized once and never
  // touched elsewhere, DiffusiveVarSet / ComputeDiffusiveFlux::setup).  The 4/3 and 2/3 of the reference are kept as they
  // are (they are swapped with respect to the stress tensor of getFlux).
  static void nsFluxJacobian(const Config& c, const double* avState, const double* gj, const double* n,
                             double mu, double tc, double* J) {
    for (int k = 0; k < Q*Q; ++k) J[k] = 0.;
    const double fourThird = 4./3.;
    const double twoThird = 2./3.;
    const double nx = n[0], ny = n[1];
    const double coeffTau = c.coeffTau, coeffQ = c.coeffQ;
    if (D == 2) {
      const double tauxx_uL = twoThird*mu*gj[0];
      const double tauxx_vL = -fourThird*mu*gj[1];
      const double tauxy_uL = mu*gj[1];
      const double tauxy_vL = mu*gj[0];
      const double tauyy_uL = -fourThird*mu*gj[0];
      const double tauyy_vL = twoThird*mu*gj[1];
      const double qx_T = -tc*gj[0];
      const double qy_T = -tc*gj[1];
      const double avU = avState[1], avV = avState[2];
      J[1*Q+1] = coeffTau*(tauxx_uL*nx + tauxy_uL*ny);
      J[1*Q+2] = coeffTau*(tauxx_vL*nx + tauxy_vL*ny);
      J[2*Q+1] = coeffTau*(tauxy_uL*nx + tauyy_uL*ny);
      J[2*Q+2] = coeffTau*(tauxy_vL*nx + tauyy_vL*ny);
      J[3*Q+1] = J[1*Q+1]*avU + J[2*Q+1]*avV;
      J[3*Q+2] = J[1*Q+2]*avU + J[2*Q+2]*avV;
      J[3*Q+3] = -coeffQ*(qx_T*nx + qy_T*ny);
    } else {
      const double nz = n[D-1];
      const double gz = gj[D-1];
      const double tauxx_uL = twoThird*mu*gj[0];
      const double tauxx_vL = -fourThird*mu*gj[1];
      const double tauxx_wL = -fourThird*mu*gz;
      const double tauyy_uL = -fourThird*mu*gj[0];
      const double tauyy_vL = twoThird*mu*gj[1];
      const double tauyy_wL = -fourThird*mu*gz;
      const double tauzz_uL = -fourThird*mu*gj[0];
      const double tauzz_vL = -fourThird*mu*gj[1];
      const double tauzz_wL = twoThird*mu*gz;
      const double tauxy_uL = mu*gj[1];
      const double tauxy_vL = mu*gj[0];
      const double tauxz_uL = mu*gz;
      const double tauxz_wL = mu*gj[0];
      const double tauyz_vL = mu*gz;
      const double tauyz_wL = mu*gj[1];
      const double qx_T = -tc*gj[0];
      const double qy_T = -tc*gj[1];
      const double qz_T = -tc*gz;
      const double avU = avState[1], avV = avState[2], avW = avState[D];
      const int q4 = Q - 1;
      J[1*Q+1] = coeffTau*(tauxx_uL*nx + tauxy_uL*ny + tauxz_uL*nz);
      J[1*Q+2] = coeffTau*(tauxx_vL*nx + tauxy_vL*ny);
      J[1*Q+3] = coeffTau*(tauxx_wL*nx + tauxz_wL*nz);
      J[2*Q+1] = coeffTau*(tauxy_uL*nx + tauyy_uL*ny);
      J[2*Q+2] = coeffTau*(tauxy_vL*nx + tauyy_vL*ny + tauyz_vL*nz);
      J[2*Q+3] = coeffTau*(tauyy_wL*ny + tauyz_wL*nz);
      J[3*Q+1] = coeffTau*(tauxz_uL*nx + tauzz_uL*nz);
      J[3*Q+2] = coeffTau*(tauyz_vL*ny + tauzz_vL*nz);
      J[3*Q+3] = coeffTau*(tauxz_wL*nx + tauyz_wL*ny + tauzz_wL*nz);
      J[q4*Q+1] = J[1*Q+1]*avU + J[2*Q+1]*avV + J[3*Q+1]*avW;
      J[q4*Q+2] = J[1*Q+2]*avU + J[2*Q+2]*avV + J[3*Q+2]*avW;
      J[q4*Q+3] = J[1*Q+3]*avU + J[2*Q+3]*avV + J[3*Q+3]*avW;
      J[q4*Q+q4] = -coeffQ*(qx_T*nx + qy_T*ny + qz_T*nz);
    }
  }
};

// MathFunctions.hh:167-179
static inline void crossProd(const double* v1, const double* v2, double* r) {
  r[0] =  v1[1]*v2[2] - v1[2]*v2[1];
  r[1] = -v1[0]*v2[2] + v1[2]*v2[0];
  r[2] =  v1[0]*v2[1] - v1[1]*v2[0];
}
// MathFunctions.hh:92-104
template <int D> static inline double getDistance(const double* a, const double* b) {
  double dist = 0.;
  for (int i = 0; i < D; ++i) { const double diff = a[i] - b[i]; dist += diff*diff; }
  return std::sqrt(dist);
}
// NormalsCalculator.hh:209-231
static inline void faceTriagNormal(const double* ci, const double* cj, const double* ck, double* out) {
  double v1[3], v2[3], v3[3];
  for (int d = 0; d < 3; ++d) { v1[d] = cj[d] - ci[d]; v2[d] = ck[d] - cj[d]; }
  crossProd(v1, v2, v3);
  for (int d = 0; d < 3; ++d) out[d] = v3[d]*0.5;
}

} // namespace cfo

namespace cfo {

// ---------------------------------------------------------------------------
// MHD3DProjection in conservative variables (a13): rho, rho u, rho v, rho w, Bx, By, Bz, rho E, phi.
// plugins/MHD/MHD3DProjectionCons.cxx:540-574 (computePhysicalData), :1132-1148 (computeStateFromPhysicalData),
// plugins/MHD/MHD3DProjectionVarSet.cxx:42-78 (computeFlux), :398-436 (getMaxEigenValue), :499-548 (computeEigenValues);
// no background potential field (potentialBType unset: B0 = 0).  Only the centred splitters (LaxFried, HLL) apply.
struct MhdData { double rho, p, A, V, B, phi, v[3], b[3], x[3]; };   // x: data[XP YP ZP], the point the data belong to

struct MhdPhys {
  static constexpr int D = 3;
  static constexpr int Q = 9;
  static constexpr int TID = 7;
  typedef MhdData Data;

  static void computePhysicalData(const Config& c, const double* s, MhdData& d) {
    const double rho = s[0];
    const double u = s[1]/rho, v = s[2]/rho, w = s[3]/rho;
    const double Bx = s[4], By = s[5], Bz = s[6];
    const double V2 = u*u + v*v + w*w;
    const double B2 = Bx*Bx + By*By + Bz*Bz;
    d.v[0] = u; d.v[1] = v; d.v[2] = w;
    d.b[0] = Bx; d.b[1] = By; d.b[2] = Bz;
    d.B = std::sqrt(B2);
    d.rho = rho;
    d.V = std::sqrt(V2);
    d.p = (c.gamma - 1.)*(s[7] - 0.5*(rho*V2 + B2));
    d.phi = s[8];
    d.A = std::sqrt(c.gamma*d.p/rho);
    d.x[0] = 0.; d.x[1] = 0.; d.x[2] = 0.;   // data[XP YP ZP] = state.getCoordinates() (:570-573): set by the caller
  }
  static void stateFromPhysicalData(const Config& c, const MhdData& d, double* s) {
    const double rho = d.rho;
    s[0] = rho;
    s[1] = rho*d.v[0]; s[2] = rho*d.v[1]; s[3] = rho*d.v[2];
    s[4] = d.b[0]; s[5] = d.b[1]; s[6] = d.b[2];
    const double B = d.B;
    s[7] = d.p/(c.gamma - 1.0) + 0.5*(rho*d.V*d.V + B*B);
    s[8] = d.phi;
  }
  static void farFieldGhost(const Config&, const double*, const double*, const double*, double*) {
    throw std::runtime_error("FarFieldEuler on MHD3DProjection");
  }
  static void farFieldCharGhost(const Config&, const double*, const double*, const double*, double*) {
    throw std::runtime_error("FarFieldEulerChar on MHD3DProjection");
  }
  // FiniteVolumeMHD/MirrorMHD3DProjection.cxx:56-96
  static void mirrorGhost(const Config& c, const double* nr, const double* inner, double* ghost) {
    double nx = nr[0], ny = nr[1], nz = nr[2];
    const double invFaceLength = 1./std::sqrt(nx*nx + ny*ny + nz*nz);
    nx *= invFaceLength; ny *= invFaceLength; nz *= invFaceLength;
    MhdData di, dg;
    computePhysicalData(c, inner, di);
    const double vn = di.v[0]*nx + di.v[1]*ny + di.v[2]*nz;
    const double bn = di.b[0]*nx + di.b[1]*ny + di.b[2]*nz;
    dg.rho = di.rho;
    dg.v[0] = di.v[0] - 2.0*vn*nx; dg.v[1] = di.v[1] - 2.0*vn*ny; dg.v[2] = di.v[2] - 2.0*vn*nz;
    dg.b[0] = di.b[0] - 2.0*bn*nx; dg.b[1] = di.b[1] - 2.0*bn*ny; dg.b[2] = di.b[2] - 2.0*bn*nz;
    dg.V = di.V; dg.A = di.A; dg.p = di.p; dg.B = di.B; dg.phi = di.phi;
    stateFromPhysicalData(c, dg, ghost);
  }
  // FiniteVolumeMHD/SuperOutletMHD3DProjection.cxx:76-239.  coronalBC (the default, :45): the "_dimless" lines :99-226, the
  // state entries read as (rho, V, B, p, phi) whatever the update variables are; otherwise the inner physical data with
  // phi = refPhi through computeStateFromPhysicalData (:228-238).  xI / xG: inner and ghost state coordinates.
  static void superOutletGhost(const Config& c, bool coronalBC, double refPhi, const double* xIc, const double* xGc,
                               const double* inner, double* ghost) {
    if (!coronalBC) {
      MhdData dg;
      computePhysicalData(c, inner, dg);
      dg.phi = refPhi;
      stateFromPhysicalData(c, dg, ghost);
      return;
    }
    const double xI = xIc[0], yI = xIc[1], zI = xIc[2];
    double sq = 0.; sq += xI*xI; sq += yI*yI; sq += zI*zI;            // RealVector::norm2 (ArrayT.hh:93-124)
    const double rI = std::sqrt(sq);
    const double rhoI = std::sqrt(xI*xI + yI*yI);
    const double xG = xGc[0], yG = xGc[1], zG = xGc[2];
    sq = 0.; sq += xG*xG; sq += yG*yG; sq += zG*zG;
    const double rG = std::sqrt(sq);
    const double rhoG = std::sqrt(xG*xG + yG*yG);
    const double densityI = inner[0];
    const double VxI = inner[1], VyI = inner[2], VzI = inner[3];
    const double BxI = inner[4], ByI = inner[5], BzI = inner[6];
    const double BrI = xI/rI*BxI + yI/rI*ByI + zI/rI*BzI;
    const double BthetaI = xI*zI/(rhoI*rI)*BxI + yI*zI/(rhoI*rI)*ByI - rhoI/rI*BzI;
    const double BphiI = -yI/rhoI*BxI + xI/rhoI*ByI;
    const double VrI = xI/rI*VxI + yI/rI*VyI + zI/rI*VzI;
    const double VthetaI = xI*zI/(rhoI*rI)*VxI + yI*zI/(rhoI*rI)*VyI - rhoI/rI*VzI;
    const double VphiI = -yI/rhoI*VxI + xI/rhoI*VyI;
    const double BrG = rI*rI/(rG*rG)*BrI;
    const double BthetaG = BthetaI;
    const double BphiG = (rI/rG)*BphiI;
    const double densityG = rI*rI*rI*densityI/(rG*rG*rG);
    ghost[0] = densityG;
    const double VrG = rI*rI*VrI*densityI/(rG*rG*densityG);
    const double VthetaG = densityI*VthetaI/densityG;
    const double VphiG = (rI/rG)*VphiI;
    const double BxG = xG/rG*BrG + xG*zG/(rhoG*rG)*BthetaG - yG/rhoG*BphiG;
    const double ByG = yG/rG*BrG + yG*zG/(rhoG*rG)*BthetaG + xG/rhoG*BphiG;
    const double BzG = zG/rG*BrG - rhoG/rG*BthetaG;
    const double VxG = xG/rG*VrG + xG*zG/(rhoG*rG)*VthetaG - yG/rhoG*VphiG;
    const double VyG = yG/rG*VrG + yG*zG/(rhoG*rG)*VthetaG + xG/rhoG*VphiG;
    const double VzG = zG/rG*VrG - rhoG/rG*VthetaG;
    ghost[4] = BxG; ghost[5] = ByG; ghost[6] = BzG;
    ghost[1] = VxG; ghost[2] = VyG; ghost[3] = VzG;
    ghost[7] = inner[7]*rI*rI*rI/(rG*rG*rG);
    ghost[8] = -inner[8];
  }
  static void convFlux(const Config& c, const MhdData& d, const double* n, double* f) {
    const double nx = n[0], ny = n[1], nz = n[2];
    const double rho = d.rho, u = d.v[0], v = d.v[1], w = d.v[2], Bx = d.b[0], By = d.b[1], Bz = d.b[2];
    const double Vn = u*nx + v*ny + w*nz;
    const double Bn = Bx*nx + By*ny + Bz*nz;
    const double B2 = Bx*Bx + By*By + Bz*Bz;
    const double VdotB = u*Bx + v*By + w*Bz;
    const double p = d.p, phi = d.phi;
    const double P = p + 0.5*B2;
    const double E = p/(c.gamma - 1.0) + 0.5*(rho*(u*u + v*v + w*w) + B2);
    const double refSpeedSq = c.refSpeed*c.refSpeed;
    f[0] = Vn*rho;
    f[1] = Vn*rho*u - Bn*Bx + P*nx;
    f[2] = Vn*rho*v - Bn*By + P*ny;
    f[3] = Vn*rho*w - Bn*Bz + P*nz;
    f[4] = (v*Bx - By*u)*ny + (w*Bx - Bz*u)*nz + phi*nx;
    f[5] = (u*By - Bx*v)*nx + (w*By - Bz*v)*nz + phi*ny;
    f[6] = (u*Bz - Bx*w)*nx + (v*Bz - By*w)*ny + phi*nz;
    f[7] = Vn*(E + P) - Bn*VdotB;
    f[8] = refSpeedSq*Bn;
  }
  // MHD3DProjectionVarSet::computeMagneticDipole (MHD3DProjectionVarSet.cxx:479-495); B0 = 0 without potentialBType
  static void backgroundB(const Config& c, const double* x, double* B0) {
    B0[0] = 0.0; B0[1] = 0.0; B0[2] = 0.0;
    if (c.mhdPotentialBType == 1) {
      const double mX = c.mhdDipole[0], mY = c.mhdDipole[1], mZ = c.mhdDipole[2];
      const double xCoord = x[0], yCoord = x[1], zCoord = x[2];
      const double r = std::sqrt(xCoord*xCoord + yCoord*yCoord + zCoord*zCoord);
      const double rCubeInv = 1.0/(r*r*r);
      const double rSqInv = 1.0/(r*r);
      const double mDotr = mX*xCoord + mY*yCoord + mZ*zCoord;
      const double coeff = 3.0*rSqInv*mDotr;
      B0[0] = rCubeInv*(coeff*xCoord - mX);
      B0[1] = rCubeInv*(coeff*yCoord - mY);
      B0[2] = rCubeInv*(coeff*zCoord - mZ);
    }
  }
  static void setCoordinates(MhdData& d, const double* x) { d.x[0] = x[0]; d.x[1] = x[1]; d.x[2] = x[2]; }
  static void eigenValues(const Config& c, const MhdData& d, const double* n, double* ev) {
    double B0[3];
    backgroundB(c, d.x, B0);
    const double Bx = d.b[0] + B0[0], By = d.b[1] + B0[1], Bz = d.b[2] + B0[2];    // B1 + B0
    const double invRho = 1./d.rho;
    const double sqB = Bx*Bx + By*By + Bz*Bz;
    const double p = d.p;
    const double Vn = d.v[0]*n[0] + d.v[1]*n[1] + d.v[2]*n[2];
    const double Bn = Bx*n[0] + By*n[1] + Bz*n[2];
    const double sqrbar = std::sqrt(d.rho);
    const double astar2 = (c.gamma*p + sqB)*invRho;
    const double astarb = std::sqrt(astar2*astar2 - 4.0*c.gamma*p*Bn*Bn*invRho*invRho);
    const double cf2 = 0.5*(astar2 + astarb);
    const double cs2 = 0.5*(astar2 - astarb);
    const double ca = std::abs(Bn)/sqrbar;
    const double cf = std::sqrt(cf2);
    const double cs = std::sqrt(cs2);
    ev[0] = Vn - cf; ev[1] = Vn - ca; ev[2] = Vn - cs; ev[3] = Vn;
    ev[4] = c.refSpeed; ev[5] = -c.refSpeed;
    ev[6] = Vn + cs; ev[7] = Vn + ca; ev[8] = Vn + cf;
  }
  // MHD3DProjectionVarSet.cxx:398-436 (what the RIGHT side of the update coefficient uses; the left side uses
  // _leftEv.max(): LaxFriedFlux.cxx:131, HLLFlux.cxx:141)
  static double maxEigenValue(const Config& c, const MhdData& d, const double* n) {
    double B0[3];
    backgroundB(c, d.x, B0);
    const double Bx = d.b[0] + B0[0], By = d.b[1] + B0[1], Bz = d.b[2] + B0[2];
    const double invRho = 1./d.rho;
    const double sqB = Bx*Bx + By*By + Bz*Bz;
    const double p = d.p;
    const double Vn = d.v[0]*n[0] + d.v[1]*n[1] + d.v[2]*n[2];
    const double Bn = Bx*n[0] + By*n[1] + Bz*n[2];
    const double astar2 = (c.gamma*p + sqB)*invRho;
    const double astarb = std::sqrt(astar2*astar2 - 4.0*c.gamma*p*Bn*Bn*invRho*invRho);
    const double cf2 = 0.5*(astar2 + astarb);
    const double cf = std::sqrt(cf2);
    return std::max(c.refSpeed, (Vn + cf));
  }
  static double evMaxOfVector(const Config& c, const MhdData& d, const double* n) {
    double ev[Q];
    eigenValues(c, d, n, ev);
    double emax = ev[0];
    for (int i = 1; i < Q; ++i) emax = std::max(emax, ev[i]);
    return emax;
  }
  static void toCons(const Config&, const double* s, double* r) { for (int i = 0; i < Q; ++i) r[i] = s[i]; }
  static void updateToSolutionMatrix(const Config&, const double*, double* M) {      // update = solution = Cons
    for (int i = 0; i < Q*Q; ++i) M[i] = 0.;
    for (int i = 0; i < Q; ++i) M[i*Q + i] = 1.;
  }

  // LaxFriedFlux.cxx:79-118 and HLLFlux.cxx:73-133 on the MHD variable set (update = solution = Cons)
  static bool centredFlux(const Config& c, const double* uL, const double* uR, const MhdData& pL, const MhdData& pR,
                          const double* n, double* result) {
    double fL[Q], fR[Q], evL[Q], evR[Q];
    convFlux(c, pR, n, fR); convFlux(c, pL, n, fL);
    eigenValues(c, pR, n, evR); eigenValues(c, pL, n, evL);
    if (c.fluxSplitter == FS_CENTRED) {          // CentredFlux.cxx:62-64
      for (int i = 0; i < Q; ++i) { const double sumFlux = fR[i] + fL[i]; result[i] = 0.5*sumFlux; }
      return true;
    }
    if (c.fluxSplitter == FS_LAXFRIED) {
      double aR = 0.0, aL = 0.0;
      for (int i = 0; i < Q; ++i) { aR = std::max(aR, std::abs(evR[i])); aL = std::max(aL, std::abs(evL[i])); }
      const double a = std::max(aR, aL);
      const double aDiff = a*c.diffRedCoeff;
      for (int i = 0; i < Q; ++i) {
        const double sumFlux = fR[i] + fL[i];
        result[i] = 0.5*(sumFlux - aDiff*(uR[i] - uL[i]));
      }
    } else {
      double aR = 0.0, aL = 0.0;
      for (int i = 0; i < Q; ++i) { aR = std::max(aR, evR[i]); aL = std::max(aL, evL[i]); }
      const double amax = std::max(0.0, std::max(aR, aL));
      aR = 0.0; aL = 0.0;
      for (int i = 0; i < Q; ++i) { aR = std::min(aR, evR[i]); aL = std::min(aL, evL[i]); }
      const double amin = std::min(0.0, std::min(aR, aL));
      const double c1 = (amax + amin)/(amax - amin);
      double c2;
      if (!c.hllUseAlpha) c2 = amax*amin/(amax - amin);
      else {
        const double alpha = std::max(std::abs(amin), std::abs(amax))/(amax - amin);
        c2 = alpha*2.0*amax*amin/(amax - amin);
      }
      for (int i = 0; i < Q; ++i)
        result[i] = 0.5*(fL[i] + fR[i] - c1*(fR[i] - fL[i]) + c2*(uR[i] - uL[i]));
    }
    return true;
  }
  // not available for this model (the configuration is rejected before the solver runs)
  static bool roeFlux(const Config&, const double*, const double*, const MhdData&, const MhdData&, const double*, double*, double = 0.) { return false; }
  static void eigenValues(const MhdData&, const double*, double*) {}     // RoeSA is not offered on MHD3DProjection
  static bool ausmFlux(const Config&, const MhdData&, const MhdData&, const double*, double*) { return false; }
  static void ausmPlusUpJacobCons(const Config&, const MhdData&, const MhdData&, const double*, int, double*) {}
  static void ausmPlusUpJacobPvt(const Config&, const MhdData&, const MhdData&, const double*, int, double*) {}
  static double dynViscosity(const Config&, const double*) { return 0.; }
  static double thermConductivity(const Config&, double) { return 0.; }
  static double nsDensity(const Config&, const double*) { return 1.; }
  static void nsFluxUnit(const Config&, const double*, const double (*)[3], const double*, double, double, double* f) {
    for (int i = 0; i < Q; ++i) f[i] = 0.;
  }  static void nsFluxJacobian(const Config&, const double*, const double*, const double*, double, double, double* J) {
    for (int k = 0; k < Q*Q; ++k) J[k] = 0.;
  }
};

} // namespace cfo
