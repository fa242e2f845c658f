"""CPU tier: the oracle's restatement of the analytic AUSM+up flux Jacobian (a19, AUSMPlusUpFlux.ci:675-1203) against
central differences of the flux that routine differentiates (its own sigmoid-blended mass / pressure flux, written
independently below in numpy), plus the three places where the reference's formulas are NOT the derivative and are
replicated as they are:
  1. d(pu)/d(qn) is multiplied by da12 instead of a12            (AUSMPlusUpFlux.ci:930, 986, ...);
  2. the rho column adds +prho2a2*drho12 (the derivative of 1/rho12 has the opposite sign)   (:916);
  3. entry (rho w, rho w) uses the rho v column's d(mdot12)      (:1141).
Agreement to ~1e-8 (finite-difference noise) pins the transcription of the 500 formula lines."""
import ctypes as C

import numpy as np
import pytest

from coolfluid_b200 import config as B

g = 1.4


def sig(x,y,K): return 0.5*(1.0-(1.0+K)*(x-y)/(-abs(x-y)-K))
def smooth(uL,uR,n,cfg,dim):
    """The flux the Jacobian routine differentiates (AUSMPlusUpFlux.ci:675-790): returns F, a12, qnL, qnR, extras"""
    K=cfg.coeffSigmoid; beta=cfg.ausmBeta; Kp=cfg.ausmKp; Ku=cfg.ausmKu; sg=cfg.ausmSigma; mInf=cfg.machInf
    def prim(u):
        rho=u[0]; v=u[1:1+dim]/rho; p=(g-1)*(u[-1]-0.5*rho*(v@v)); H=(u[-1]+p)/rho
        return rho,v,p,H
    rL,vL,pL,HL=prim(uL); rR,vR,pR,HR=prim(uR)
    qL=vL@n; qR=vR@n
    acL=np.sqrt(2*(g-1)/(g+1)*HL); acR=np.sqrt(2*(g-1)/(g+1)*HR)
    sL=sig(acL,qL,K); sR=sig(acR,-qR,K)
    asL=sL*acL+(1-sL)*qL; asR=sR*acR-(1-sR)*qR
    ahL=acL*acL/max(acL,qL); ahR=acR*acR/max(acR,-qR)
    s12=sig(ahL,ahR,K)
    a12=(1-s12)*acL*acL/asL+s12*acR*acR/asR
    r12=0.5*(rL+rR); ML=qL/a12; MR=qR/a12
    MB=0.5*(ML*ML+MR*MR); s1=sig(MB,mInf*mInf,K); s2=s1*MB+(1-s1)*mInf*mInf
    z=sig(1.0,s2,K); m0=np.sqrt((1-z)+z*s2); fa=m0*(2-m0)
    mps=sig(1-sg*MB,0.0,K)
    m1p=lambda m:0.5*(m+abs(m)); m1m=lambda m:0.5*(m-abs(m)); m2p=lambda m:0.25*(m+1)**2; m2m=lambda m:-0.25*(m-1)**2
    M4p=m1p(ML) if abs(ML)>=1 else m2p(ML)*(1-16*beta*m2m(ML))
    M4m=m1m(MR) if abs(MR)>=1 else m2m(MR)*(1+16*beta*m2p(MR))
    mP=-Kp*mps*(pR-pL)/(r12*a12*a12)
    M12=M4p+M4m+mP
    sr=sig(M12,0.0,K); rs=sr*rL+(1-sr)*rR; md=M12*a12*rs
    sm=sig(md,0.0,K)
    al=(3/16)*(-4+5*fa*fa)
    P5p=m1p(ML)/ML if abs(ML)>=1 else m2p(ML)*((2-ML)-16*al*ML*m2m(ML))
    P5m=m1m(MR)/MR if abs(MR)>=1 else m2m(MR)*((-2-MR)+16*al*MR*m2p(MR))
    pu=0.5*Ku*P5p*P5m*r12*fa*a12*(qR-qL)
    p12=P5p*pL+P5m*pR+pu
    F=np.zeros(dim+2); F[0]=md
    for i in range(dim): F[1+i]=md*(sm*vL[i]+(1-sm)*vR[i])+p12*n[i]
    F[-1]=md*(sm*HL+(1-sm)*HR)
    return F,dict(a12=a12,qL=qL,qR=qR,pudqn=P5p*P5m*r12*fa,sm=sm,md=md,vL=vL,vR=vR,HL=HL,HR=HR,dMpCon=-Kp*mps,prho2a2=(pR-pL)/(r12*r12*a12*a12),a12rs=a12*rs)


def _pd(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def _state(rho, vel, p):
    return np.array([rho] + [rho*v for v in vel] + [p/0.4 + 0.5*rho*sum(v*v for v in vel)])


@pytest.mark.parametrize("dim", [2, 3])
@pytest.mark.parametrize("mach_inf,scale", [(0.5, 1.0), (2.0, 1.0), (0.5, 3.0), (0.5, -3.0)])
def test_analytic_jacobian_is_the_derivative_plus_the_reference_quirks(dim, mach_inf, scale, oracle):
    L = oracle.lib(False)
    q = dim + 2
    cfg = B.make_config(dim, flux=B.FS_AUSMPLUSUP, jacobian="analytic", mach_inf=mach_inf, coeff_sigmoid=1e-8)
    rng = np.random.RandomState(dim)
    n = rng.standard_normal(dim); n /= np.linalg.norm(n)
    uL = _state(1.0, [scale*x for x in [0.5, 0.1, 0.05][:dim]], 1.0)
    uR = _state(0.9, [scale*x for x in [0.45, 0.12, 0.02][:dim]], 0.95)
    checked = 0
    for side in (0, 1):
        J = np.zeros(q*q)
        L.cfo_ausm_jacob_cons(C.byref(cfg), _pd(uL), _pd(uR), _pd(n), side, _pd(J))
        J = J.reshape(q, q)
        F0, e0 = smooth(uL, uR, n, cfg, dim)
        E = np.zeros((q, q)); da12 = np.zeros(q); dq = np.zeros(q); dmd = np.zeros(q)
        for j in range(q):
            h = 1e-6*max(1, abs([uL, uR][side][j]))
            a = [uL.copy(), uR.copy()]; b = [uL.copy(), uR.copy()]
            a[side][j] += h; b[side][j] -= h
            Fa, ea = smooth(a[0], a[1], n, cfg, dim); Fb, eb = smooth(b[0], b[1], n, cfg, dim)
            E[:, j] = (Fa - Fb)/(2*h)
            da12[j] = (ea['a12'] - eb['a12'])/(2*h)
            dq[j] = ((ea['qR'] - ea['qL']) - (eb['qR'] - eb['qL']))/(2*h)
            dmd[j] = (ea['md'] - eb['md'])/(2*h)
        sm = e0['sm']
        for j in range(q):                                   # quirk 1
            corr = 0.5*cfg.ausmKu*e0['pudqn']*(da12[j] - e0['a12'])*dq[j]
            for i in range(dim):
                E[1+i, j] += corr*n[i]
        dlt = e0['dMpCon']*2*e0['prho2a2']*0.5*e0['a12rs']   # quirk 2
        E[0, 0] += dlt
        for i in range(dim):
            E[1+i, 0] += dlt*(sm*e0['vL'][i] + (1 - sm)*e0['vR'][i])
        E[-1, 0] += dlt*(sm*e0['HL'] + (1 - sm)*e0['HR'])
        if dim == 3:                                         # quirk 3
            E[3, 3] += (dmd[2] - dmd[3])*(sm*e0['vL'][2] + (1 - sm)*e0['vR'][2])
        if np.abs(E).max() > 1e-3:
            assert np.abs(J - E).max() <= 1e-6*np.abs(E).max()
            checked += 1
        else:                                                # fully upwinded from the other side: zero block
            assert np.abs(J).max() <= 1e-6
    assert checked >= 1


# ---- Puvt / Pvt variables (ComputedFluxdPvt, AUSMPlusUpFlux.ci:160-669) ------------------------------------------------
def smooth_pvt(prim, n, cfg, dim, frozen=None):
    """The sigmoid-blended flux ComputedFluxdPvt differentiates, as a function of the INDEPENDENT per-side quantities
    prim = [(rho, v, H, p) left, (rho, v, H, p) right].  `frozen` (base values) reproduces two simplifications of the
    reference's derivative so that central differences of this function ARE its formulas:
      * the factor (1 - sigma MBarSq) of mP is not differentiated and is left out of d mP (dMpCon, :220): mP is
        linearised as -Kp mPSig (c1 X0 + (X - X0)), X = (pR - pL)/(fa rho12 a12^2);
      * d fa carries a spurious factor (1 - sigma MBarSq) (dmZeroCon, :288): fa is replaced by fa0 + c1 (fa - fa0)."""
    K=cfg.coeffSigmoid; beta=cfg.ausmBeta; Kp=cfg.ausmKp; Ku=cfg.ausmKu; sg=cfg.ausmSigma; mInf=cfg.machInf
    (rL,vL,HL,pL),(rR,vR,HR,pR)=prim
    qL=vL@n; qR=vR@n
    acL=np.sqrt(2*(g-1)/(g+1)*HL); acR=np.sqrt(2*(g-1)/(g+1)*HR)
    sL=sig(acL,qL,K); sR=sig(acR,-qR,K)
    asL=sL*acL+(1-sL)*qL; asR=sR*acR-(1-sR)*qR
    ahL=acL*acL/max(acL,qL); ahR=acR*acR/max(acR,-qR)
    s12=sig(ahL,ahR,K)
    a12=(1-s12)*acL*acL/asL+s12*acR*acR/asR
    r12=0.5*(rL+rR); ML=qL/a12; MR=qR/a12
    MB=0.5*(ML*ML+MR*MR); s1=sig(MB,mInf*mInf,K); s2=s1*MB+(1-s1)*mInf*mInf
    z=sig(1.0,s2,K); m0=np.sqrt((1-z)+z*s2); fa=m0*(2-m0)
    c1=1-sg*MB
    if frozen is not None:
        fa=frozen['fa']+frozen['c1']*(fa-frozen['fa'])
    mps=sig(c1,0.0,K)
    m1p=lambda m:0.5*(m+abs(m)); m1m=lambda m:0.5*(m-abs(m)); m2p=lambda m:0.25*(m+1)**2; m2m=lambda m:-0.25*(m-1)**2
    M4p=m1p(ML) if abs(ML)>=1 else m2p(ML)*(1-16*beta*m2m(ML))
    M4m=m1m(MR) if abs(MR)>=1 else m2m(MR)*(1+16*beta*m2p(MR))
    X=(pR-pL)/(fa*r12*a12*a12)
    mP=-Kp*mps*c1*X if frozen is None else -Kp*mps*(frozen['c1']*frozen['X']+(X-frozen['X']))
    M12=M4p+M4m+mP
    sr=sig(M12,0.0,K); rs=sr*rL+(1-sr)*rR; md=M12*a12*rs
    sm=sig(md,0.0,K)
    al=(3/16)*(-4+5*fa*fa)
    P5p=m1p(ML)/ML if abs(ML)>=1 else m2p(ML)*((2-ML)-16*al*ML*m2m(ML))
    P5m=m1m(MR)/MR if abs(MR)>=1 else m2m(MR)*((-2-MR)+16*al*MR*m2p(MR))
    pu=0.5*Ku*P5p*P5m*r12*fa*a12*(qR-qL)
    p12=P5p*pL+P5m*pR+pu
    F=np.zeros(dim+2); F[0]=md
    for i in range(dim): F[1+i]=md*(sm*vL[i]+(1-sm)*vR[i])+p12*n[i]
    F[-1]=md*(sm*HL+(1-sm)*HR)
    return F,dict(a12=a12,dq=qR-qL,pudqn=P5p*P5m*r12*fa,sm=sm,md=md,fa=fa,c1=c1,X=X,dMpCon=-Kp*mps,
                  prho2a2=-(pR-pL)/(fa*r12*r12*a12*a12),a12rs=a12*rs,
                  vs=sm*vL+(1-sm)*vR,Hs=sm*HL+(1-sm)*HR)


@pytest.mark.parametrize("dim", [2, 3])
@pytest.mark.parametrize("mach_inf,scale,sigma", [(0.5, 1.0, 1.0), (0.1, 0.3, 1.0), (2.0, 1.0, 1.0), (0.5, 3.0, 1.0),
                                                  (0.5, -3.0, 1.0), (0.5, 1.0, 0.0)])
def test_analytic_jacobian_pvt_is_the_derivative_plus_the_reference_quirks(dim, mach_inf, scale, sigma, oracle):
    """ComputedFluxdPvt against central differences of smooth_pvt, chained with the reference's own seeds
      p column:  d rho = Mmass/(Rgas T), d H = -gamma p/((gamma-1) rho^2), d p = 1        (:323-328)
      u_m column: d qn = n_m, d H = u_m, d u_m = 1                                         (:397-402)
      T column:  d rho = -Mmass p/(Rgas T^2), d H = gamma/(rho (gamma-1))                  (:601-604)
    (air's dimensional Mmass and Rgas whatever the model's R), plus the places where its formulas are not that derivative:
      1. d(pu)/d(qn) is multiplied by da12 instead of a12                                  (:445);
      2. the T column leaves the d rho12 term out of d mP                                  (:632-634);
      3. the p column has no mdot12 dH term in the energy row                              (:393);
      4. the two simplifications reproduced by smooth_pvt(frozen=...)                      (:220, :288)."""
    L = oracle.lib(False)
    q = dim + 2
    Rair = 8.3145/28.965338E-3
    cfg = B.make_config(dim, update_var=B.UV_PVT, flux=B.FS_AUSMPLUSUP, jacobian="analytic", mach_inf=mach_inf,
                        coeff_sigmoid=1e-11, R=Rair)
    cfg.ausmSigma = sigma
    rng = np.random.RandomState(10 + dim)
    n = rng.standard_normal(dim); n /= np.linalg.norm(n)
    a0 = np.sqrt(g*Rair*300.0)
    sts = [np.array([101325.0] + [scale*a0*x for x in [0.5, 0.1, 0.05][:dim]] + [300.0]),
           np.array([96000.0] + [scale*a0*x for x in [0.45, 0.12, 0.02][:dim]] + [310.0])]
    def prim_of(s):
        rho = s[0]/(Rair*s[-1]); v = s[1:1+dim].copy()
        return [rho, v, g/(g - 1)*s[0]/rho + 0.5*(v@v), s[0]]
    prim0 = [prim_of(sts[0]), prim_of(sts[1])]
    F0, e0 = smooth_pvt(prim0, n, cfg, dim)
    # the mass and energy flux this derivative belongs to are those the splitter computes (its interface pressure is not:
    # "new pressure interface for consistency", AUSMPlusUpFlux.ci:264-267, has another velocity-diffusion term)
    Fref = np.zeros(q)
    L.cfo_ausm_flux(C.byref(cfg), _pd(sts[0]), _pd(sts[1]), _pd(n), _pd(Fref))
    assert abs(F0[0] - Fref[0]) <= 1e-6*abs(Fref[0]) and abs(F0[-1] - Fref[-1]) <= 1e-6*abs(Fref[-1])
    frozen = dict(fa=e0['fa'], c1=e0['c1'], X=e0['X'])
    Mmass, Rgas = 28.965338E-3, 8.3145

    def dG(side, what, m=None):
        """central difference of (F, a12, qnR - qnL) with respect to one independent quantity of one side"""
        x0 = prim0[side][what] if m is None else prim0[side][what][m]
        h = 1e-4*max(abs(x0), 1e-3*a0 if what == 1 else 1e-30)      # d/dH is multiplied by seeds of 3e5: keep the noise down
        out = []
        for sgn in (+1, -1):
            pr = [[r, v.copy(), H, p] for (r, v, H, p) in prim0]
            if m is None: pr[side][what] = x0 + sgn*h
            else: pr[side][what][m] = x0 + sgn*h
            out.append(smooth_pvt(pr, n, cfg, dim, frozen))
        (Fa, ea), (Fb, eb) = out
        return (Fa - Fb)/(2*h), (ea['a12'] - eb['a12'])/(2*h), (ea['dq'] - eb['dq'])/(2*h)

    checked = 0
    for side in (0, 1):
        J = np.zeros(q*q)
        L.cfo_ausm_jacob_pvt(C.byref(cfg), _pd(sts[0]), _pd(sts[1]), _pd(n), side, _pd(J))
        J = J.reshape(q, q)
        rho, v, H, p = prim0[side]; T = sts[side][-1]
        dFr, _, _ = dG(side, 0); dFH, dAH, _ = dG(side, 2); dFp, _, _ = dG(side, 3)
        E = np.zeros((q, q)); S = np.zeros((q, q))         # S: size of the terms summed (the p column cancels heavily)
        # p column
        dh = -g*p/((g - 1)*rho*rho)
        E[:, 0] = dFp + dFr*(Mmass/(Rgas*T)) + dFH*dh
        S[:, 0] = np.abs(dFp) + np.abs(dFr*(Mmass/(Rgas*T))) + np.abs(dFH*dh)
        E[-1, 0] -= e0['md']*(e0['sm'] if side == 0 else 1 - e0['sm'])*dh                     # quirk 3
        # velocity columns
        for m in range(dim):
            dFv, dAv, dqv = dG(side, 1, m)
            E[:, 1+m] = dFv + dFH*v[m]
            S[:, 1+m] = np.abs(dFv) + np.abs(dFH*v[m])
            da12 = dAv + dAH*v[m]
            corr = 0.5*cfg.ausmKu*e0['pudqn']*(da12 - e0['a12'])*dqv                          # quirk 1
            for i in range(dim):
                E[1+i, 1+m] += corr*n[i]
        # T column
        drho = -Mmass*p/(Rgas*T*T)
        E[:, -1] = dFr*drho + dFH*(g/(rho*(g - 1)))
        S[:, -1] = np.abs(dFr*drho) + np.abs(dFH*(g/(rho*(g - 1))))
        dlt = -e0['dMpCon']*e0['prho2a2']*(0.5*drho)*e0['a12rs']                              # quirk 2
        E[0, -1] += dlt
        for i in range(dim):
            E[1+i, -1] += dlt*e0['vs'][i]
        E[-1, -1] += dlt*e0['Hs']
        # compare column by column (the columns have different units)
        for j in range(q):
            ref = S[:, j].max()
            if ref > 0:
                assert np.abs(J[:, j] - E[:, j]).max() <= 1e-6*ref, (side, j)
                checked += 1
            else:
                assert not J[:, j].any()
    assert checked >= q


def test_analytic_diffusive_jacobian_is_the_derivative_of_a_laplacian_like_flux(oracle):
    """NavierStokes{2DPuvt,3DPvt}::computeFluxJacobian (a19): momentum rows are linear in gj with the reference's
    coefficients (2/3 on the diagonal component, -4/3 on the others: NOT the stress tensor of getFlux, kept as is), the
    energy row is (momentum rows).avVelocity and the conduction term coeffQ*lambda*(gj.n)."""
    L = oracle.lib(False)
    rng = np.random.RandomState(5)
    for dim in (2, 3):
        q = dim + 2
        cfg = B.make_config(dim, update_var=B.UV_PVT, diffusive=B.DIFF_NS, **cases_ns())
        n = rng.standard_normal(dim); n /= np.linalg.norm(n)
        gj = rng.standard_normal(dim); av = rng.standard_normal(q) + 2.0
        mu, tc = 0.37, 1.9
        J = np.zeros(q*q)
        L.cfo_ns_flux_jacob(C.byref(cfg), _pd(av), _pd(gj), _pd(n), C.c_double(mu), C.c_double(tc), _pd(J))
        J = J.reshape(q, q)
        assert not J[0].any() and not J[:, 0].any() and not J[1:1+dim, -1].any()
        # tau_ab,u_c of the reference: diagonal stress (a == b): 2/3 mu gj_c if c == a else -4/3 mu gj_c;
        # shear stress (a != b): mu gj_b if c == a, mu gj_a if c == b
        E = np.zeros((q, q))
        for a in range(dim):
            for c in range(dim):
                t = 0.0
                for b in range(dim):
                    if a == b:
                        coef = (2./3.)*mu*gj[c] if c == a else -(4./3.)*mu*gj[c]
                    else:
                        coef = mu*gj[b] if c == a else (mu*gj[a] if c == b else 0.0)
                    t += coef*n[b]
                E[1+a, 1+c] = cfg.coeffTau*t
        E[-1, 1:1+dim] = av[1:1+dim] @ E[1:1+dim, 1:1+dim]
        E[-1, -1] = cfg.coeffQ*tc*(gj @ n)
        assert np.abs(J - E).max() <= 1e-14*np.abs(E).max()


def cases_ns():
    import cases
    return cases.ns_kwargs()
