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
