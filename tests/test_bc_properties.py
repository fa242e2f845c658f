"""CPU tier: properties of the boundary commands and of the source term added late in round 2, checked on the ORACLE against
formulas written independently of it (numpy, from the physics the reference's comments state), so that oracle == product
parity is not the only thing that pins them.

* SuperOutletMHD3DProjection (coronalBC): "Continuous BCs for r^2*Br, Btheta, Bphi", "r^2*density*Vr, density*Vtheta,
  r*Vphi" (SuperOutletMHD3DProjection.cxx:155-170), r^3 rho and r^3 p, phi mirrored.
* FarFieldEulerChar2D/3D: the boundary state 0.5 (inner + ghost) carries the mean Riemann invariant as its normal velocity
  and keeps the tangential velocity, density and pressure of the upwind side (FarFieldEulerChar2D.cxx:190-245).
* MHDConsACAST: S[phi] = -(refSpeed/dissipCoeff)^2 phi V and its Jacobian on the diagonal block (MHDConsACASourceTerm.cxx:197-205)."""
import numpy as np

import cases
from cases import B, meshgen


def _spherical(x, v):
    r = np.linalg.norm(x, axis=1); rho = np.hypot(x[:, 0], x[:, 1])
    er = x/r[:, None]
    et = np.stack([x[:, 0]*x[:, 2]/(rho*r), x[:, 1]*x[:, 2]/(rho*r), -rho/r], axis=1)
    ep = np.stack([-x[:, 1]/rho, x[:, 0]/rho, np.zeros_like(r)], axis=1)
    return r, (er*v).sum(1), (et*v).sum(1), (ep*v).sum(1)


def test_coronal_outlet_keeps_its_invariants(oracle):
    fm = oracle.build_mesh(meshgen.hex_box(5, 4, 4))
    cfg = cases.mhd3d_cfg()
    bcs = cases.bcs_mhd3d_outlet(True)
    U = cases.init_state(fm, 9, "mhd")
    oracle.apply_bc(cfg, fm, bcs, U)
    U = U.reshape(-1, 9)
    x = fm.centres.reshape(-1, 3)
    g = np.nonzero(np.isin(fm.faceBC[fm.nInner:], [1, 3]))[0]
    assert len(g) > 10
    I, G = U[fm.faceL[fm.nInner + g]], U[fm.nCells + g]
    xI, xG = x[fm.faceL[fm.nInner + g]], x[fm.nCells + g]
    rI, BrI, BtI, BpI = _spherical(xI, I[:, 4:7]); rG, BrG, BtG, BpG = _spherical(xG, G[:, 4:7])
    _, VrI, VtI, VpI = _spherical(xI, I[:, 1:4]); _, VrG, VtG, VpG = _spherical(xG, G[:, 1:4])
    close = lambda a, b: np.allclose(a, b, rtol=1e-12, atol=1e-14)
    assert close(rG**2*BrG, rI**2*BrI) and close(BtG, BtI) and close(rG*BpG, rI*BpI)
    assert close(rG**3*G[:, 0], rI**3*I[:, 0]) and close(rG**3*G[:, 7], rI**3*I[:, 7])
    assert close(rG**2*G[:, 0]*VrG, rI**2*I[:, 0]*VrI) and close(G[:, 0]*VtG, I[:, 0]*VtI) and close(rG*VpG, rI*VpI)
    assert np.array_equal(G[:, 8], -I[:, 8])
    # the refPhi form: the inner state with phi replaced
    bcs = cases.bcs_mhd3d_outlet(False, 0.02)
    U2 = cases.init_state(fm, 9, "mhd")
    oracle.apply_bc(cfg, fm, bcs, U2)
    U2 = U2.reshape(-1, 9)
    G2 = U2[fm.nCells + g]
    assert np.allclose(G2[:, :8], I[:, :8], rtol=1e-13, atol=0) and (G2[:, 8] == 0.02).all()


def test_char_far_field_boundary_state(oracle):
    for dim, em, cfg in ((2, meshgen.quad_channel(10, 8), cases.ns2d_cfg()), (3, meshgen.hex_box(5, 4, 4), cases.ns3d_cfg())):
        fm = oracle.build_mesh(em)
        bcs = cases.bcs_farfield_char(dim, True)
        q = dim + 2
        inf = np.array([bcs[0].value[i] for i in range(q)])
        U = cases.init_state(fm, q, True)
        oracle.apply_bc(cfg, fm, bcs, U)
        U = U.reshape(-1, q)
        nb = fm.nFaces - fm.nInner
        I, G = U[fm.faceL[fm.nInner:]], U[fm.nCells:fm.nCells + nb]
        n = fm.normals.reshape(-1, dim)[fm.nInner:]; n = n/np.linalg.norm(n, axis=1)[:, None]
        gam, R = cfg.gamma, cfg.R
        p, v, T = I[:, 0], I[:, 1:1+dim], I[:, q-1]
        a = np.sqrt(gam*R*T); vn = (v*n).sum(1)
        aInf = np.sqrt(gam*R*inf[q-1]); vnInf = (inf[1:1+dim]*n).sum(1)
        mean = 0.5*(vn + 2*a/(gam - 1) + vnInf - 2*aInf/(gam - 1))
        assert (np.abs(vn/a) < 1).all()                       # the subsonic state of these cases
        out = mean > 0
        assert out.any() and (~out).any()
        Vb = 0.5*(v + G[:, 1:1+dim])                          # boundary velocity
        assert np.allclose((Vb*n).sum(1), mean, rtol=1e-12, atol=1e-13)
        ref = np.where(out[:, None], v, inf[None, 1:1+dim])   # tangential part of the upwind side
        tang = lambda w: w - (w*n).sum(1)[:, None]*n
        assert np.allclose(tang(Vb), tang(ref), rtol=1e-11, atol=1e-13)
        # ghost p, T: the inner ones on outflow, the far-field ones on inflow (rho = p/(R T) of that side)
        assert np.allclose(G[out, 0], p[out], rtol=1e-13) and np.allclose(G[~out, 0], inf[0], rtol=1e-13)
        assert np.allclose(G[out, q-1], T[out], rtol=1e-12) and np.allclose(G[~out, q-1], inf[q-1], rtol=1e-12)


def test_aca_source_term_against_its_formula(oracle):
    fm = oracle.build_mesh(meshgen.hex_box(5, 4, 4))
    U = cases.init_state(fm, 9, "mhd")
    rf, d = 0.8, 0.35
    a = oracle.compute(cases.mhd3d_cfg(res_factor=rf), fm, cases.bcs_mhd3d(), U.copy(), jacobian=True)
    b = oracle.compute(cases.mhd3d_cfg(source_term=B.ST_MHD_CONS_ACA, mhd_dissip_coeff=d, res_factor=rf), fm, cases.bcs_mhd3d(),
                       U.copy(), jacobian=True)
    dr = (b.rhs - a.rhs).reshape(-1, 9)
    phi = U.reshape(-1, 9)[:fm.nCells, 8]
    assert not dr[:, :8].any()
    assert np.allclose(dr[:, 8], rf*(-(1.0/d**2))*phi*fm.volumes, rtol=1e-12, atol=1e-18)       # refSpeed = 1
    dj = (b.bsrVal - a.bsrVal).reshape(-1, 9, 9)
    diag = np.array([fm.bsrRowPtr[i] + np.searchsorted(fm.bsrColInd[fm.bsrRowPtr[i]:fm.bsrRowPtr[i+1]], i) for i in range(fm.nCells)])
    off = np.ones(len(dj), dtype=bool); off[diag] = False
    assert not dj[off].any()                                                                   # diagonal blocks only
    want = np.zeros((fm.nCells, 9, 9)); want[:, 8, 8] = rf*(1.0/d**2)*fm.volumes               # J = -d rhs / dU
    assert np.allclose(dj[diag], want, rtol=1e-7, atol=1e-12)                                  # a finite difference of a linear term
    # "Hyperbolic" correction: no term at all
    c = oracle.compute(cases.mhd3d_cfg(source_term=B.ST_MHD_CONS_ACA, res_factor=rf), fm, cases.bcs_mhd3d(), U.copy(), jacobian=True)
    assert np.array_equal(c.rhs, a.rhs) and np.array_equal(c.bsrVal, a.bsrVal)
