"""Shared parity cases: (mesh, configuration, boundary conditions, seeded state) used by the CPU checks of the
kernel logic (tests/hostsim) and by the GPU parity tests through the C ABI."""
import numpy as np
from coolfluid_b200 import meshgen, config as B


def relerr(a, b):
    s = np.abs(b).max()
    return float(np.abs(a - b).max()/(s if s > 0 else 1.0))


def relerr_rows(a, b, width):
    """Largest error relative to the max-abs of its own row (one cell's residual, one Jacobian block): the per-cell /
    per-entry measure BASELINE.json's 1e-12 is stated in (SURVEY.md 8d parity gates)."""
    a = np.asarray(a).reshape(-1, width); b = np.asarray(b).reshape(-1, width)
    scale = np.maximum(np.abs(b).max(axis=1, keepdims=True), 1e-300)
    return float((np.abs(a - b)/scale).max())


def init_state(fm, q, pvt, seed=67890):
    """Smooth field + seeded noise, physically admissible (SURVEY.md §8d synthetic inputs)."""
    d = fm.dim
    x = fm.centres.reshape(-1, d)[:fm.nCells]
    rng = np.random.RandomState(seed)
    n = fm.nCells
    U = np.zeros((n + fm.nGhosts, q))
    s = 1.0 + 0.05*np.sin(2*np.pi*x[:, 0])*np.cos(2*np.pi*x[:, 1]) + 0.01*rng.standard_normal(n)
    fast = 1.0
    if pvt in ("cons_fast", "pvt_fast"):       # normal Mach numbers beyond +-1 at the x boundaries (FarFieldEuler branches)
        fast = 2.0 if pvt == "cons_fast" else 3.5
        pvt = (pvt == "pvt_fast")
    if pvt == "mhd":
        # rho, rho V, B (dipole-like smooth field + uniform part), rho E, phi
        rho = s
        u = 0.8 + 0.1*np.cos(2*np.pi*x[:, 1]); v = 0.05*np.sin(2*np.pi*x[:, 0]); w = 0.02*np.sin(2*np.pi*x[:, 2])
        r2 = ((x - np.array([0.5, 0.5, -0.6]))**2).sum(axis=1)
        bx = 0.3 + 0.05*(x[:, 0] - 0.5)/r2**1.5; by = 0.1 + 0.05*(x[:, 1] - 0.5)/r2**1.5; bz = 0.05 + 0.05*(x[:, 2] + 0.6)/r2**1.5
        p = 1.0 + 0.05*np.cos(2*np.pi*x[:, 0])
        U[:n, 0] = rho; U[:n, 1] = rho*u; U[:n, 2] = rho*v; U[:n, 3] = rho*w
        U[:n, 4] = bx; U[:n, 5] = by; U[:n, 6] = bz
        U[:n, 7] = p/0.05 + 0.5*rho*(u*u + v*v + w*w) + 0.5*(bx*bx + by*by + bz*bz)
        U[:n, 8] = 0.01*np.sin(2*np.pi*x[:, 0])*np.sin(2*np.pi*x[:, 1])
    elif pvt:
        U[:n, 0] = s
        U[:n, 1] = fast*(0.4 + 0.05*np.cos(2*np.pi*x[:, 1]))
        U[:n, 2] = 0.03*np.sin(2*np.pi*x[:, 0])
        if d == 3:
            U[:n, 3] = 0.02*np.sin(2*np.pi*x[:, 2])
        U[:n, q-1] = 1.0 + 0.05*np.cos(2*np.pi*x[:, 0]) + 0.01*rng.standard_normal(n)
    else:
        rho = s
        u = fast*(0.8 + 0.1*np.cos(2*np.pi*x[:, 1]))
        v = 0.05*np.sin(2*np.pi*x[:, 0])
        p = 1.0 + 0.05*np.cos(2*np.pi*x[:, 0])
        U[:n, 0] = rho; U[:n, 1] = rho*u; U[:n, 2] = rho*v
        ke = 0.5*rho*(u*u + v*v)
        if d == 3:
            w = 0.02*np.sin(2*np.pi*x[:, 2])
            U[:n, 3] = rho*w
            ke = ke + 0.5*rho*w*w
        U[:n, q-1] = p/0.4 + ke
    return U.reshape(-1).copy()


def ns_kwargs():
    kw = B.ns_reference_values()
    for k in ("ref_speed", "rho_ref", "u_inf"):
        kw.pop(k)
    return kw


def bcs_ns3d():
    """one no-slip wall, one inlet, one outlet, mirrors (+ one supersonic outlet)."""
    wall = B.make_bc(B.BC_DIRICHLET, mask=[0, 1, 1, 1, 0], value=[0, 0, 0, 0, 0])
    inlet = B.make_bc(B.BC_DIRICHLET, mask=[0, 1, 1, 1, 1], value=[0, 0.4, 0, 0, 1.0])
    outlet = B.make_bc(B.BC_DIRICHLET, mask=[1, 0, 0, 0, 0], value=[1.0, 0, 0, 0, 0])
    mirror = B.make_bc(B.BC_MIRROR_VELOCITY, zero_grad=[1, 0, 0, 0, 1])
    return [inlet, outlet, wall, mirror, mirror, B.make_bc(B.BC_SUPER_OUTLET)]


def bcs_ns3d_isothermal(t_wall=1.1, t_ghost_min=0.1):
    """bcs_ns3d with the adiabatic wall replaced by NoSlipWallIsothermalNSvt (a moving wall to exercise uWall)."""
    b = bcs_ns3d()
    b[2] = B.make_bc(B.BC_NOSLIP_ISOTHERMAL, value=[t_ghost_min, 0.02, -0.01, 0.0, t_wall])
    return b


def bcs_ns2d_isothermal():
    b = bcs_ns2d()
    b[2] = B.make_bc(B.BC_NOSLIP_ISOTHERMAL, value=[0.1, 0.0, 0.0, 0.9])
    return b


def bcs_ns2d():
    wall = B.make_bc(B.BC_DIRICHLET, mask=[0, 1, 1, 0], value=[0, 0, 0, 0])
    inlet = B.make_bc(B.BC_DIRICHLET, mask=[0, 1, 1, 1], value=[0, 0.4, 0, 1.0])
    outlet = B.make_bc(B.BC_DIRICHLET, mask=[1, 0, 0, 0], value=[1.0, 0, 0, 0])
    return [inlet, outlet, wall, B.make_bc(B.BC_MIRROR_VELOCITY)]


def bcs_euler2d():
    mir = B.make_bc(B.BC_MIRROR_EULER_CONS)
    inl = B.make_bc(B.BC_DIRICHLET, mask=[1, 1, 1, 1], value=[1.0, 0.8, 0.0, 2.82])
    return [inl, B.make_bc(B.BC_SUPER_OUTLET), mir, mir]


def bcs_euler3d():
    mir = B.make_bc(B.BC_MIRROR_EULER_CONS)
    return [B.make_bc(B.BC_SUPER_OUTLET)]*2 + [mir]*4


def bcs_farfield(dim, pvt):
    """FarFieldEuler2D / FarFieldEuler3D on every boundary: value = p, u, v, [w,] T at infinity."""
    R = 1.0 if pvt else B.make_config(dim).R
    vel = [0.4, 0.0, 0.0][:dim] if pvt else [0.8, 0.0, 0.0][:dim]
    ff = B.make_bc(B.BC_FARFIELD_EULER, value=[1.0] + vel + [1.0/R])
    return [ff]*(2*dim)


def bcs_farfield_char(dim, pvt):
    """FarFieldEulerChar2D / 3D on every boundary (value = p, u, v, [w,] T at infinity)."""
    b = bcs_farfield(dim, pvt)
    for x in b:
        x.type = B.BC_FARFIELD_EULER_CHAR
    return b


def bcs_mhd3d():
    """SuperInlet (constant Def) / SuperOutlet / MirrorMHD3DProjection."""
    inl = B.make_bc(B.BC_DIRICHLET, mask=[1]*9, value=[1.0, 0.8, 0.0, 0.0, 0.3, 0.1, 0.05, 2.5, 0.0])
    mir = B.make_bc(B.BC_MIRROR_EULER_CONS)      # dispatched by the physical model: MirrorMHD3DProjection
    return [inl, B.make_bc(B.BC_SUPER_OUTLET), mir, mir, mir, mir]


def bcs_mhd3d_outlet(coronal, ref_phi=0.0):
    """bcs_mhd3d with SuperOutletMHD3DProjection (coronalBC or refPhi form) at xmax and on one side."""
    b = bcs_mhd3d()
    out = B.make_bc(B.BC_SUPER_OUTLET_MHD, mask=[1 if coronal else 0] + [0]*8, value=[ref_phi] + [0.]*8)
    b[1] = out
    b[3] = out
    return b


def mhd3d_cfg(**over):
    """BASELINE config 4: MHD3DProjection, q = 9, gamma 1.05, refSpeed 1."""
    kw = dict(physical_model=B.PM_MHD3D_PROJECTION, flux=B.FS_HLL, poly_rec=B.PR_LINEARLS, limiter=B.LIM_VENKTN,
              jacobian=True, gamma=1.05, ref_speed=1.0)
    kw.update(over)
    return B.make_config(3, **kw)


def with_stencil(em, kind):
    """The element mesh with the reconstruction stencil of `Setup1.stencil = kind` (a3; the default is FaceVertexPlusGhost)."""
    em.stencil_kind = kind
    return em


def ns3d_cfg(**over):
    kw = dict(update_var=B.UV_PVT, poly_rec=B.PR_LINEARLS, limiter=B.LIM_VENKTN, diffusive=B.DIFF_NS, jacobian=True)
    kw.update(ns_kwargs())
    kw.update(over)
    return B.make_config(3, **kw)


def ns2d_cfg(**over):
    kw = dict(update_var=B.UV_PVT, poly_rec=B.PR_LINEARLS, limiter=B.LIM_VENKTN, diffusive=B.DIFF_NS, jacobian=True)
    kw.update(ns_kwargs())
    kw.update(over)
    return B.make_config(2, **kw)


# name -> (mesh factory, config factory, bcs factory, primitive update variables?)
CASES = {
    # BASELINE config 3: 3D Navier-Stokes, Roe + viscous, LSQ + Venkatakrishnan, implicit (FD Jacobian)
    "ns3d_hex": (lambda: meshgen.hex_box(7, 6, 5), ns3d_cfg, bcs_ns3d, True),
    "ns3d_tet": (lambda: meshgen.tet_box(4, 3, 3), ns3d_cfg, bcs_ns3d, True),
    "ns3d_hex_barth_fullstencil": (lambda: meshgen.hex_box(6, 5, 4),
                                   lambda: ns3d_cfg(limiter=B.LIM_BARTHJESP, use_full_stencil=True), bcs_ns3d, True),
    "ns3d_hex_resfactor": (lambda: meshgen.hex_box(5, 5, 4), lambda: ns3d_cfg(res_factor=0.5, gradient_factor=0.7),
                           bcs_ns3d, True),
    # BASELINE config 1: 2D Euler, first-order Roe
    "euler2d_1st_quad": (lambda: meshgen.quad_channel(12, 9), lambda: B.make_config(2, jacobian=True), bcs_euler2d, False),
    "euler2d_1st_tri": (lambda: meshgen.quad_channel(12, 9, triangles=True), lambda: B.make_config(2, jacobian=True),
                        bcs_euler2d, False),
    # BASELINE config 2: 2D Euler, LSQ + Venkatakrishnan, implicit
    "euler2d_2nd_quad": (lambda: meshgen.quad_channel(12, 9),
                         lambda: B.make_config(2, poly_rec=B.PR_LINEARLS, limiter=B.LIM_VENKTN, jacobian=True),
                         bcs_euler2d, False),
    "euler2d_2nd_tri_fullstencil": (lambda: meshgen.quad_channel(12, 9, triangles=True),
                                    lambda: B.make_config(2, poly_rec=B.PR_LINEARLS, limiter=B.LIM_VENKTN, jacobian=True,
                                                          use_full_stencil=True), bcs_euler2d, False),
    "ns2d_quad": (lambda: meshgen.quad_channel(10, 8), ns2d_cfg, bcs_ns2d, True),
    "ns2d_tri_nolimiter": (lambda: meshgen.quad_channel(10, 8, triangles=True), lambda: ns2d_cfg(limiter=B.LIM_NONE),
                           bcs_ns2d, True),
    "euler3d_cons_tet": (lambda: meshgen.tet_box(3, 3, 3),
                         lambda: B.make_config(3, poly_rec=B.PR_LINEARLS, limiter=B.LIM_VENKTN, jacobian=True),
                         bcs_euler3d, False),
    # a12: AUSM+ (2D Euler, cons, 1st order; 3D NS Pvt 2nd order) and AUSM+up (jets2DFVMImplAUSMAnalytic-like options,
    # numerical Jacobian), all three perfect-gas interface sound speeds
    "euler2d_ausmplus_1st": (lambda: meshgen.quad_channel(12, 9),
                             lambda: B.make_config(2, flux=B.FS_AUSMPLUS, jacobian=True), bcs_euler2d, False),
    "euler2d_ausmplusup_2nd_tri": (lambda: meshgen.quad_channel(12, 9, triangles=True),
                                   lambda: B.make_config(2, flux=B.FS_AUSMPLUSUP, poly_rec=B.PR_LINEARLS,
                                                         limiter=B.LIM_VENKTN, jacobian=True, mach_inf=2.0, choice_a12=1),
                                   bcs_euler2d, False),
    "ns3d_hex_ausmplusup": (lambda: meshgen.hex_box(6, 5, 4), lambda: ns3d_cfg(flux=B.FS_AUSMPLUSUP, mach_inf=0.35),
                            bcs_ns3d, True),
    "ns3d_tet_ausmplus_a2": (lambda: meshgen.tet_box(3, 3, 3), lambda: ns3d_cfg(flux=B.FS_AUSMPLUS, choice_a12=2),
                             bcs_ns3d, True),
    "euler3d_ausmplusup_a3": (lambda: meshgen.tet_box(3, 3, 3),
                              lambda: B.make_config(3, flux=B.FS_AUSMPLUSUP, poly_rec=B.PR_LINEARLS, limiter=B.LIM_VENKTN,
                                                    jacobian=True, mach_inf=0.8, choice_a12=3), bcs_euler3d, False),
    # a19 / BASELINE config 2: AUSM+up with the ANALYTIC Jacobian (NumJacobAnalytic), conservative variables
    "euler2d_ausmplusup_analytic": (lambda: meshgen.quad_channel(12, 9, triangles=True),
                                    lambda: B.make_config(2, flux=B.FS_AUSMPLUSUP, poly_rec=B.PR_LINEARLS,
                                                          limiter=B.LIM_VENKTN, jacobian="analytic", mach_inf=2.0),
                                    bcs_euler2d, False),
    "euler3d_ausmplusup_analytic": (lambda: meshgen.hex_box(5, 4, 4),
                                    lambda: B.make_config(3, flux=B.FS_AUSMPLUSUP, poly_rec=B.PR_LINEARLS,
                                                          limiter=B.LIM_VENKTN, jacobian="analytic", mach_inf=0.8,
                                                          res_factor=0.5), bcs_euler3d, False),
    # a19 in Puvt / Pvt variables (ComputedFluxdPvt) and with the analytic diffusive Jacobian (NSFluxAndJacob), all three
    # DiamondVolume control volumes + Corrected2D
    "euler2d_pvt_ausmplusup_analytic": (lambda: meshgen.quad_channel(12, 9, triangles=True),
                                        lambda: B.make_config(2, update_var=B.UV_PVT, flux=B.FS_AUSMPLUSUP,
                                                              poly_rec=B.PR_LINEARLS, limiter=B.LIM_VENKTN,
                                                              jacobian="analytic", mach_inf=0.5, **ns_kwargs()),
                                        bcs_ns2d, True),
    "ns2d_quad_ausmplusup_analytic": (lambda: meshgen.quad_channel(10, 8),
                                      lambda: ns2d_cfg(flux=B.FS_AUSMPLUSUP, jacobian="analytic", mach_inf=0.35),
                                      bcs_ns2d, True),
    "ns2d_tri_ausmplusup_analytic_corrected": (lambda: meshgen.quad_channel(10, 8, triangles=True),
                                               lambda: ns2d_cfg(flux=B.FS_AUSMPLUSUP, jacobian="analytic", mach_inf=0.35,
                                                                derivative="corrected", res_factor=0.5),
                                               bcs_ns2d, True),
    "ns3d_hex_ausmplusup_analytic": (lambda: meshgen.hex_box(6, 5, 4),
                                     lambda: ns3d_cfg(flux=B.FS_AUSMPLUSUP, jacobian="analytic", mach_inf=0.35),
                                     bcs_ns3d, True),
    "ns3d_tet_ausmplusup_analytic": (lambda: meshgen.tet_box(3, 3, 3),
                                     lambda: ns3d_cfg(flux=B.FS_AUSMPLUSUP, jacobian="analytic", mach_inf=0.35,
                                                      choice_a12=1), bcs_ns3d, True),
    # "RoeEntropyFix" (RoeEntropyFixFlux.cxx), both correction types
    "euler2d_roe_efix1": (lambda: meshgen.quad_channel(12, 9, triangles=True),
                          lambda: B.make_config(2, poly_rec=B.PR_LINEARLS, limiter=B.LIM_VENKTN, jacobian=True,
                                                entropy_fix_id=1), bcs_euler2d, False),
    "ns3d_hex_roe_efix2": (lambda: meshgen.hex_box(5, 4, 4), lambda: ns3d_cfg(entropy_fix_id=2), bcs_ns3d, True),
    # "RoeSA" (RoeSAFlux.cxx): Sanders' H-correction over the faces of the two cells, all four entropyFixID forms, both
    # stencils; the cylinder case of the reference (cyl_Pg_M15_FVM_1st2nd.CFcase:104) is Euler2D Puvt + RoeSA
    "euler2d_puvt_roesa_quad": (lambda: meshgen.quad_channel(12, 9),
                                lambda: B.make_config(2, update_var=B.UV_PVT, flux=B.FS_ROESA, poly_rec=B.PR_LINEARLS,
                                                      limiter=B.LIM_VENKTN, jacobian=True, **ns_kwargs()), bcs_ns2d, True),
    "euler2d_roesa_tri_fix1": (lambda: meshgen.quad_channel(12, 9, triangles=True),
                               lambda: B.make_config(2, flux=B.FS_ROESA, poly_rec=B.PR_LINEARLS, limiter=B.LIM_VENKTN,
                                                     jacobian=True, entropy_fix_id=1), bcs_euler2d, False),
    "euler2d_roesa_1st_oldstencil": (lambda: meshgen.quad_channel(12, 9),
                                     lambda: B.make_config(2, flux=B.FS_ROESA, jacobian=True, entropy_fix_id=3,
                                                           roe_sa_old_stencil=1), bcs_euler2d, False),
    "ns3d_hex_roesa_fix2": (lambda: meshgen.hex_box(5, 4, 4), lambda: ns3d_cfg(flux=B.FS_ROESA, entropy_fix_id=2),
                            bcs_ns3d, True),
    "ns3d_tet_roesa": (lambda: meshgen.tet_box(3, 3, 3), lambda: ns3d_cfg(flux=B.FS_ROESA), bcs_ns3d, True),
    # a13 (Euler part): Lax-Friedrichs and HLL
    "euler2d_laxfried_1st": (lambda: meshgen.quad_channel(12, 9),
                             lambda: B.make_config(2, flux=B.FS_LAXFRIED, jacobian=True), bcs_euler2d, False),
    "euler3d_hll_2nd_tet": (lambda: meshgen.tet_box(3, 3, 3),
                            lambda: B.make_config(3, flux=B.FS_HLL, poly_rec=B.PR_LINEARLS, limiter=B.LIM_VENKTN,
                                                  jacobian=True), bcs_euler3d, False),
    "ns3d_hex_hll_alpha": (lambda: meshgen.hex_box(5, 4, 4), lambda: ns3d_cfg(flux=B.FS_HLL, hll_use_alpha=1),
                           bcs_ns3d, True),
    "ns2d_quad_laxfried": (lambda: meshgen.quad_channel(10, 8), lambda: ns2d_cfg(flux=B.FS_LAXFRIED, diff_red_coeff=0.8),
                           bcs_ns2d, True),
    # "Centred" (CentredFlux.cxx): 0.5*(fR + fL), on the Euler variable sets and on MHD3DProjection
    "euler2d_centred_2nd_tri": (lambda: meshgen.quad_channel(12, 9, triangles=True),
                                lambda: B.make_config(2, flux=B.FS_CENTRED, poly_rec=B.PR_LINEARLS, limiter=B.LIM_VENKTN,
                                                      jacobian=True), bcs_euler2d, False),
    "ns3d_hex_centred": (lambda: meshgen.hex_box(5, 4, 4), lambda: ns3d_cfg(flux=B.FS_CENTRED), bcs_ns3d, True),
    "mhd3d_hex_centred": (lambda: meshgen.hex_box(4, 4, 3), lambda: mhd3d_cfg(flux=B.FS_CENTRED), bcs_mhd3d, "mhd"),
    # BASELINE config 4: 3D ideal MHD with hyperbolic divergence cleaning (q = 9), HLL / Lax-Friedrichs
    "mhd3d_hex_hll": (lambda: meshgen.hex_box(5, 4, 4), mhd3d_cfg, bcs_mhd3d, "mhd"),
    # background potential field B0 = "Dipole" (MHDTerm potentialBType): enters the wave speeds and the update coefficient
    # Data.SourceTerm = MHDConsACAST (26 of the reference's MHD cases, solar terms off): phi damping of the "Mixed" correction
    # in rhs (first touch) and its FD Jacobian on the diagonal blocks; "Hyperbolic": the term and its Jacobian are zero
    "mhd3d_hex_hll_aca_mixed": (lambda: meshgen.hex_box(5, 4, 4),
                                lambda: mhd3d_cfg(source_term=B.ST_MHD_CONS_ACA, mhd_dissip_coeff=0.35, res_factor=0.8),
                                bcs_mhd3d, "mhd"),
    "mhd3d_tet_laxfried_aca_hyperbolic": (lambda: meshgen.tet_box(3, 3, 2),
                                          lambda: mhd3d_cfg(flux=B.FS_LAXFRIED, source_term=B.ST_MHD_CONS_ACA), bcs_mhd3d, "mhd"),
    # SuperOutletMHD3DProjection (16 of the reference's cases): the coronal form (default) and the refPhi form, with the FD
    # Jacobian of the boundary faces (the BC is re-applied on the perturbed inner state)
    "mhd3d_hex_hll_outlet_coronal": (lambda: meshgen.hex_box(5, 4, 4),
                                     lambda: mhd3d_cfg(source_term=B.ST_MHD_CONS_ACA, mhd_dissip_coeff=0.5),
                                     lambda: bcs_mhd3d_outlet(True), "mhd"),
    "mhd3d_tet_laxfried_outlet_refphi": (lambda: meshgen.tet_box(3, 3, 2), lambda: mhd3d_cfg(flux=B.FS_LAXFRIED),
                                         lambda: bcs_mhd3d_outlet(False, 0.02), "mhd"),
    "mhd3d_hex_hll_dipole": (lambda: meshgen.hex_box(5, 4, 4), lambda: mhd3d_cfg(mhd_dipole=(0.002, 0.001, 0.004)),
                             bcs_mhd3d, "mhd"),
    "mhd3d_tet_laxfried_1st_dipole": (lambda: meshgen.tet_box(3, 3, 2),
                                      lambda: mhd3d_cfg(flux=B.FS_LAXFRIED, poly_rec=B.PR_CONSTANT, limiter=B.LIM_NONE,
                                                        mhd_dipole=(0.0, 0.0, 0.005)), bcs_mhd3d, "mhd"),
    "mhd3d_tet_laxfried_1st": (lambda: meshgen.tet_box(3, 3, 2),
                               lambda: mhd3d_cfg(flux=B.FS_LAXFRIED, poly_rec=B.PR_CONSTANT, limiter=B.LIM_NONE),
                               bcs_mhd3d, "mhd"),
    "mhd3d_hex_laxfried_barth": (lambda: meshgen.hex_box(4, 4, 3),
                                 lambda: mhd3d_cfg(flux=B.FS_LAXFRIED, limiter=B.LIM_BARTHJESP, res_factor=0.5),
                                 bcs_mhd3d, "mhd"),
    # a7 variants: the "Venkatakrishnan paper" branch (isMFMHD) of Venktn2D/3D and Venktn3DStrict (what the reference's MHD
    # dipole case selects, plugins/MHD/testcases/COCONUT/Dipole/map_TEST.CFcase:73)
    "ns3d_hex_venktn_mfmhd": (lambda: meshgen.hex_box(6, 5, 4), lambda: ns3d_cfg(is_mfmhd=1, coeff_eps=0.7), bcs_ns3d, True),
    "euler2d_quad_venktn_nolimiterid": (lambda: meshgen.quad_channel(12, 9),
                                        lambda: B.make_config(2, poly_rec=B.PR_LINEARLS, limiter=B.LIM_VENKTN, jacobian=True,
                                                              no_limiter_id=[0, 3]), bcs_euler2d, False),
    "euler2d_tri_venktn_mfmhd": (lambda: meshgen.quad_channel(12, 9, triangles=True),
                                 lambda: B.make_config(2, poly_rec=B.PR_LINEARLS, limiter=B.LIM_VENKTN, jacobian=True,
                                                       is_mfmhd=1, magnitude_values=[1.0, 0.9, 0.2, 2.5]), bcs_euler2d, False),
    "mhd3d_hex_venktn_strict": (lambda: meshgen.hex_box(5, 4, 4),
                                lambda: mhd3d_cfg(limiter=B.LIM_VENKTN_STRICT, strict_coeff=0.6,
                                                  magnitude_values=[1.0, 0.8, 0.8, 0.8, 0.3, 0.3, 0.3, 20.0, 0.1]),
                                bcs_mhd3d, "mhd"),
    "mhd3d_tet_venktn_strict_psimin1": (lambda: meshgen.tet_box(3, 3, 2),
                                        lambda: mhd3d_cfg(limiter=B.LIM_VENKTN_STRICT, psi_min_equal_1=1, use_full_stencil=True),
                                        bcs_mhd3d, "mhd"),
    # a16: NoSlipWallIsothermalNSvt (FiniteVolumeNavierStokes/NoSlipWallIsothermalNSvt.ci:178-243), hot and cold wall
    "ns3d_hex_isothermal_wall": (lambda: meshgen.hex_box(6, 5, 4), ns3d_cfg, bcs_ns3d_isothermal, True),
    "ns3d_tet_isothermal_wall_corrected": (lambda: meshgen.tet_box(3, 3, 3), lambda: ns3d_cfg(derivative="corrected"),
                                           bcs_ns3d_isothermal, True),
    "ns2d_quad_isothermal_wall": (lambda: meshgen.quad_channel(10, 8), ns2d_cfg, bcs_ns2d_isothermal, True),
    # a15: the Corrected2D / Corrected3D derivative computer (what most reference test cases select) on every face type
    "ns3d_hex_corrected": (lambda: meshgen.hex_box(6, 5, 4), lambda: ns3d_cfg(derivative="corrected"), bcs_ns3d, True),
    "ns3d_tet_corrected_ausmplusup": (lambda: meshgen.tet_box(3, 3, 3),
                                      lambda: ns3d_cfg(derivative="corrected", flux=B.FS_AUSMPLUSUP, mach_inf=0.35),
                                      bcs_ns3d, True),
    "ns2d_quad_corrected": (lambda: meshgen.quad_channel(10, 8), lambda: ns2d_cfg(derivative="corrected"), bcs_ns2d, True),
    "ns2d_tri_corrected_fullstencil": (lambda: meshgen.quad_channel(10, 8, triangles=True),
                                       lambda: ns2d_cfg(derivative="corrected", use_full_stencil=True, res_factor=0.5),
                                       bcs_ns2d, True),
    # a15, third derivative computer: "StateDiff" (two-point difference along the line of centres)
    "ns2d_quad_statediff": (lambda: meshgen.quad_channel(10, 8), lambda: ns2d_cfg(derivative="statediff"), bcs_ns2d, True),
    "ns3d_tet_statediff_1st": (lambda: meshgen.tet_box(3, 3, 3),
                               lambda: ns3d_cfg(derivative="statediff", poly_rec=B.PR_CONSTANT, limiter=B.LIM_NONE,
                                                flux=B.FS_AUSMPLUS), bcs_ns3d, True),
    "ns3d_hex_statediff": (lambda: meshgen.hex_box(5, 4, 4), lambda: ns3d_cfg(derivative="statediff", res_factor=0.5),
                           bcs_ns3d, True),
    # a15, fourth derivative computer: "CorrectedGG2D/3D" (Green-Gauss cell gradients over the nodal states; boundary least
    # squares over the face neighbours; the shifted "average state" of the reference)
    "ns2d_quad_correctedgg": (lambda: meshgen.quad_channel(10, 8), lambda: ns2d_cfg(derivative="correctedgg"), bcs_ns2d, True),
    "ns2d_tri_correctedgg_weights": (lambda: meshgen.quad_channel(10, 8, triangles=True),
                                     lambda: ns2d_cfg(derivative="correctedgg", gg_use_weights=1, res_factor=0.5),
                                     bcs_ns2d, True),
    "ns3d_hex_correctedgg": (lambda: meshgen.hex_box(5, 4, 4), lambda: ns3d_cfg(derivative="correctedgg"), bcs_ns3d, True),
    "ns3d_tet_correctedgg_ausmplus": (lambda: meshgen.tet_box(3, 3, 3),
                                      lambda: ns3d_cfg(derivative="correctedgg", flux=B.FS_AUSMPLUS, gg_use_weights=1),
                                      bcs_ns3d, True),
    "ns3d_single_cell_corrected": (lambda: meshgen.hex_box(1, 1, 1, jitter=0.0), lambda: ns3d_cfg(derivative="corrected"),
                                   bcs_ns3d, True),
    # a16: FarFieldEuler2D (conservative) / FarFieldEuler3D (Cons, Pvt): subsonic in/outflow and, with the fast initial
    # field, supersonic inflow at xmin and outflow at xmax
    "euler2d_farfield_sub": (lambda: meshgen.quad_channel(12, 9),
                             lambda: B.make_config(2, poly_rec=B.PR_LINEARLS, limiter=B.LIM_VENKTN, jacobian=True),
                             lambda: bcs_farfield(2, False), False),
    "euler2d_farfield_super": (lambda: meshgen.quad_channel(12, 9, triangles=True),
                               lambda: B.make_config(2, poly_rec=B.PR_LINEARLS, limiter=B.LIM_VENKTN, jacobian=True),
                               lambda: bcs_farfield(2, False), "cons_fast"),
    "euler3d_farfield_cons_tet": (lambda: meshgen.tet_box(3, 3, 3),
                                  lambda: B.make_config(3, poly_rec=B.PR_LINEARLS, limiter=B.LIM_VENKTN, jacobian=True),
                                  lambda: bcs_farfield(3, False), "cons_fast"),
    "ns3d_farfield_pvt_sub": (lambda: meshgen.hex_box(5, 4, 4), ns3d_cfg, lambda: bcs_farfield(3, True), True),
    "ns3d_farfield_pvt_super": (lambda: meshgen.hex_box(5, 4, 4), lambda: ns3d_cfg(derivative="corrected"),
                                lambda: bcs_farfield(3, True), "pvt_fast"),
    # a3: the other ComputeStencil strategies (Setup1.stencil = FaceVertex: 59 of the reference's cases; FaceEdge; Face)
    "euler2d_2nd_quad_facevertex": (lambda: with_stencil(meshgen.quad_channel(12, 9), "FaceVertex"),
                                    lambda: B.make_config(2, poly_rec=B.PR_LINEARLS, limiter=B.LIM_VENKTN, jacobian=True,
                                                          use_full_stencil=True), bcs_euler2d, False),
    "ns3d_hex_facevertex": (lambda: with_stencil(meshgen.hex_box(6, 5, 4), "FaceVertex"), ns3d_cfg, bcs_ns3d, True),
    "ns3d_tet_faceedge": (lambda: with_stencil(meshgen.tet_box(3, 3, 3), "FaceEdge"), ns3d_cfg, bcs_ns3d, True),
    "euler2d_2nd_quad_face": (lambda: with_stencil(meshgen.quad_channel(10, 8), "Face"),
                              lambda: B.make_config(2, poly_rec=B.PR_LINEARLS, limiter=B.LIM_BARTHJESP, jacobian=True,
                                                    use_full_stencil=True), bcs_euler2d, False),
    # FarFieldEulerChar2D / 3D: subsonic in / out by the mean Riemann invariant (2D Puvt, 3D Pvt), supersonic in / out (Cons)
    "ns2d_farfield_char_sub": (lambda: meshgen.quad_channel(10, 8), ns2d_cfg, lambda: bcs_farfield_char(2, True), True),
    "ns3d_farfield_char_sub": (lambda: meshgen.hex_box(5, 4, 4), ns3d_cfg, lambda: bcs_farfield_char(3, True), True),
    "ns3d_farfield_char_super": (lambda: meshgen.tet_box(3, 3, 2), ns3d_cfg, lambda: bcs_farfield_char(3, True), "pvt_fast"),
    "euler2d_farfield_char_super": (lambda: meshgen.quad_channel(12, 9, triangles=True),
                                    lambda: B.make_config(2, poly_rec=B.PR_LINEARLS, limiter=B.LIM_VENKTN, jacobian=True),
                                    lambda: bcs_farfield_char(2, False), "cons_fast"),
    # single cell / single row edge cases
    "ns3d_single_cell": (lambda: meshgen.hex_box(1, 1, 1, jitter=0.0), ns3d_cfg, bcs_ns3d, True),
    "euler2d_single_row": (lambda: meshgen.quad_channel(5, 1), lambda: B.make_config(2, poly_rec=B.PR_LINEARLS,
                           limiter=B.LIM_BARTHJESP, jacobian=True), bcs_euler2d, False),
}

FIELDS = ("uX", "uY", "uZ", "limiter", "nstates")


def run_case(name, oracle, make_engine, updatable=None, limiter_init=None, cfg_mut=None, shuffle=False):
    """Runs the oracle and an engine on one case; returns {quantity: (engine array, oracle array)}.
    shuffle: the cells of the mesh are renumbered at random first (the oracle then takes THAT numbering as the reference's)."""
    mesh_f, cfg_f, bcs_f, pvt = CASES[name]
    em = mesh_f()
    stencil = getattr(em, "stencil_kind", "FaceVertexPlusGhost")
    if shuffle:
        em, _ = meshgen.shuffle_cells(em)
    fm = oracle.build_mesh(em, stencil=stencil)
    cfg = cfg_f()
    if cfg_mut:
        cfg_mut(cfg)
    bcs = bcs_f()
    q = cfg.nEqs
    if updatable is not None:
        fm.updatable = np.ascontiguousarray(updatable(fm), dtype=np.uint8)
    U = init_state(fm, q, pvt)
    jac = bool(cfg.computeJacobian)
    uc0 = 0.1*np.arange(fm.nCells, dtype=np.float64)/max(fm.nCells, 1)      # updateCoeff accumulates onto its input
    Ug = U.copy()                                  # the oracle fills the ghost states in place
    r = oracle.compute(cfg, fm, bcs, Ug, jacobian=jac, limiter_init=limiter_init, update_coeff_init=uc0)
    assert r.status == 0
    e = make_engine(cfg)
    out = {}
    try:
        e.set_mesh(fm, bcs)
        e.set_states(U[:fm.nCells*q])
        e.set_update_coeff(uc0)
        if limiter_init is not None:
            e.set_limiter(limiter_init)
        rhs, uc = e.compute_rhs()
        out["rhs"] = (rhs, r.rhs.copy())
        out["updateCoeff"] = (uc, r.updateCoeff.copy())
        for f in FIELDS:
            if f == "uZ" and fm.dim == 2:
                continue
            out[f] = (e.field(f), getattr(r, f).copy())
        out["gstates"] = (e.field("gstates"), Ug[fm.nCells*q:].copy())
        if q == 9:
            out["divB"] = (e.field("divB"), r.divB.copy())      # a22
        if jac:
            out["jacobian"] = (e.bsr_values(), r.bsrVal.copy())
            rp, ci = e.bsr_pattern()
            out["bsrRowPtr"] = (rp, fm.bsrRowPtr)
            out["bsrColInd"] = (ci, fm.bsrColInd)
            # pseudo-steady time term on top (FVMCC_PseudoSteadyTimeRhs)
            past = U[:fm.nCells*q]*(1 + 1e-3)
            Uo = U.copy()
            oracle.apply_bc(cfg, fm, bcs, Uo)
            oracle.time_rhs(cfg, fm, bcs, Uo, np.concatenate([past, np.zeros(fm.nGhosts*q)]), r)
            e.set_past_states(past)
            out["time_rhs"] = (e.time_rhs(), r.rhs.copy())
            out["time_jacobian"] = (e.bsr_values(), r.bsrVal.copy())
        out["_fm"] = fm
    finally:
        e.close()
    return out


def check_newton_update(engine_factory, build_mesh):
    """StdUpdateSol (NewtonMethod/StdUpdateSol.cxx:110-141,256) through cfb200_update_sol_newton against numpy:
    U += alpha*dU on the updatable states only, dU zeroed elsewhere, updateCoeff zeroed, log10 L2 norm of dU[monitorVar]
    (ComputeL2Norm.cxx:111-143).  One multiplication and one addition per entry: bit-exact."""
    from coolfluid_b200 import lib as cl
    fm = build_mesh(meshgen.hex_box(6, 5, 4))
    x = fm.centres.reshape(-1, 3)[:fm.nCells, 0]
    fm.updatable = np.ascontiguousarray((x < 0.7).astype(np.uint8))
    e = engine_factory(ns3d_cfg())
    e.set_mesh(fm, bcs_ns3d())
    U = init_state(fm, 5, True)[:fm.nCells*5]
    e.set_states(U)
    rng = np.random.RandomState(7)
    dU = 1e-3*rng.standard_normal(fm.nCells*5)
    alpha = np.array([1.0, 0.9, 0.8, 0.7, 0.5])
    e.set_update_coeff(np.ones(fm.nCells))
    res = e.update_sol_newton(dU, alpha, 4)
    upd = fm.updatable.astype(bool)
    ref = U.reshape(-1, 5).copy()
    ref[upd] += alpha*dU.reshape(-1, 5)[upd]
    assert np.array_equal(e.get_states().reshape(-1, 5), ref)
    d4 = np.where(upd, dU.reshape(-1, 5)[:, 4], 0.0)
    assert abs(res - np.log10(np.sqrt((d4*d4).sum()))) <= 1e-12
    out = np.empty(fm.nCells)
    e._chk(e.L.get_field(e.h, b"updateCoeff", out.ctypes.data_as(cl.P_D)), "get_field")
    assert not out.any()
    rhs = np.empty(fm.nCells*5)
    e._chk(e.L.get_field(e.h, b"rhs", rhs.ctypes.data_as(cl.P_D)), "get_field")
    assert np.array_equal(rhs.reshape(-1, 5), np.where(upd[:, None], dU.reshape(-1, 5), 0.0))
    # the increment already on the device, unit relaxation
    before = e.get_states()
    e.update_sol_newton(None, None, 0)
    after = before.reshape(-1, 5).copy()
    after[upd] += dU.reshape(-1, 5)[upd]
    assert np.array_equal(e.get_states().reshape(-1, 5), after)
    e.close()


def check_bc_face_values(oracle, engine_factory):
    """Position-dependent SuperInlet (Def functions of x, y, z evaluated at the boundary point of every face,
    SuperInlet.cxx:156-209): per-face Dirichlet data through cfb200_set_bc_face_values against the oracle, with the
    FD Jacobian (the boundary blocks re-apply the BC on the perturbed inner state)."""
    fm = oracle.build_mesh(meshgen.hex_box(5, 4, 4))
    cfg = ns3d_cfg()
    bcs = bcs_ns3d()
    q = 5
    # boundary point of each face of record 0 (the inlet): 0.5*(inner centre + ghost centre)
    gb = fm.faceBC[fm.nInner:]
    faces = np.nonzero(gb == 0)[0]
    x = fm.centres.reshape(-1, 3)
    xb = 0.5*(x[fm.faceL[fm.nInner + faces]] + x[fm.nCells + faces])
    vals = np.zeros((len(faces), q))
    vals[:, 1] = 0.4*(1.0 + 0.3*np.sin(2*np.pi*xb[:, 1])*np.cos(np.pi*xb[:, 2]))
    vals[:, 2] = 0.02*xb[:, 1]
    vals[:, 3] = -0.01*xb[:, 2]
    vals[:, 4] = 1.0 + 0.05*xb[:, 1]*xb[:, 2]
    fm.bcFaceVal = np.zeros(fm.nGhosts*q)
    fm.bcFaceFlag = np.zeros(fm.nGhosts, dtype=np.uint8)
    fm.bcFaceVal.reshape(-1, q)[faces] = vals
    fm.bcFaceFlag[faces] = 1
    U = init_state(fm, q, True)
    Ug = U.copy()
    r = oracle.compute(cfg, fm, bcs, Ug, jacobian=True)
    e = engine_factory(cfg)
    try:
        e.set_mesh(fm, bcs)
        e.set_bc_face_values(0, vals.reshape(-1))
        e.set_states(U[:fm.nCells*q])
        rhs, uc = e.compute_rhs()
        assert np.array_equal(e.field("gstates"), Ug[fm.nCells*q:])
        assert np.array_equal(rhs, r.rhs) and np.array_equal(uc, r.updateCoeff)
        assert np.array_equal(e.bsr_values(), r.bsrVal)
        # back to the constant of the record
        e.set_bc_face_values(0, None)
        e.set_update_coeff(None)
        fm.bcFaceFlag[:] = 0
        Ug2 = U.copy()
        r2 = oracle.compute(cfg, fm, bcs, Ug2, jacobian=True)
        rhs2, _ = e.compute_rhs()
        assert np.array_equal(rhs2, r2.rhs) and not np.array_equal(rhs2, rhs)
    finally:
        e.close()


def check_nodal_prescribed(oracle, engine_factory):
    """NodalExtrapolation = DistanceBasedGMove (DistanceBasedExtrapolatorGMove.cxx): velocity and temperature imposed on
    the nodes that belong to the wall TRS (`TRSName = Wall`, `ValuesIdx = 1 2 3 4`, `Values = 0 0 0 Tw`, the options of
    the reference's blunt-body cases): nodal states, viscous flux (rhs) and FD Jacobian against the oracle, on hexahedra
    (DiamondVolume3D) and triangles (Corrected2D), then back to the plain extrapolation."""
    for em, cfg, bcs, idx, vals in (
            (meshgen.hex_box(5, 4, 4), ns3d_cfg(), bcs_ns3d(), [1, 2, 3, 4], [0., 0., 0., 1.1]),
            (meshgen.quad_channel(8, 6, triangles=True), ns2d_cfg(derivative="corrected"), bcs_ns2d(), [3, 1], [0.9, 0.05])):
        fm = oracle.build_mesh(em)
        q = cfg.nEqs
        wall = 2                                                    # record index of the no-slip wall in bcs_ns*()
        # socket trsID: a boundary node belongs to the first TRS that lists it (NodalStatesExtrapolator.ci:529-560)
        trs_id = np.full(fm.nNodes, -1)
        for f in range(fm.nInner, fm.nFaces):
            for n in fm.faceNodes[fm.faceNodePtr[f]:fm.faceNodePtr[f+1]]:
                if trs_id[n] == -1:
                    trs_id[n] = fm.faceBC[f]
        flag = (trs_id == wall).astype(np.uint8)
        assert flag.any() and not flag.all()
        U = init_state(fm, q, True)
        base = oracle.compute(cfg, fm, bcs, U.copy(), jacobian=True)
        fm.nodePresc = flag
        fm.nodePrescIdx = np.asarray(idx, dtype=np.int32)
        fm.nodePrescVal = np.asarray(vals, dtype=np.float64)
        r = oracle.compute(cfg, fm, bcs, U.copy(), jacobian=True)
        assert r.status == 0
        ns = r.nstates.reshape(-1, q)
        assert all((ns[flag == 1, i] == v).all() for i, v in zip(idx, vals))
        assert not np.array_equal(r.rhs, base.rhs)                  # the wall nodes feed the viscous flux
        e = engine_factory(cfg)
        try:
            e.set_mesh(fm, bcs)
            e.set_nodal_prescribed(flag, idx, vals)
            e.set_states(U[:fm.nCells*q])
            rhs, uc = e.compute_rhs()
            assert np.array_equal(e.field("nstates"), r.nstates)
            assert np.array_equal(rhs, r.rhs) and np.array_equal(uc, r.updateCoeff)
            assert np.array_equal(e.bsr_values(), r.bsrVal)
            e.set_nodal_prescribed(None)
            e.set_update_coeff()
            rhs, uc = e.compute_rhs()
            assert np.array_equal(e.field("nstates"), base.nstates)
            assert np.array_equal(rhs, base.rhs) and np.array_equal(e.bsr_values(), base.bsrVal)
        finally:
            e.close()


def check_time_term_variants(oracle, engine_factory):
    """a21 FVMCC_PseudoSteadyTimeRhs in all its branches (steady updateCoeff/CFL, dt > 0, useGlobalDT, zeroDiagValue,
    numerical / analytical dU/dP), as a separate call (cfb200_time_rhs) and fused into compute_rhs (cfg.fuseTimeRhs):
    rhs and the block Jacobian bit-identical to the oracle's compute + time_rhs in every variant."""
    variants = [
        ("ns3d_hex", {}),
        ("ns3d_hex", {"dt": 1e-3}),
        ("ns3d_hex", {"timeAnalyticalMatrix": 1}),
        ("ns3d_hex", {"timeAnalyticalMatrix": 1, "zero": [0, 1, 0, 0, 1]}),
        ("ns3d_hex", {"timeUseGlobalDT": 1}),
        ("ns2d_quad", {"timeAnalyticalMatrix": 1, "dt": 2e-3}),
        ("euler2d_2nd_quad", {"zero": [0, 0, 1, 0]}),
        ("euler2d_2nd_quad", {"timeAnalyticalMatrix": 1}),
        ("mhd3d_hex_hll", {"zero": [0, 0, 0, 0, 0, 0, 0, 0, 1]}),     # the phi equation of the divergence cleaning
        ("mhd3d_hex_hll", {"timeAnalyticalMatrix": 1, "zero": [0, 0, 0, 0, 0, 0, 0, 0, 1]}),
    ]
    for name, opts in variants:
        mesh_f, cfg_f, bcs_f, pvt = CASES[name]
        fm = oracle.build_mesh(mesh_f())
        bcs = bcs_f()

        def cfg_with(fuse):
            c = cfg_f()
            c.cfl = 2.5
            for k, v in opts.items():
                if k == "zero":
                    for i, z in enumerate(v):
                        c.timeZeroDiag[i] = z
                else:
                    setattr(c, k, v)
            c.fuseTimeRhs = fuse
            return c
        q = cfg_f().nEqs
        U = init_state(fm, q, pvt)
        past = U[:fm.nCells*q]*(1 + 1e-3)
        Ug = U.copy()
        c0 = cfg_with(0)
        r = oracle.compute(c0, fm, bcs, Ug, jacobian=True)
        oracle.time_rhs(c0, fm, bcs, Ug, np.concatenate([past, np.zeros(fm.nGhosts*q)]), r)
        for fuse in (0, 1):
            if fuse and opts.get("timeUseGlobalDT"):
                continue
            e = engine_factory(cfg_with(fuse))
            try:
                e.set_mesh(fm, bcs)
                e.set_states(U[:fm.nCells*q])
                e.set_past_states(past)
                rhs, uc = e.compute_rhs()
                if not fuse:
                    rhs = e.time_rhs()
                jac = e.bsr_values()
            finally:
                e.close()
            assert np.array_equal(rhs, r.rhs), (name, opts, fuse, "rhs")
            assert np.array_equal(uc, r.updateCoeff), (name, opts, fuse, "updateCoeff")
            assert np.array_equal(jac, r.bsrVal), (name, opts, fuse, "jacobian")


def check_renumbering_layer(oracle, engine_factory, names=("ns3d_hex", "ns3d_tet", "euler2d_2nd_tri_fullstencil", "ns2d_quad_corrected",
                                                            "mhd3d_hex_hll", "euler2d_ausmplusup_analytic", "ns3d_hex_roe_efix2", "ns3d_hex_roesa_fix2", "ns3d_hex_correctedgg", "ns2d_tri_correctedgg_weights")):
    """cfg.renumber: the engine stores cells along a space-filling curve and faces by their cells, yet everything it returns
    is in the caller's numbering and bit-identical to the oracle on the caller's (here: randomly shuffled) numbering --
    ghost states, gradients, limiter, nodal states, residual, update coefficient, divB, the BSR pattern and every block,
    the pseudo-steady term; then a ForwardEuler update and its ordered residual norm."""
    for name in names:
        def mut(cfg):
            cfg.renumber = 1
        out = run_case(name, oracle, engine_factory, cfg_mut=mut, shuffle=True)
        for k, ab in out.items():
            if k.startswith("_") or ab[1] is None:
                continue
            assert np.array_equal(ab[0], ab[1]), (name, k)
    # update + norm: renumbered vs plain storage on the same shuffled mesh
    mesh_f, cfg_f, bcs_f, pvt = CASES["euler2d_2nd_tri_fullstencil"]
    em, _ = meshgen.shuffle_cells(mesh_f())
    fm = oracle.build_mesh(em)
    res = []
    for ren in (0, 1):
        cfg = cfg_f()
        cfg.computeJacobian = 0
        cfg.renumber = ren
        e = engine_factory(cfg)
        e.set_mesh(fm, bcs_f())
        e.set_states(init_state(fm, 4, pvt)[:fm.nCells*4])
        h = []
        for _ in range(4):
            e.compute_rhs(apply_bc=True, download=False)
            h.append(e.update_sol_fwd_euler(0.5, 0))
        res.append((h, e.get_states()))
        e.close()
    assert res[0][0] == res[1][0] and np.array_equal(res[0][1], res[1][1])


def check_ghost_reposition_is_reported(oracle, engine_factory):
    from coolfluid_b200 import lib as cl
    fm = oracle.build_mesh(meshgen.hex_box(4, 4, 3))
    e = engine_factory(ns3d_cfg())
    try:
        e.set_mesh(fm, bcs_ns3d_isothermal(t_wall=0.3, t_ghost_min=0.5))      # 2*0.3 - ~1.0 < 0.5
        e.set_states(init_state(fm, 5, True)[:fm.nCells*5])
        try:
            e.compute_rhs()
            raise AssertionError("no error reported")
        except cl.Cfb200Error as err:
            assert err.code == 8
    finally:
        e.close()
