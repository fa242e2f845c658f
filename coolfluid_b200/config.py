"""Run configuration of the FVMCC ComputeRHS path: the values the reference reads from a .CFcase
file and from SubSystemStatus, flattened into the C struct ``cfb200_config`` (include/cfb200.h).

``ns_reference_values`` restates NavierStokesPhysicalModel<DIM>::setReferenceValuesCompressible
(/root/reference/plugins/NavierStokes/NavierStokesPhysicalModel.ci:69-170) on the host: it is
option processing, not hot-path arithmetic.
"""
from __future__ import annotations
import ctypes as C
import math

MAXEQ = 9

# enums (must match include/cfb200.h and oracle/cfo_types.hpp)
ABI_VERSION = 3                      # CFB200_ABI_VERSION of include/cfb200.h (the struct mirrors below follow it)
PM_EULER, PM_MHD3D_PROJECTION = 0, 1
ST_NONE, ST_MHD_CONS_ACA = 0, 1      # Data.SourceTerm: none, MHDConsACAST (phi damping)
DV_DIAMOND, DV_CORRECTED, DV_STATEDIFF, DV_CORRECTED_GG = 0, 1, 2, 3
UV_CONS, UV_PVT = 0, 1
FS_ROE, FS_AUSMPLUS, FS_AUSMPLUSUP, FS_LAXFRIED, FS_HLL, FS_ROESA, FS_CENTRED = 0, 1, 2, 3, 4, 5, 6
PR_CONSTANT, PR_LINEARLS = 0, 1
LIM_NONE, LIM_VENKTN, LIM_BARTHJESP, LIM_VENKTN_STRICT = 0, 1, 2, 3
DIFF_NONE, DIFF_NS = 0, 1
VISC_SIMPLER_SUTHERLAND, VISC_SUTHERLAND, VISC_FIXED = 0, 1, 2
BC_SUPER_OUTLET, BC_DIRICHLET, BC_MIRROR_VELOCITY, BC_MIRROR_EULER_CONS, BC_FARFIELD_EULER, BC_NOSLIP_ISOTHERMAL, BC_SUPER_OUTLET_MHD, BC_FARFIELD_EULER_CHAR = 0, 1, 2, 3, 4, 5, 6, 7


class ConfigStruct(C.Structure):
    _fields_ = [
        ("dim", C.c_int32), ("nEqs", C.c_int32), ("updateVar", C.c_int32), ("fluxSplitter", C.c_int32),
        ("polyRec", C.c_int32), ("limiter", C.c_int32), ("diffusive", C.c_int32), ("viscLaw", C.c_int32),
        ("computeJacobian", C.c_int32), ("useFullStencil", C.c_int32), ("freezeLimiter", C.c_int32),
        ("stopLimiting", C.c_int32), ("isAdimensional", C.c_int32), ("freezeDiffCoeff", C.c_int32),
        ("iter", C.c_int32), ("limitIter", C.c_int32),
        ("gamma", C.c_double), ("R", C.c_double), ("pOffset", C.c_double),
        ("refValues", C.c_double*MAXEQ), ("fdTol", C.c_double), ("resFactor", C.c_double),
        ("residual", C.c_double), ("limitRes", C.c_double), ("gradientFactor", C.c_double),
        ("coeffEps", C.c_double), ("length", C.c_double), ("magnitudeValues", C.c_double*MAXEQ),
        ("diffRedCoeff", C.c_double),
        ("coeffTau", C.c_double), ("coeffQ", C.c_double), ("cpOverPrandtl", C.c_double),
        ("muRef", C.c_double), ("tempRef", C.c_double), ("pRef", C.c_double), ("artDiffCoeff", C.c_double),
        ("viscRef", C.c_double), ("viscTRef", C.c_double), ("viscSuthConst", C.c_double),
        ("fixedVisc", C.c_double), ("cfl", C.c_double),
        ("choiceA12", C.c_int32), ("hllUseAlpha", C.c_int32),
        ("ausmBeta", C.c_double), ("ausmAlpha", C.c_double),
        ("ausmKu", C.c_double), ("ausmKp", C.c_double), ("ausmSigma", C.c_double), ("machInf", C.c_double),
        ("physicalModel", C.c_int32), ("entropyFixID", C.c_int32), ("refSpeed", C.c_double), ("coeffSigmoid", C.c_double),
        ("derivativeStrategy", C.c_int32), ("fuseTimeRhs", C.c_int32),
        ("timeZeroDiag", C.c_int32*MAXEQ), ("timeAnalyticalMatrix", C.c_int32), ("timeUseGlobalDT", C.c_int32),
        ("renumber", C.c_int32), ("dt", C.c_double),
        ("venktnIsMFMHD", C.c_int32), ("psiMinEqual1", C.c_int32), ("strictCoeff", C.c_double),
        ("roeSaOldStencil", C.c_int32), ("mhdPotentialBType", C.c_int32),
        ("mhdDipole", C.c_double*3), ("noLimiterVar", C.c_int32*MAXEQ), ("correctedGGUseWeights", C.c_int32),
        ("sourceTerm", C.c_int32), ("mhdDissipCoeff", C.c_double),
    ]


class BCStruct(C.Structure):
    _fields_ = [("type", C.c_int32), ("mask", C.c_int32*MAXEQ), ("value", C.c_double*MAXEQ),
                ("zeroGrad", C.c_int32*MAXEQ)]


def make_bc(bc_type, mask=None, value=None, zero_grad=None, n_eqs=MAXEQ):
    b = BCStruct()
    b.type = bc_type
    for i in range(MAXEQ):
        b.mask[i] = int(mask[i]) if mask is not None and i < len(mask) else 0
        b.value[i] = float(value[i]) if value is not None and i < len(value) else 0.0
        b.zeroGrad[i] = int(zero_grad[i]) if zero_grad is not None and i < len(zero_grad) else 0
    return b


def bc_array(bcs):
    arr = (BCStruct*len(bcs))()
    for i, b in enumerate(bcs):
        arr[i] = b
    return arr


def sutherland_simpler(T, visc_ref=0.000001458, t_ref=110.4):
    """SimplerSutherlandDynViscosity defaults (SimplerSutherlandDynViscosity.cxx)."""
    return visc_ref*T**1.5/(t_ref + T)


def make_config(dim, update_var=UV_CONS, flux=FS_ROE, poly_rec=PR_CONSTANT, limiter=LIM_NONE,
                diffusive=DIFF_NONE, jacobian=False, gamma=1.4, R=287.046, ref_values=None,
                res_factor=1.0, cfl=1.0, limit_res=-2.0, coeff_eps=1.0, length=1.0,
                magnitude_values=None, use_full_stencil=False, **kw):
    """Defaults follow the reference: gamma/RDim EulerTerm.cxx:46-50, tol NumericalJacobian.cxx:31,
    limitRes/gradientFactor/limitIter PolyReconstructor.ci:57-67, coeffEps Venktn2D.cxx:58."""
    c = ConfigStruct()
    c.physicalModel = int(kw.get("physical_model", PM_EULER))
    q = 9 if c.physicalModel == PM_MHD3D_PROJECTION else dim + 2
    c.refSpeed = kw.get("ref_speed", 1.0)
    c.entropyFixID = int(kw.get("entropy_fix_id", 0))
    c.dim, c.nEqs = dim, q
    c.updateVar, c.fluxSplitter, c.polyRec, c.limiter, c.diffusive = update_var, flux, poly_rec, limiter, diffusive
    c.viscLaw = kw.get("visc_law", VISC_SIMPLER_SUTHERLAND)
    c.computeJacobian = (2 if jacobian == "analytic" else 1) if jacobian else 0
    c.coeffSigmoid = kw.get("coeff_sigmoid", 1e-8)
    c.derivativeStrategy = {"diamond": DV_DIAMOND, "corrected": DV_CORRECTED, "statediff": DV_STATEDIFF,
                            "correctedgg": DV_CORRECTED_GG}[kw.get("derivative", "diamond")]
    c.useFullStencil = 1 if use_full_stencil else 0
    c.freezeLimiter = int(kw.get("freeze_limiter", 0))
    c.stopLimiting = int(kw.get("stop_limiting", 0))
    c.isAdimensional = int(kw.get("is_adimensional", 0))
    c.freezeDiffCoeff = int(kw.get("freeze_diff_coeff", 0))
    c.iter = int(kw.get("iter", 1))
    c.limitIter = int(kw.get("limit_iter", 1000000000))
    c.gamma, c.R, c.pOffset = gamma, R, kw.get("p_offset", 0.0)
    rv = ref_values if ref_values is not None else [1.0]*q
    mv = magnitude_values if magnitude_values is not None else rv
    for i in range(MAXEQ):
        c.refValues[i] = float(rv[i]) if i < len(rv) else 1.0
        c.magnitudeValues[i] = float(mv[i]) if i < len(mv) else 1.0
    c.fdTol = kw.get("fd_tol", 10e-7)
    c.resFactor = res_factor
    c.residual = kw.get("residual", 0.0)
    c.limitRes = limit_res
    c.gradientFactor = kw.get("gradient_factor", 1.0)
    c.coeffEps, c.length = coeff_eps, length
    c.diffRedCoeff = kw.get("diff_red_coeff", 1.0)
    c.coeffTau = kw.get("coeff_tau", 1.0)
    c.coeffQ = kw.get("coeff_q", 1.0)
    c.cpOverPrandtl = kw.get("cp_over_prandtl", gamma/(gamma - 1.0)*R/0.72)
    c.muRef = kw.get("mu_ref", 1.0)
    c.tempRef = kw.get("temp_ref", 1.0)
    c.pRef = kw.get("p_ref", 1.0)
    c.artDiffCoeff = kw.get("art_diff_coeff", 1.0)
    c.viscRef = kw.get("visc_ref", 0.000001458)
    c.viscTRef = kw.get("visc_tref", 110.4)
    c.viscSuthConst = kw.get("visc_suth_const", 120.0)
    c.fixedVisc = kw.get("fixed_visc", 1.0)
    c.cfl = cfl
    # AUSM family defaults: AUSMFlux.ci:55, AUSMPlusFlux.ci:17-23, AUSMPlusUpFlux.ci:20-33
    c.choiceA12 = int(kw.get("choice_a12", 1))
    c.hllUseAlpha = int(kw.get("hll_use_alpha", 0))
    c.ausmBeta = kw.get("ausm_beta", 1.0/8.0)
    c.ausmAlpha = kw.get("ausm_alpha", 3.0/16.0)
    c.ausmKu = kw.get("coeff_ku", 0.75)
    c.ausmKp = kw.get("coeff_kp", 0.25)
    c.ausmSigma = kw.get("sigma", 1.0)
    c.machInf = kw.get("mach_inf", 0.0)
    # pseudo-steady time term (FVMCC_StdComputeTimeRhs.cxx:42-57, FVMCC_PseudoSteadyTimeRhs.cxx:40-46)
    c.fuseTimeRhs = int(kw.get("fuse_time_rhs", 0))
    zd = kw.get("time_zero_diag", None)
    for i in range(MAXEQ):
        c.timeZeroDiag[i] = int(zd[i]) if zd is not None and i < len(zd) else 0
    c.timeAnalyticalMatrix = int(kw.get("time_analytical_matrix", 0))
    c.timeUseGlobalDT = int(kw.get("time_use_global_dt", 0))
    c.dt = float(kw.get("dt", 0.0))
    c.renumber = int(kw.get("renumber", 0))
    c.venktnIsMFMHD = int(kw.get("is_mfmhd", 1 if limiter == LIM_VENKTN_STRICT else 0))
    c.psiMinEqual1 = int(kw.get("psi_min_equal_1", 0))
    c.strictCoeff = float(kw.get("strict_coeff", 1.0))
    c.roeSaOldStencil = int(kw.get("roe_sa_old_stencil", 0))
    c.correctedGGUseWeights = int(kw.get("gg_use_weights", 0))
    for i in kw.get("no_limiter_id", ()):            # Venktn2D "2DNoLimiterID"
        c.noLimiterVar[int(i)] = 1
    c.sourceTerm = int(kw.get("source_term", 0))      # ST_MHD_CONS_ACA: MHDConsACAST
    c.mhdDissipCoeff = float(kw.get("mhd_dissip_coeff", 0.0))
    dip = kw.get("mhd_dipole")                       # (mX, mY, mZ): potentialBType = Dipole
    c.mhdPotentialBType = 1 if dip is not None else 0
    for i in range(3):
        c.mhdDipole[i] = float(dip[i]) if dip is not None else 0.0
    return c


def ns_reference_values(gamma=1.4, r_dim=287.046, temp_ref=298.15, mach_inf=0.35, reynolds=76000.0,
                        ref_length=1.0, prandtl=0.72, visc_ref=0.000001458, visc_tref=110.4):
    """Adimensional set-up used by e.g. flatPlate3DFVMBlasius.CFcase (tempRef, machInf, Reynolds given).
    Returns the keyword arguments for :func:`make_config`."""
    u_inf = mach_inf*math.sqrt(gamma*r_dim*temp_ref)
    ref_speed = math.sqrt(r_dim*temp_ref)
    mu_ref = visc_ref*temp_ref**1.5/(visc_tref + temp_ref)     # getDynViscosityDim(pRef, tempRef)
    rho_ref = reynolds*mu_ref/(u_inf*ref_length)
    p_ref = rho_ref*r_dim*temp_ref
    inv_re = 1.0/reynolds
    coeff_tau = math.sqrt(gamma)*mach_inf*inv_re
    coeff_q = coeff_tau*gamma/((gamma - 1.0)*prandtl)
    return dict(R=1.0, gamma=gamma, is_adimensional=1, coeff_tau=coeff_tau, coeff_q=coeff_q,
                cp_over_prandtl=1.0, mu_ref=mu_ref, temp_ref=temp_ref, p_ref=p_ref,
                visc_ref=visc_ref, visc_tref=visc_tref, ref_speed=ref_speed, rho_ref=rho_ref, u_inf=u_inf)
