"""GPU tier (-m gpu): the CUDA path through the C ABI (libcfb200.so on cuda:0) against the oracle on the same
seeded inputs.  Tolerance: BASELINE.json asks for 1e-12 relative per-cell residual and Jacobian entry; the kernels
are built without FMA contraction and with the reference's operation order, so the tests demand 1e-13 relative to
the field's max-abs and additionally report whether the match is bit-exact."""
import numpy as np
import pytest

import cases
from coolfluid_b200 import lib as cl, meshgen, config as B

pytestmark = pytest.mark.gpu
TOL = 1e-13


def _factory(product_lib):
    return lambda cfg: cl.Engine(cfg, 0, lib=product_lib)


def _check(out, tol=TOL):
    worst = 0.0
    for k, ab in out.items():
        if k.startswith("_") or ab[1] is None:
            continue
        a, b = ab
        assert a.shape == b.shape, k
        if a.dtype.kind == "i":
            assert np.array_equal(a, b), k
            continue
        err = cases.relerr(a, b)
        worst = max(worst, err)
        assert err <= tol, "%s: rel err %.3e" % (k, err)
    assert np.abs(out["rhs"][1]).max() > 0.0
    return worst


@pytest.mark.parametrize("name", sorted(cases.CASES))
def test_cuda_matches_oracle(name, oracle, product_lib):
    _check(cases.run_case(name, oracle, _factory(product_lib)))


def test_cuda_bit_exact_on_flagship_case(oracle, product_lib):
    """3D NS hex (BASELINE config 3) is expected to match the oracle bit for bit (no FMA, IEEE div/sqrt)."""
    out = cases.run_case("ns3d_hex", oracle, _factory(product_lib))
    for k in ("rhs", "updateCoeff", "jacobian", "time_jacobian"):
        assert np.array_equal(out[k][0], out[k][1]), k


def test_partition_overlap_rows(oracle, product_lib):
    def upd(fm):
        x = fm.centres.reshape(-1, fm.dim)[:fm.nCells, 0]
        return (x < 0.6).astype(np.uint8)
    for name in ("ns3d_hex", "euler2d_2nd_tri_fullstencil"):
        out = cases.run_case(name, oracle, _factory(product_lib), updatable=upd)
        _check(out)
        fm = out["_fm"]
        jac = out["jacobian"][0].reshape(len(fm.bsrColInd), -1)
        for c in np.nonzero(fm.updatable == 0)[0]:
            assert not jac[fm.bsrRowPtr[c]:fm.bsrRowPtr[c+1]].any()


def test_historical_limiter(oracle, product_lib):
    rng = np.random.RandomState(3)

    def mut(cfg):
        cfg.residual = -3.0
    n = cases.CASES["ns3d_hex"][0]().n_cells*5
    lim0 = 0.3 + 0.7*rng.random_sample(n)
    _check(cases.run_case("ns3d_hex", oracle, _factory(product_lib), limiter_init=lim0, cfg_mut=mut))


def _medium(oracle, product_lib, em, cfg, bcs, pvt):
    fm = cl.build_mesh(em, product_lib)
    q = cfg.nEqs
    U = cases.init_state(fm, q, pvt)
    r = oracle.compute(cfg, fm, bcs, U.copy(), jacobian=True, omp=True)
    assert r.status == 0
    e = cl.Engine(cfg, 0, lib=product_lib)
    e.set_mesh(fm, bcs)
    e.set_states(U[:fm.nCells*q])
    rhs, uc = e.compute_rhs()
    jac = e.bsr_values()
    # determinism: a second evaluation is bitwise identical (no atomics, fixed gather order)
    e.set_update_coeff(None)
    rhs2, uc2 = e.compute_rhs()
    jac2 = e.bsr_values()
    e.close()
    assert np.array_equal(rhs, rhs2) and np.array_equal(uc, uc2) and np.array_equal(jac, jac2)
    assert cases.relerr(rhs, r.rhs) <= TOL
    assert cases.relerr(uc, r.updateCoeff) <= TOL
    # per-entry Jacobian error relative to the block's max-abs (SURVEY.md §8d parity gates)
    qq = q*q
    a = jac.reshape(-1, qq); b = r.bsrVal.reshape(-1, qq)
    scale = np.maximum(np.abs(b).max(axis=1, keepdims=True), 1e-300)
    assert (np.abs(a - b)/scale).max() <= 1e-12


def test_medium_hex_ns(oracle, product_lib):
    _medium(oracle, product_lib, meshgen.hex_box(24, 22, 20), cases.ns3d_cfg(), cases.bcs_ns3d(), True)


def test_medium_tet_ns(oracle, product_lib):
    _medium(oracle, product_lib, meshgen.tet_box(10, 9, 8), cases.ns3d_cfg(), cases.bcs_ns3d(), True)


def test_medium_tri_euler(oracle, product_lib):
    cfg = B.make_config(2, poly_rec=B.PR_LINEARLS, limiter=B.LIM_VENKTN, jacobian=True)
    _medium(oracle, product_lib, meshgen.quad_channel(90, 70, triangles=True), cfg, cases.bcs_euler2d(), False)


def test_forward_euler_history(oracle, product_lib):
    """BASELINE config 1/2 explicit loop, device resident (applyBC -> ComputeRHS -> UpdateSol + L2 norm) against the
    oracle's ForwardEuler driver: residual history to 1e-10 (BASELINE.json north_star)."""
    em = meshgen.quad_channel(40, 30, triangles=True)
    fm = cl.build_mesh(em, product_lib)
    cfg = B.make_config(2, poly_rec=B.PR_LINEARLS, limiter=B.LIM_VENKTN, limit_res=-1.42, cfl=0.7)
    bcs = cases.bcs_euler2d()
    U0 = cases.init_state(fm, 4, False)
    nit = 60
    ref = oracle.run_fwd_euler(B.make_config(2, poly_rec=B.PR_LINEARLS, limiter=B.LIM_VENKTN, limit_res=-1.42, cfl=0.7),
                               fm, bcs, U0.copy(), 0.7, -99.0, nit, omp=True)
    e = cl.Engine(cfg, 0, lib=product_lib)
    e.set_mesh(fm, bcs)
    e.set_states(U0[:fm.nCells*4])
    hist = []
    cfg.residual = 0.0
    for it in range(nit):
        cfg.iter = it + 1
        e.set_step(cfg)
        e.compute_rhs(apply_bc=True, download=False)
        res = e.update_sol_fwd_euler(0.7, 0)
        hist.append(res)
        cfg.residual = res
    e.close()
    assert np.abs(np.array(hist) - ref.history).max() <= 1e-10


def test_position_dependent_inlet(oracle, product_lib):
    cases.check_bc_face_values(oracle, _factory(product_lib))


def test_newton_update(product_lib):
    cases.check_newton_update(_factory(product_lib), lambda em: cl.build_mesh(em, product_lib))


def test_large_size_properties(product_lib):
    """At a size the oracle does not finish quickly: size-independent properties.  (i) determinism; (ii) block-row
    sums: every flux is added to L and subtracted from R, so over a mesh whose cells are all updatable the column
    sums of the Jacobian restricted to interior-face terms vanish: sum_rows J[:, c] == boundary blocks only, i.e.
    for interior cells (no boundary face) diag(c) + sum_n block(n, c) == 0 exactly."""
    em = meshgen.hex_box(64, 64, 48)
    fm = cl.build_mesh(em, product_lib)
    cfg = cases.ns3d_cfg()
    e = cl.Engine(cfg, 0, lib=product_lib)
    e.set_mesh(fm, cases.bcs_ns3d())
    U = cases.init_state(fm, 5, True)
    e.set_states(U[:fm.nCells*5])
    rhs, uc = e.compute_rhs()
    jac = e.bsr_values().reshape(-1, 25)
    assert np.isfinite(rhs).all() and np.isfinite(jac).all() and (uc > 0).all()
    rp, ci = e.bsr_pattern()
    rows = np.repeat(np.arange(fm.nCells), np.diff(rp))
    colsum = np.zeros((fm.nCells, 25))
    np.add.at(colsum, ci, jac)
    interior = np.ones(fm.nCells, dtype=bool)
    interior[fm.faceL[fm.nInner:]] = False
    scale = np.abs(jac).max()
    assert np.abs(colsum[interior]).max() <= 1e-12*scale
    # (i) a second evaluation of the same state is bit-identical (no atomics, fixed gather order)
    e.set_update_coeff(None)          # the updateCoeff socket accumulates until the updater zeroes it
    rhs2, uc2 = e.compute_rhs()
    assert np.array_equal(rhs, rhs2) and np.array_equal(uc, uc2)
    assert np.array_equal(jac, e.bsr_values().reshape(-1, 25))
    e.close()


def test_full_size_freestream_and_determinism(product_lib):
    """BASELINE config 3 at its full size (220^3 = 10.6 M hexahedra): a uniform state is preserved by every cell
    without a boundary face (closed control volumes, zero gradients: the residual is rounding noise of the flux sums),
    and two evaluations of a non-uniform state give bit-identical residuals."""
    n = 220
    fm = cl.build_mesh(meshgen.hex_box(n, n, n), product_lib)
    cfg = cases.ns3d_cfg()
    e = cl.Engine(cfg, 0, lib=product_lib)
    e.set_mesh(fm, cases.bcs_ns3d())
    U = np.empty((fm.nCells, 5))
    U[:] = [1.0, 0.4, 0.03, 0.02, 1.0]                    # p u v w T in the units of cases.ns3d_cfg()
    e.set_states(U.reshape(-1))
    rhs, uc = e.compute_rhs()
    # two cell layers next to the boundary see the ghost states through their gradients
    x = fm.centres.reshape(-1, 3)[:fm.nCells]
    interior = ((x > 2.0/n) & (x < 1.0 - 2.0/n)).all(axis=1)
    assert interior.sum() == (n - 4)**3
    flux_scale = 3.5*1.0*0.4*(1.0/n)**2                   # ~ rho u H * face area
    assert np.abs(rhs.reshape(fm.nCells, 5)[interior]).max() <= 1e-11*flux_scale
    assert (uc > 0).all()
    U2 = cases.init_state(fm, 5, True)
    e.set_states(U2[:fm.nCells*5])
    e.set_update_coeff(None)
    a, ua = e.compute_rhs()
    e.set_update_coeff(None)
    b, ub = e.compute_rhs()
    assert np.isfinite(a).all() and np.array_equal(a, b) and np.array_equal(ua, ub)
    e.close()


def test_cuda_matches_committed_golden_vectors(product_lib):
    """The CUDA path against tests/golden/cases.npz (vectors generated by tests/golden/make_golden.py): no oracle
    run involved on the GPU box."""
    import os
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "cases.npz"))
    for name in sorted(cases.CASES):
        mesh_f, cfg_f, bcs_f, pvt = cases.CASES[name]
        fm = cl.build_mesh(mesh_f(), product_lib)
        c = cfg_f()
        U = cases.init_state(fm, c.nEqs, pvt)
        uc0 = 0.1*np.arange(fm.nCells, dtype=np.float64)/max(fm.nCells, 1)
        e = cl.Engine(c, 0, lib=product_lib)
        e.set_mesh(fm, bcs_f())
        e.set_states(U[:fm.nCells*c.nEqs])
        e.set_update_coeff(uc0)
        rhs, uc = e.compute_rhs()
        got = {"rhs": rhs, "updateCoeff": uc, "limiter": e.field("limiter"), "jacobian": e.bsr_values()}
        e.close()
        for k, a in got.items():
            assert cases.relerr(a, g[name + "/" + k]) <= TOL, (name, k)


def test_div_sqrt_bitexact(product_lib):
    """The kernels' branch-free FP64 divide / square root (csrc/cfb_math.cuh) against the compiler's IEEE `/` and
    sqrt() on the device: 2^28 operand pairs per exponent span, zeros, exact quotients and shared reciprocals included.
    Bar: zero mismatching bit patterns."""
    import ctypes as C
    for span, seed in ((4, 1), (60, 2), (300, 3), (480, 4)):
        bad = C.c_int64(-1)
        rc = product_lib.selftest_divsqrt(0, 1 << 28, seed, span, C.byref(bad))
        assert rc == 0
        assert bad.value == 0, "span 2^+-%d: %d mismatches" % (span, bad.value)
