"""CPU tier: the product's kernel functors + engine sequencing, compiled for the host by tests/hostsim (TEST ONLY),
against the oracle.  This checks the kernel LOGIC (gather orders, slots, signs, operation order) without a GPU;
the same cases run on the B200 through the C ABI in test_gpu_parity.py.  Bar: bit-exact (the FD Jacobian
amplifies flux rounding by ~1e6, so the 1e-12 target of BASELINE.json is only robust if fluxes agree bit for bit)."""
import numpy as np
import pytest

import cases
from coolfluid_b200 import lib as cl


def _engine_factory(hostsim_lib):
    return lambda cfg: cl.Engine(cfg, 0, lib=hostsim_lib)


def _check(out):
    for k, ab in out.items():
        if k.startswith("_") or ab[1] is None:
            continue
        a, b = ab
        assert a.shape == b.shape, k
        assert np.array_equal(a, b), "%s differs: rel err %.3e" % (k, cases.relerr(a, b))
    assert np.abs(out["rhs"][1]).max() > 0.0


@pytest.mark.parametrize("name", sorted(cases.CASES))
def test_kernel_logic_bit_exact(name, oracle, hostsim_lib):
    _check(cases.run_case(name, oracle, _engine_factory(hostsim_lib)))


def test_partition_overlap_rows(oracle, hostsim_lib):
    """Non-updatable (overlap) cells: faces with no updatable side are skipped, rows of non-updatable cells stay
    zero, one-sided faces fill only the updatable row (FVMCC_ComputeRhsJacob.cxx:108-119,184-236)."""
    def upd(fm):
        x = fm.centres.reshape(-1, fm.dim)[:fm.nCells, 0]
        return (x < 0.6).astype(np.uint8)
    for name in ("ns3d_hex", "euler2d_2nd_tri_fullstencil", "euler2d_ausmplusup_analytic", "mhd3d_hex_hll",
                 "ns3d_hex_roesa_fix2", "ns2d_quad_ausmplusup_analytic"):
        out = cases.run_case(name, oracle, _engine_factory(hostsim_lib), updatable=upd)
        _check(out)
        fm = out["_fm"]
        jac = out["jacobian"][0].reshape(len(fm.bsrColInd), -1)
        for c in np.nonzero(fm.updatable == 0)[0]:
            assert not jac[fm.bsrRowPtr[c]:fm.bsrRowPtr[c+1]].any()


def test_historical_and_frozen_limiter(oracle, hostsim_lib):
    """FVMCC_PolyRec.cxx:131-165: residual <= limitRes -> limiter = min(new, stored); freezeLimiter keeps it."""
    rng = np.random.RandomState(3)
    for freeze in (0, 1):
        def mut(cfg):
            cfg.residual = -3.0
            cfg.freezeLimiter = freeze
        mesh_f = cases.CASES["ns3d_hex"][0]
        n = mesh_f().n_cells*5
        lim0 = 0.3 + 0.7*rng.random_sample(n)
        out = cases.run_case("ns3d_hex", oracle, _engine_factory(hostsim_lib), limiter_init=lim0, cfg_mut=mut)
        _check(out)
        if freeze:
            assert np.array_equal(out["limiter"][0], lim0)
        else:
            assert (out["limiter"][0] <= lim0).all() and (out["limiter"][0] < lim0).any()


def test_stop_limiting(oracle, hostsim_lib):
    def mut(cfg):
        cfg.stopLimiting = 1
    _check(cases.run_case("euler2d_2nd_quad", oracle, _engine_factory(hostsim_lib), cfg_mut=mut))


def test_bad_value_is_reported(oracle, hostsim_lib):
    """a^2 < 0 in the Roe linearisation -> CFB200_ERR_BAD_VALUE (BadValueException in the reference)."""
    from coolfluid_b200 import meshgen, config as B
    for jac in (False, True):       # the pipelined finite-difference loop carries the flag instead of branching
        fm = oracle.build_mesh(meshgen.quad_channel(4, 4))
        cfg = B.make_config(2, jacobian=jac)
        U = cases.init_state(fm, 4, False)
        U[4*5 + 3] = -50.0          # negative total energy in cell 5
        e = cl.Engine(cfg, 0, lib=hostsim_lib)
        e.set_mesh(fm, cases.bcs_euler2d())
        e.set_states(U[:fm.nCells*4])
        with pytest.raises(cl.Cfb200Error) as ei:
            e.compute_rhs()
        assert ei.value.code == 1
        e.close()


def test_time_term_variants_separate_and_fused(oracle, hostsim_lib):
    cases.check_time_term_variants(oracle, _engine_factory(hostsim_lib))


def test_renumbering_layer_is_a_pure_permutation(oracle, hostsim_lib):
    cases.check_renumbering_layer(oracle, _engine_factory(hostsim_lib))


def test_isothermal_wall_reports_ghost_reposition(oracle, hostsim_lib):
    """ghost T = 2 TWall - T_inner below TGhostMin: the reference moves the ghost node (repositionNode); the engine
    reports CFB200_ERR_GHOST_REPOSITION instead of returning something else."""
    cases.check_ghost_reposition_is_reported(oracle, _engine_factory(hostsim_lib))


def test_newton_update(hostsim_lib):
    from oracle import binding as ob
    cases.check_newton_update(_engine_factory(hostsim_lib), ob.build_mesh)


def test_position_dependent_inlet(oracle, hostsim_lib):
    cases.check_bc_face_values(oracle, _engine_factory(hostsim_lib))


def test_nodal_values_prescribed_on_wall_nodes(oracle, hostsim_lib):
    cases.check_nodal_prescribed(oracle, _engine_factory(hostsim_lib))


def test_set_step_cannot_change_the_derivative_strategy(oracle, hostsim_lib):
    """The boundary gradient stencil of Corrected2D/3D is built at set_mesh: switching afterwards is refused."""
    from coolfluid_b200 import meshgen
    fm = oracle.build_mesh(meshgen.hex_box(3, 3, 2))
    e = cl.Engine(cases.ns3d_cfg(), 0, lib=hostsim_lib)
    e.set_mesh(fm, cases.bcs_ns3d())
    with pytest.raises(cl.Cfb200Error) as ei:
        e.set_step(cases.ns3d_cfg(derivative="corrected"))
    assert ei.value.code == 2
    e.set_step(cases.ns3d_cfg(gradient_factor=0.5, res_factor=0.3))       # per-step scalars may change
    e.close()
