"""GPU tier, collected last on purpose: the allocation-free residual norm `cfb200_l2_norm_noalloc` (no counterpart in the
reference; its scratch is a file-scope device array of the library) against the existing entry point
`cfb200_update_sol_newton`, which forms the same norm with the engine's cudaMalloc'ed scratch.  Both go through the C ABI.
Bar: the sum of squares has the same reduction shape, so it is compared bit for bit through the log10 both report
(1e-12: device log10/sqrt against the host's libm), on two calls in a row (the scratch keeps the previous contents)."""
import ctypes as C

import numpy as np
import pytest

import cases
from coolfluid_b200 import lib as cl, config as B, meshgen

pytestmark = pytest.mark.gpu

TOL = 1e-12


def _noalloc(L, d_dU, d_upd, n, q, var, d_out, device=0):
    err = C.create_string_buffer(512)
    rc = L.l2_norm_noalloc(device, d_dU, d_upd, n, q, var, None, d_out, err, len(err))
    return rc, err.value.decode()


def test_noalloc_norm_matches_update_sol_newton(oracle, product_lib):
    import torch
    L = product_lib
    fm = oracle.build_mesh(meshgen.hex_box(13, 11, 9))            # 1287 cells: two blocks of partial sums, one ragged
    cfg = cases.ns3d_cfg()
    q = cfg.nEqs
    n = fm.nCells
    e = cl.Engine(cfg, 0, lib=L)
    try:
        e.set_mesh(fm, cases.bcs_ns3d())
        U = cases.init_state(fm, q, True)
        e.set_states(U[:n*q])
        out = torch.zeros(2, dtype=torch.float64, device="cuda:0")
        rc, msg = _noalloc(L, None, None, 0, q, 0, None)          # set-up call: resolves the symbol address
        assert rc == 0, msg
        rng = np.random.RandomState(11)
        zero = np.zeros(q)
        for call in range(2):                                      # two calls in a row share the scratch array
            dU = rng.standard_normal(n*q)*10.0**rng.randint(-6, 3)
            var = (1 + 2*call) % q
            ref = e.update_sol_newton(dU, alpha=zero, monitor_var=var)      # existing entry point (alpha = 0: U untouched)
            d_rhs = L.device_ptr(e.h, b"rhs")                      # the increments it reduced, on the device
            assert d_rhs
            rc, msg = _noalloc(L, d_rhs, None, n, q, var, out.data_ptr())
            assert rc == 0, msg
            got = out.cpu().numpy()
            assert abs(got[1] - ref) <= TOL*max(1.0, abs(ref)), (call, got, ref)
            sq = (dU.reshape(n, q)[:, var]**2).sum()
            assert abs(got[0] - sq) <= 1e-12*sq
            assert abs(got[0] - 10.0**(2.0*ref)) <= 1e-10*got[0]
            # the updatable mask: cells masked out do not count (rows of non-updatable cells, StdUpdateSol.cxx:110-141)
            mask = (rng.random_sample(n) < 0.7).astype(np.uint8)
            d_mask = torch.from_numpy(mask).to("cuda:0")
            rc, msg = _noalloc(L, d_rhs, d_mask.data_ptr(), n, q, var, out.data_ptr())
            assert rc == 0, msg
            got = out.cpu().numpy()
            sqm = (dU.reshape(n, q)[mask != 0, var]**2).sum()
            assert abs(got[0] - sqm) <= 1e-12*sqm
            assert abs(got[1] - np.log10(np.sqrt(sqm))) <= TOL*max(1.0, abs(ref))
    finally:
        e.close()


def test_noalloc_two_level_and_limits(product_lib):
    """More than 1024 block partials (the second reduction level) and the size limit."""
    import torch
    L = product_lib
    n, q, var = 1024*1024 + 4097, 2, 1
    g = torch.Generator(device="cpu").manual_seed(5)
    dU = torch.randn(n*q, dtype=torch.float64, generator=g)
    d = dU.to("cuda:0")
    out = torch.zeros(2, dtype=torch.float64, device="cuda:0")
    for _ in range(2):
        rc, msg = _noalloc(L, d.data_ptr(), None, n, q, var, out.data_ptr())
        assert rc == 0, msg
        got = out.cpu().numpy()
        sq = float((dU.view(n, q)[:, var]**2).sum())
        assert abs(got[0] - sq) <= 1e-12*sq
        assert abs(got[1] - np.log10(np.sqrt(sq))) <= TOL
    rc, msg = _noalloc(L, d.data_ptr(), None, 16777216 + 1, q, var, out.data_ptr())
    assert rc == 2 and "CFB200_NOALLOC_MAX_CELLS" in msg
    rc, msg = _noalloc(L, d.data_ptr(), None, n, q, q, out.data_ptr())
    assert rc == 2 and msg
    rc, msg = _noalloc(L, None, None, n, q, var, out.data_ptr())
    assert rc == 2 and msg
