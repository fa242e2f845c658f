"""Size-independent properties of the path, checked on the oracle (CPU tier); tests/test_gpu_parity.py checks the
same properties on the CUDA path at BASELINE sizes."""
import numpy as np

import cases
from coolfluid_b200 import meshgen
from oracle import binding as ob


def test_oracle_preserves_a_uniform_state():
    """Closed control volumes + zero gradients: away from the boundary (two cell layers see the ghost states through
    their least-squares stencils) the residual of a uniform state is rounding noise of the face-flux sums."""
    n = 14
    fm = ob.build_mesh(meshgen.hex_box(n, n, n))
    U = np.zeros((fm.nCells + fm.nGhosts, 5))
    U[:fm.nCells] = [1.0, 0.4, 0.03, 0.02, 1.0]
    r = ob.compute(cases.ns3d_cfg(), fm, cases.bcs_ns3d(), U.reshape(-1))
    x = fm.centres.reshape(-1, 3)[:fm.nCells]
    interior = ((x > 2.0/n) & (x < 1.0 - 2.0/n)).all(axis=1)
    assert interior.sum() == (n - 4)**3
    rhs = r.rhs.reshape(fm.nCells, 5)
    assert np.abs(rhs[interior]).max() <= 1e-11*3.5*0.4/n**2
    assert np.abs(rhs[~interior]).max() > 1e-6          # the boundary layers do see the boundary conditions
