"""Generates the committed golden fixtures of tests/golden/ (run in the build container, where /root/reference is
mounted):

  wedge_history.npz   residual history of the reference's own regression case
                      plugins/NavierStokes/testcases/Wedge/wedgeFVM.CFcase (Euler2D, Roe, LinearLS2D + Venktn2D,
                      ForwardEuler CFL 0.7, stop at Norm -4.0) computed by the oracle on the reference's mesh
                      wedge.thor/wedge.SP; the reference pins the final value to `### Residual = -4.0017502` at 0.1 %.
  cases.npz           rhs / updateCoeff / limiter / Jacobian of the small parity cases of tests/cases.py computed by
                      the golden (-O2, no FP contraction) oracle build: guards the oracle itself against drift and
                      gives the GPU tests vectors that do not depend on running the oracle.

usage: python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import cases  # noqa: E402
from coolfluid_b200 import meshgen, config as B  # noqa: E402
from oracle import binding as ob  # noqa: E402

# the reference's mesh of the case (plugins/NavierStokes/testcases/Wedge/wedge.thor, wedge.SP), committed beside this script
# as a data fixture so that the GPU box (no /root/reference) can run the case
WEDGE = HERE + "/"
WEDGE_INLET = [1., 2.366431913, 0.0, 5.3]          # wedgeFVM.CFcase InField.Def / Inlet.Def


def wedge_setup():
    em = meshgen.read_thor(WEDGE + "wedge.thor", WEDGE + "wedge.SP")
    cfg = B.make_config(2, poly_rec=B.PR_LINEARLS, limiter=B.LIM_VENKTN, limit_res=-1.42, coeff_eps=1.0, cfl=0.7)
    bcmap = {"SlipWall": B.make_bc(B.BC_MIRROR_EULER_CONS),
             "SuperInlet": B.make_bc(B.BC_DIRICHLET, mask=[1, 1, 1, 1], value=WEDGE_INLET),
             "SuperOutlet": B.make_bc(B.BC_SUPER_OUTLET)}
    return em, cfg, bcmap


def main():
    ob.build()
    em, cfg, bcmap = wedge_setup()
    fm = ob.build_mesh(em)
    bcs = [bcmap[n] for n in fm.trs_names]
    U = np.zeros((fm.nCells + fm.nGhosts, 4))
    U[:fm.nCells] = WEDGE_INLET
    U = U.reshape(-1).copy()
    r = ob.run_fwd_euler(cfg, fm, bcs, U, 0.7, -4.0, 5000)
    print("wedge: %d iterations, final residual %.7f (reference pins -4.0017502)" % (len(r.history), r.history[-1]))
    np.savez_compressed(os.path.join(HERE, "wedge_history.npz"), history=r.history, n_cells=fm.nCells,
                        reference_residual=-4.0017502)
    out = {}
    for name in sorted(cases.CASES):
        mesh_f, cfg_f, bcs_f, pvt = cases.CASES[name]
        fm = ob.build_mesh(mesh_f())
        c = cfg_f()
        U = cases.init_state(fm, c.nEqs, pvt)
        uc0 = 0.1*np.arange(fm.nCells, dtype=np.float64)/max(fm.nCells, 1)
        r = ob.compute(c, fm, bcs_f(), U, jacobian=True, update_coeff_init=uc0)
        assert r.status == 0
        out[name + "/rhs"] = r.rhs
        out[name + "/updateCoeff"] = r.updateCoeff
        out[name + "/limiter"] = r.limiter
        out[name + "/jacobian"] = r.bsrVal
    np.savez_compressed(os.path.join(HERE, "cases.npz"), **out)
    print("cases:", len(out), "arrays")


if __name__ == "__main__":
    main()
