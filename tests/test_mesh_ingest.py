"""CPU tier, SURVEY.md 8(f) rank 4 (mesh ingest): the readers of the reference's two mesh formats on the reference's own
fixtures (committed as data under tests/golden/, byte for byte the files of plugins/NavierStokes/testcases/):
  * read_thor(jets2DFVM.thor, .SP) produces exactly the element / node / TRS tables of jets2DFVM.CFmesh, the file the
    reference's OWN converter (plugins/THOR2CFmesh) wrote from it: a reference-held pin of the THOR reader's ordering;
  * read_cfmesh(flatPlateQD.CFmesh), the mesh of flatPlateFVMBlasius.CFcase, goes through the host mesh builder and one
    ComputeRHS + Jacobian evaluation of that case's scheme (NavierStokes2D Puvt, Roe + NS flux, LinearLS2D, NumJacob):
    engine (host build) == oracle bit for bit."""
import os

import numpy as np

import cases
from coolfluid_b200 import meshgen, config as B, lib as cl

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
REF = "/root/reference/plugins/NavierStokes/testcases/"


def test_fixtures_are_the_reference_files():
    if not os.path.isdir(REF):
        return
    for name, src in (("flatPlateQD.CFmesh", "FlatPlate/flatPlateQD.CFmesh"), ("jets2DFVM.thor", "Jets2D/jets2DFVM.thor"),
                      ("jets2DFVM.SP", "Jets2D/jets2DFVM.SP"), ("jets2DFVM.CFmesh", "Jets3D/jets2DFVM.CFmesh")):
        assert open(os.path.join(GOLD, name), "rb").read() == open(REF + src, "rb").read(), name


def test_thor_reader_matches_the_reference_converter_output():
    a = meshgen.read_thor(os.path.join(GOLD, "jets2DFVM.thor"), os.path.join(GOLD, "jets2DFVM.SP"))
    b, states = meshgen.read_cfmesh(os.path.join(GOLD, "jets2DFVM.CFmesh"))
    assert states is None and a.trs_names == b.trs_names == ["SuperOutlet", "SuperInlet"]
    assert a.n_cells == 3178 and a.n_nodes == 1670
    for k in ("nodes", "cell_node_ptr", "cell_nodes", "cell_shape", "trs_face_ptr", "trs_face_cell", "trs_face_node_ptr",
              "trs_face_nodes"):
        assert np.array_equal(getattr(a, k), getattr(b, k)), k


def test_flat_plate_cfmesh_through_the_engine(oracle, hostsim_lib):
    em, states = meshgen.read_cfmesh(os.path.join(GOLD, "flatPlateQD.CFmesh"))
    assert em.n_cells == 1850 and em.n_nodes == 1938 and em.trs_names == ["SlipWall", "NoSlipWall", "Inlet", "Outlet"]
    fm = oracle.build_mesh(em)
    assert fm.nGhosts == 174 and (fm.volumes > 0).all()
    # the product's threaded builder gives the same tables
    fm2 = cl.build_mesh(em, cl.load()) if os.path.exists(cl.LIB_PATH) else None
    if fm2 is not None:
        for k in ("faceL", "faceR", "normals", "volumes", "stencil", "bsrColInd"):
            assert np.array_equal(getattr(fm, k), getattr(fm2, k)), k
    cfg = cases.ns2d_cfg()
    wall = B.make_bc(B.BC_DIRICHLET, mask=[0, 1, 1, 0], value=[0, 0, 0, 0])
    inlet = B.make_bc(B.BC_DIRICHLET, mask=[0, 1, 1, 1], value=[0, 0.4, 0, 1.0])
    outlet = B.make_bc(B.BC_DIRICHLET, mask=[1, 0, 0, 0], value=[1.0, 0, 0, 0])
    bcs = [B.make_bc(B.BC_MIRROR_VELOCITY), wall, inlet, outlet]
    U = cases.init_state(fm, 4, True)
    Ug = U.copy()
    r = oracle.compute(cfg, fm, bcs, Ug, jacobian=True)
    assert r.status == 0
    e = cl.Engine(cfg, 0, lib=hostsim_lib)
    try:
        e.set_mesh(fm, bcs)
        e.set_states(U[:fm.nCells*4])
        rhs, uc = e.compute_rhs()
        assert np.array_equal(rhs, r.rhs) and np.array_equal(uc, r.updateCoeff) and np.array_equal(e.bsr_values(), r.bsrVal)
    finally:
        e.close()
