"""Host mesh-data construction of the product (cfb200_mesh_build: hash-based face matcher, threaded loops) against
the oracle's literal restatement of the reference's FVMCC_MeshDataBuilder + geometry + stencil algorithms.
Integer tables must be identical (IDs are the parity key) and geometry bit-exact."""
import numpy as np
import pytest

from coolfluid_b200 import meshgen, lib as cl

MESHES = {
    "hex": lambda: meshgen.hex_box(9, 7, 5),
    "hex_flat": lambda: meshgen.hex_box(6, 5, 1, jitter=0.0),
    "tet": lambda: meshgen.tet_box(5, 4, 3),
    "quad": lambda: meshgen.quad_channel(13, 9),
    "tri": lambda: meshgen.quad_channel(11, 8, triangles=True),
    "one_hex": lambda: meshgen.hex_box(1, 1, 1),
    "one_quad": lambda: meshgen.quad_channel(1, 1),
}


@pytest.mark.parametrize("name", sorted(MESHES))
def test_mesh_data_identical(name, oracle, product_lib):
    em = MESHES[name]()
    a = oracle.build_mesh(em)
    b = cl.build_mesh(em, product_lib)
    assert (a.nFaces, a.nInner, a.nGhosts) == (b.nFaces, b.nInner, b.nGhosts)
    assert a.equal_to(b) == []


@pytest.mark.parametrize("kind", ["FaceVertex", "FaceEdge", "Face"])
@pytest.mark.parametrize("name", ["hex", "tet", "quad", "tri", "one_hex"])
def test_stencil_variants_identical(name, kind, oracle, product_lib):
    """Setup1.stencil = FaceVertex | FaceEdge | Face (cfb200_mesh_set_stencil: sort + run lengths over the cells around
    the nodes) against the oracle's literal restatement of ComputeFaceVertexNeighbors / ComputeFaceEdgeNeighbors /
    ComputeFaceNeighbors; everything but the stencil stays what the build left."""
    em = MESHES[name]()
    a = oracle.build_mesh(em, stencil=kind)
    b = cl.build_mesh(em, product_lib, stencil=kind)
    assert a.equal_to(b) == []
    base = cl.build_mesh(em, product_lib)
    assert b.stencil.max(initial=-1) < b.nCells                      # no ghost states in these
    if kind == "FaceVertex":                                         # = FaceVertexPlusGhost without the ghosts
        for e in range(b.nCells):
            full = base.stencil[base.stencilPtr[e]:base.stencilPtr[e+1]]
            assert sorted(full[full < b.nCells]) == list(b.stencil[b.stencilPtr[e]:b.stencilPtr[e+1]])
    if kind == "Face":                                               # one entry per interior face
        assert len(b.stencil) == b.nInner


def test_stencil_kind_is_validated(product_lib):
    em = meshgen.quad_channel(3, 3)
    h = product_lib.mesh_build(em.dim, em.n_nodes, cl._pd(em.nodes), em.n_cells, cl._pi(em.cell_node_ptr),
                               cl._pi(em.cell_nodes), cl._pi(em.cell_shape), em.n_trs, cl._pi(em.trs_face_ptr),
                               cl._pi(em.trs_face_cell), cl._pi(em.trs_face_node_ptr), cl._pi(em.trs_face_nodes))
    try:
        assert product_lib.mesh_set_stencil(h, 7) == 2 and product_lib.mesh_set_stencil(h, 0) == 2
        assert product_lib.mesh_set_stencil(h, 1) == 0
    finally:
        product_lib.mesh_free(h)


def test_face_orientation_and_closure(product_lib):
    """Size-independent properties: L < R on inner faces (L is the first cell that creates the face,
    FVMCC_MeshDataBuilder.cxx:369-499), and the outward normals of every cell sum to zero."""
    em = meshgen.hex_box(12, 10, 8)
    fm = cl.build_mesh(em, product_lib)
    assert (fm.faceL[:fm.nInner] < fm.faceR[:fm.nInner]).all()
    nrm = fm.normals.reshape(-1, 3)
    acc = np.zeros((fm.nCells, 3))
    np.add.at(acc, fm.faceL[:fm.nInner], nrm[:fm.nInner])
    np.add.at(acc, fm.faceR[:fm.nInner], -nrm[:fm.nInner])
    # boundary quad normals are replaced by the 4-triangle average (ComputeDummyStates.cxx:206-358): skip boundary cells
    interior = np.ones(fm.nCells, dtype=bool)
    interior[fm.faceL[fm.nInner:]] = False
    assert np.abs(acc[interior]).max() < 1e-14
    assert abs(fm.volumes.sum() - 1.0) < 1e-12


def test_bad_input_returns_null(product_lib):
    em = meshgen.quad_channel(3, 3)
    bad = em.trs_face_cell.copy()
    bad[0] = 4       # a cell that does not own this boundary face
    h = product_lib.mesh_build(em.dim, em.n_nodes, cl._pd(em.nodes), em.n_cells, cl._pi(em.cell_node_ptr),
                               cl._pi(em.cell_nodes), cl._pi(em.cell_shape), em.n_trs, cl._pi(em.trs_face_ptr),
                               cl._pi(bad), cl._pi(em.trs_face_node_ptr), cl._pi(em.trs_face_nodes))
    assert not h
    msg = product_lib.mesh_build_error().decode()
    assert msg and "face" in msg.lower()                 # the reason is reported, not only the NULL
    ok = cl.build_mesh(em, product_lib)
    assert ok.nFaces > 0 and product_lib.mesh_build_error() == b""


def test_diamond_volumes_of_a_caller_built_mesh(product_lib):
    """cfb200_mesh_diamond_volumes (what the COOLFluiD plugin calls after flattening the sockets) reproduces the
    cvol / cvSign of the mesh builder on hexahedra (pyramids) and tetrahedra, and gives zeros in 2D."""
    import ctypes as C
    for em in (meshgen.hex_box(5, 4, 3), meshgen.tet_box(3, 3, 2), meshgen.quad_channel(6, 4)):
        fm = cl.build_mesh(em, product_lib)
        ms = cl.mesh_struct(fm)
        cvol = np.full(fm.nFaces, -1.0)
        sign = np.full(fm.nFaces, 7, dtype=np.uint8)
        product_lib.mesh_diamond_volumes(C.byref(ms), cl._pd(cvol), cl._pu8(sign))
        assert np.array_equal(cvol, fm.cvol) and np.array_equal(sign, fm.cvSign)
        if em.dim == 3:
            assert (cvol > 0).all()
