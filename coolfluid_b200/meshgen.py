"""Synthetic unstructured meshes and fixture readers (host side, numpy only).

The generators return the *element description* a CFmesh file would hold (nodes, element-node
table, boundary TRS faces); face numbering, geometry, stencils etc. are then built by the C++
host layer (``cfb200_mesh_*``) exactly as the reference's FVMCC_MeshDataBuilder does
(/root/reference/src/Framework/FVMCC_MeshDataBuilder.cxx:91-499).

Shapes: 0 = Triag, 1 = Quad, 2 = Tetra, 3 = Hexa (P1).
"""
from __future__ import annotations
import numpy as np

SH_TRIAG, SH_QUAD, SH_TETRA, SH_HEXA = 0, 1, 2, 3
NB_NODES = {SH_TRIAG: 3, SH_QUAD: 4, SH_TETRA: 4, SH_HEXA: 8}


class ElementMesh:
    """Element-level mesh: what the reference reads from a CFmesh file."""

    def __init__(self, dim, nodes, cell_node_ptr, cell_nodes, cell_shape, trs_names, trs_face_ptr,
                 trs_face_cell, trs_face_node_ptr, trs_face_nodes):
        self.dim = int(dim)
        self.nodes = np.ascontiguousarray(nodes, dtype=np.float64)
        self.cell_node_ptr = np.ascontiguousarray(cell_node_ptr, dtype=np.int32)
        self.cell_nodes = np.ascontiguousarray(cell_nodes, dtype=np.int32)
        self.cell_shape = np.ascontiguousarray(cell_shape, dtype=np.int32)
        self.trs_names = list(trs_names)
        self.trs_face_ptr = np.ascontiguousarray(trs_face_ptr, dtype=np.int32)
        self.trs_face_cell = np.ascontiguousarray(trs_face_cell, dtype=np.int32)
        self.trs_face_node_ptr = np.ascontiguousarray(trs_face_node_ptr, dtype=np.int32)
        self.trs_face_nodes = np.ascontiguousarray(trs_face_nodes, dtype=np.int32)

    @property
    def n_cells(self):
        return len(self.cell_shape)

    @property
    def n_nodes(self):
        return self.nodes.shape[0]

    @property
    def n_trs(self):
        return len(self.trs_names)


def _pack_trs(faces_per_trs):
    """faces_per_trs: list of (cells[int], nodes[nf, k]) -> CSR arrays."""
    ptr = [0]
    cells, nptr, nodes = [], [0], []
    for c, n in faces_per_trs:
        c = np.asarray(c, dtype=np.int64)
        if len(c) == 0:                 # an empty TRS keeps its index (boundary conditions are matched by TRS index)
            cells.append(c); nodes.append(np.zeros(0, dtype=np.int64)); ptr.append(ptr[-1])
            continue
        n = np.asarray(n, dtype=np.int64).reshape(len(c), -1)
        cells.append(c)
        k = n.shape[1]
        nodes.append(n.reshape(-1))
        base = nptr[-1]
        nptr.extend((base + k*(np.arange(len(c)) + 1)).tolist())
        ptr.append(ptr[-1] + len(c))
    return (np.array(ptr, dtype=np.int32), np.concatenate(cells).astype(np.int32),
            np.array(nptr, dtype=np.int32), np.concatenate(nodes).astype(np.int32))


def _hash_uniform(ids, seed, comp):
    """U(-1, 1) as a pure function of (global node id, component, seed): splitmix64.  Any z-slab of a box can be generated
    on its own with the node positions of the whole box (bench.py: one slab per GPU / per host thread of ONE mesh)."""
    with np.errstate(over="ignore"):
        x = ids.astype(np.uint64)*np.uint64(3) + np.uint64(comp) + np.uint64(seed)*np.uint64(0x9E3779B97F4A7C15)
        x = x + np.uint64(0x9E3779B97F4A7C15)
        x = (x ^ (x >> np.uint64(30)))*np.uint64(0xBF58476D1CE4E5B9)
        x = (x ^ (x >> np.uint64(27)))*np.uint64(0x94D049BB133111EB)
        x = x ^ (x >> np.uint64(31))
    return (x >> np.uint64(11)).astype(np.float64)*(2.0/9007199254740992.0) - 1.0


def hex_box(nx, ny, nz, lengths=(1.0, 1.0, 1.0), jitter=0.15, seed=12345, k0=0, nz_total=None):
    """Unit-box hexahedral mesh, cells numbered x-fastest, interior nodes jittered by
    ``jitter*h*U(-1,1)`` so that faces are non-planar and the LSQ problem is non-trivial
    (SURVEY.md §8d config C3).  TRS order: xmin, xmax, ymin, ymax, zmin, zmax.

    ``nz_total`` given: the mesh is the slab of cell layers [k0, k0 + nz) of the box nx x ny x nz_total of extent
    ``lengths``; the jitter is then a hash of the GLOBAL node index, so that slabs generated independently are pieces of
    one mesh (node for node).  Its zmin / zmax TRSs are the cut planes when they are not the box sides."""
    lx, ly, lz = lengths
    nzt = nz if nz_total is None else int(nz_total)
    xs = np.linspace(0.0, lx, nx + 1)
    ys = np.linspace(0.0, ly, ny + 1)
    zs = np.linspace(0.0, lz, nzt + 1)[k0:k0 + nz + 1]
    Z, Y, X = np.meshgrid(zs, ys, xs, indexing="ij")
    nodes = np.stack([X, Y, Z], axis=-1).astype(np.float64)
    if jitter > 0.0:
        h = np.array([lx/nx, ly/ny, lz/nzt])
        if nz_total is None:
            rng = np.random.RandomState(seed)
            u = 2.0*rng.random_sample(nodes.shape) - 1.0
            mask = np.zeros(nodes.shape[:3], dtype=bool)
            mask[1:-1, 1:-1, 1:-1] = True
        else:
            KK, JJ, II = np.meshgrid(np.arange(k0, k0 + nz + 1), np.arange(ny + 1), np.arange(nx + 1), indexing="ij")
            gid = (KK.astype(np.int64)*(ny + 1) + JJ)*(nx + 1) + II
            u = np.stack([_hash_uniform(gid, seed, c) for c in range(3)], axis=-1)
            mask = (II > 0) & (II < nx) & (JJ > 0) & (JJ < ny) & (KK > 0) & (KK < nzt)
        nodes = nodes + jitter*h*u*mask[..., None]
    nodes = nodes.reshape(-1, 3)

    def nid(i, j, k):
        return (k*(ny + 1) + j)*(nx + 1) + i

    K, J, I = np.meshgrid(np.arange(nz), np.arange(ny), np.arange(nx), indexing="ij")
    I = I.reshape(-1); J = J.reshape(-1); K = K.reshape(-1)
    cells = np.stack([nid(I, J, K), nid(I+1, J, K), nid(I+1, J+1, K), nid(I, J+1, K),
                      nid(I, J, K+1), nid(I+1, J, K+1), nid(I+1, J+1, K+1), nid(I, J+1, K+1)], axis=1)
    cid = (K*ny + J)*nx + I

    def side(mask, ln):
        return cid[mask], cells[mask][:, ln]

    trs = [side(I == 0, [0, 4, 7, 3]), side(I == nx-1, [1, 2, 6, 5]),
           side(J == 0, [0, 1, 5, 4]), side(J == ny-1, [2, 3, 7, 6]),
           side(K == 0, [0, 3, 2, 1]), side(K == nz-1, [4, 5, 6, 7])]
    ptr, fc, fnp, fn = _pack_trs(trs)
    nc = cells.shape[0]
    return ElementMesh(3, nodes, np.arange(nc + 1, dtype=np.int32)*8, cells.reshape(-1),
                       np.full(nc, SH_HEXA), ["xmin", "xmax", "ymin", "ymax", "zmin", "zmax"], ptr, fc, fnp, fn)


def shuffle_cells(em, seed=4321):
    """The same mesh with its cells renumbered at random (nodes keep their IDs): what a mesh generator without any
    locality in its numbering hands over.  Returns (mesh, new -> old cell map)."""
    rng = np.random.RandomState(seed)
    perm = rng.permutation(em.n_cells)                         # new -> old
    inv = np.empty_like(perm); inv[perm] = np.arange(em.n_cells)
    npc = np.diff(em.cell_node_ptr).astype(np.int64)
    lnpc = npc[perm]
    ptr = np.concatenate([[0], np.cumsum(lnpc)])
    src = np.repeat(em.cell_node_ptr[perm].astype(np.int64), lnpc) + (np.arange(ptr[-1]) - np.repeat(ptr[:-1], lnpc))
    out = ElementMesh(em.dim, em.nodes, ptr.astype(np.int32), em.cell_nodes[src], em.cell_shape[perm], em.trs_names,
                      em.trs_face_ptr, inv[em.trs_face_cell].astype(np.int32), em.trs_face_node_ptr, em.trs_face_nodes)
    if hasattr(em, "stencil_kind"):
        out.stencil_kind = em.stencil_kind
    return out, perm


# Kuhn split of the reference hexahedron (local hex nodes) into 6 positively oriented tetrahedra
_KUHN = np.array([[0, 1, 2, 6], [0, 2, 3, 6], [0, 3, 7, 6], [0, 7, 4, 6], [0, 4, 5, 6], [0, 5, 1, 6]])


def tet_box(nx, ny, nz, lengths=(1.0, 1.0, 1.0), jitter=0.15, seed=12345):
    """Each hexahedron of :func:`hex_box` split into 6 tetrahedra (Kuhn)."""
    hm = hex_box(nx, ny, nz, lengths, jitter, seed)
    hexes = hm.cell_nodes.reshape(-1, 8)
    tets = hexes[:, _KUHN].reshape(-1, 4)            # hex h -> tets 6h..6h+5
    # boundary faces: every tet face that lies on a box side
    nodes = hm.nodes
    lx, ly, lz = lengths
    tol = 1e-12
    sides = [(0, 0.0), (0, lx), (1, 0.0), (1, ly), (2, 0.0), (2, lz)]
    local_faces = np.array([[0, 1, 2], [0, 3, 1], [1, 3, 2], [0, 2, 3]])
    trs = []
    for ax, val in sides:
        on = np.abs(nodes[:, ax] - val) < tol
        cs, ns = [], []
        for lf in local_faces:
            fnodes = tets[:, lf]
            m = on[fnodes].all(axis=1)
            cs.append(np.nonzero(m)[0]); ns.append(fnodes[m])
        c = np.concatenate(cs); n = np.concatenate(ns)
        order = np.argsort(c, kind="stable")
        trs.append((c[order], n[order]))
    ptr, fc, fnp, fn = _pack_trs(trs)
    nc = tets.shape[0]
    return ElementMesh(3, nodes, np.arange(nc + 1, dtype=np.int32)*4, tets.reshape(-1),
                       np.full(nc, SH_TETRA), hm.trs_names, ptr, fc, fnp, fn)


def quad_channel(nx, ny, lengths=(1.0, 1.0), jitter=0.15, seed=12345, triangles=False):
    """2D structured-as-unstructured channel of quads (or triangles: each quad split along 0-2).
    TRS order: xmin, xmax, ymin, ymax."""
    lx, ly = lengths
    xs = np.linspace(0.0, lx, nx + 1)
    ys = np.linspace(0.0, ly, ny + 1)
    Y, X = np.meshgrid(ys, xs, indexing="ij")
    nodes = np.stack([X, Y], axis=-1)
    if jitter > 0.0:
        rng = np.random.RandomState(seed)
        h = np.array([lx/nx, ly/ny])
        d = jitter*h*(2.0*rng.random_sample(nodes.shape) - 1.0)
        mask = np.zeros(nodes.shape[:2], dtype=bool)
        mask[1:-1, 1:-1] = True
        nodes = nodes + d*mask[..., None]
    nodes = nodes.reshape(-1, 2)

    def nid(i, j):
        return j*(nx + 1) + i

    J, I = np.meshgrid(np.arange(ny), np.arange(nx), indexing="ij")
    I = I.reshape(-1); J = J.reshape(-1)
    quads = np.stack([nid(I, J), nid(I+1, J), nid(I+1, J+1), nid(I, J+1)], axis=1)
    qid = J*nx + I
    if not triangles:
        def side(mask, ln):
            return qid[mask], quads[mask][:, ln]
        trs = [side(I == 0, [3, 0]), side(I == nx-1, [1, 2]), side(J == 0, [0, 1]), side(J == ny-1, [2, 3])]
        ptr, fc, fnp, fn = _pack_trs(trs)
        nc = quads.shape[0]
        return ElementMesh(2, nodes, np.arange(nc + 1, dtype=np.int32)*4, quads.reshape(-1),
                           np.full(nc, SH_QUAD), ["xmin", "xmax", "ymin", "ymax"], ptr, fc, fnp, fn)
    tris = np.stack([quads[:, [0, 1, 2]], quads[:, [0, 2, 3]]], axis=1).reshape(-1, 3)
    # tri 2q = (0,1,2) holds edges 0-1 (ymin side) and 1-2 (xmax side); tri 2q+1 = (0,2,3): 2-3 (ymax), 3-0 (xmin)
    trs = [(2*qid[I == 0] + 1, quads[I == 0][:, [3, 0]]), (2*qid[I == nx-1], quads[I == nx-1][:, [1, 2]]),
           (2*qid[J == 0], quads[J == 0][:, [0, 1]]), (2*qid[J == ny-1] + 1, quads[J == ny-1][:, [2, 3]])]
    ptr, fc, fnp, fn = _pack_trs(trs)
    nc = tris.shape[0]
    return ElementMesh(2, nodes, np.arange(nc + 1, dtype=np.int32)*3, tris.reshape(-1),
                       np.full(nc, SH_TRIAG), ["xmin", "xmax", "ymin", "ymax"], ptr, fc, fnp, fn)


def read_thor(thor_path, sp_path):
    """Reader for the THOR mesh format + .SP super-patch file, producing the same element / TRS
    ordering as the reference's converter writes into the CFmesh it then reads
    (/root/reference/plugins/THOR2CFmesh/THOR2CFmeshConverter.cxx:294-397 readTHOR, :424-490 readSP,
    :712-790 writeDiscontinuousTrsData: TRS = super patches in file order, each the concatenation of
    its patches in the listed order; indices shifted to 0-based)."""
    tok = open(thor_path).read().split()
    p = 0

    def nxt():
        nonlocal p
        p += 1
        return tok[p-1]

    dim = int(nxt()); nb_lam = int(nxt()); nb_turb = int(nxt())
    nb_cells = int(nxt()); nb_nodes = int(nxt()); nb_faces = int(nxt()); nb_patches = int(nxt())
    nb_types = int(nxt())
    types = [(int(nxt()), int(nxt())) for _ in range(nb_types)]
    cell_nodes, cell_ptr, cell_shape = [], [0], []
    for npc, ncells in types:
        arr = np.array(tok[p:p + npc*ncells], dtype=np.int64) - 1
        p += npc*ncells
        cell_nodes.append(arr)
        cell_ptr.extend((cell_ptr[-1] + npc*(np.arange(ncells) + 1)).tolist())
        if dim == 2:
            sh = SH_TRIAG if npc == 3 else SH_QUAD
        else:
            sh = SH_TETRA if npc == 4 else SH_HEXA
        cell_shape.extend([sh]*ncells)
    patches = {}
    for _ in range(nb_patches):
        code = int(nxt()); nf = int(nxt())
        cells, nodes = [], []
        for _f in range(nf):
            cid = int(nxt()) - 1; k = int(nxt())
            nodes.append([int(nxt()) - 1 for _k in range(k)])
            cells.append(cid)
        patches[code] = (cells, nodes)
    p += nb_nodes                                   # list of updatable nodes
    coords = np.array(tok[p:p + dim*nb_nodes], dtype=np.float64).reshape(dim, nb_nodes).T.copy()
    # super patches
    sp = open(sp_path).read().split()
    q = 0
    nsp = int(sp[q]); q += 1
    names, trs = [], []
    for _ in range(nsp):
        name = sp[q]; n = int(sp[q+1]); q += 2
        ids = [int(x) for x in sp[q:q+n]]; q += n
        cells, nodes = [], []
        for pid in ids:
            cells.extend(patches[pid][0]); nodes.extend(patches[pid][1])
        names.append(name)
        trs.append((cells, nodes))
    ptr, fc, fnp, fn = _pack_trs(trs)
    return ElementMesh(dim, coords, np.array(cell_ptr), np.concatenate(cell_nodes), np.array(cell_shape),
                       names, ptr, fc, fnp, fn)


_CFMESH_SHAPES = {"Triag": (SH_TRIAG, 3), "Quad": (SH_QUAD, 4), "Tetra": (SH_TETRA, 4), "Hexa": (SH_HEXA, 8)}


def read_cfmesh(path):
    """Reader for the reference's native ASCII mesh format (what CFmeshFileReader consumes,
    /root/reference/src/Framework/CFmeshFileReader.ci:60-100 keyword table, :902-951 readGeomEntList), restricted to what the
    cell-centred FV path reads: P1 geometry, P0 solution (one state per element), Triag / Quad / Tetra / Hexa elements.
    Elements keep the file order (type after type), the TRSs the file order, each the concatenation of its TRs; a boundary
    face is `nbNodes nbStates nodes... states...` and its state ID names the cell it belongs to.
    Returns (ElementMesh, states or None): states [nCells, nEq] when the file carries a solution (!LIST_STATE 1)."""
    tok = open(path).read().split()
    p = 0
    n = len(tok)
    dim = neq = nb_nodes = nb_states = 0
    coords = states = None
    types, per_type, nodes_per_type, states_per_type = [], [], [], []
    elem_nodes, elem_ptr, elem_shape, state_of_elem = [], [0], [], []
    names, trs = [], []
    geo_order = sol_order = None
    while p < n:
        key = tok[p]; p += 1
        if key == "!END":
            break
        if key == "!NB_DIM":
            dim = int(tok[p]); p += 1
        elif key == "!NB_EQ":
            neq = int(tok[p]); p += 1
        elif key == "!NB_NODES":
            nb_nodes = int(tok[p]); p += 2                      # updatable, non-updatable
        elif key == "!LIST_NODE":
            coords = np.array(tok[p:p + dim*nb_nodes], dtype=np.float64).reshape(nb_nodes, dim); p += dim*nb_nodes
        elif key == "!NB_STATES":
            nb_states = int(tok[p]); p += 2
        elif key == "!LIST_STATE":
            flag = int(tok[p]); p += 1
            if flag:
                states = np.array(tok[p:p + neq*nb_states], dtype=np.float64).reshape(nb_states, neq); p += neq*nb_states
        elif key == "!NB_ELEM":
            p += 1
        elif key == "!NB_ELEM_TYPES":
            ntypes = int(tok[p]); p += 1
        elif key == "!GEOM_POLYORDER":
            geo_order = int(tok[p]); p += 1
        elif key == "!SOL_POLYORDER":
            sol_order = int(tok[p]); p += 1
        elif key == "!ELEM_TYPES":
            types = tok[p:p + ntypes]; p += ntypes
        elif key == "!NB_ELEM_PER_TYPE":
            per_type = [int(x) for x in tok[p:p + ntypes]]; p += ntypes
        elif key == "!NB_NODES_PER_TYPE":
            nodes_per_type = [int(x) for x in tok[p:p + ntypes]]; p += ntypes
        elif key == "!NB_STATES_PER_TYPE":
            states_per_type = [int(x) for x in tok[p:p + ntypes]]; p += ntypes
        elif key == "!LIST_ELEM":
            if geo_order != 1 or sol_order != 0:
                raise ValueError("read_cfmesh: only P1 geometry with a P0 (cell-centred) solution is supported")
            for t, ne, nn, ns in zip(types, per_type, nodes_per_type, states_per_type):
                if t not in _CFMESH_SHAPES or _CFMESH_SHAPES[t][1] != nn or ns != 1:
                    raise ValueError("read_cfmesh: unsupported element type %s" % t)
                a = np.array(tok[p:p + ne*(nn + 1)], dtype=np.int64).reshape(ne, nn + 1); p += ne*(nn + 1)
                elem_nodes.append(a[:, :nn].reshape(-1))
                state_of_elem.append(a[:, nn])
                elem_ptr.extend((elem_ptr[-1] + nn*(np.arange(ne) + 1)).tolist())
                elem_shape.extend([_CFMESH_SHAPES[t][0]]*ne)
        elif key == "!NB_TRSs":
            p += 1
        elif key == "!TRS_NAME":
            names.append(tok[p]); p += 1
        elif key == "!NB_TRs":
            ntr = int(tok[p]); p += 1
        elif key == "!NB_GEOM_ENTS":
            counts = [int(x) for x in tok[p:p + ntr]]; p += ntr
        elif key == "!GEOM_TYPE":
            p += 1
        elif key == "!LIST_GEOM_ENT":
            cells, fnodes = [], []
            for _ in range(sum(counts)):
                nn, ns = int(tok[p]), int(tok[p + 1]); p += 2
                fnodes.append([int(x) for x in tok[p:p + nn]]); p += nn
                cells.append(int(tok[p])); p += ns
            trs.append((cells, fnodes))
        elif key.startswith("!"):
            # keywords without meaning for this path (e.g. !COOLFLUID_VERSION, !CFMESH_FORMAT_VERSION, !NB_EXTRA_*): skip their values
            while p < n and not tok[p].startswith("!"):
                p += 1
    soe = np.concatenate(state_of_elem)
    cell_of_state = np.empty(len(soe), dtype=np.int64); cell_of_state[soe] = np.arange(len(soe))
    # one TRS may mix face shapes only in 3D hybrid meshes, which the element types above already exclude per mesh
    packed = []
    for cells, fnodes in trs:
        k = len(fnodes[0]) if fnodes else 2
        packed.append((cell_of_state[np.array(cells, dtype=np.int64)] if cells else [], np.array(fnodes, dtype=np.int64).reshape(len(cells), k)))
    ptr, fc, fnp, fn = _pack_trs(packed)
    em = ElementMesh(dim, coords, np.array(elem_ptr), np.concatenate(elem_nodes), np.array(elem_shape), names, ptr, fc, fnp, fn)
    if states is not None:
        states = states[soe]
    return em, states
