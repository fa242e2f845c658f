"""Domain decomposition for the stand-alone harness (host side, numpy only): one partition per GPU.

Mirrors what a COOLFluiD rank holds after the parallel mesh reader has run:
  * its *updatable* cells (element -> rank map from ParMETIS, /root/reference/src/Framework/ParMetis.cxx:146-160;
    here: an imported ``part[]`` or recursive coordinate bisection),
  * ``overlap`` layers of non-updatable cells grown by node adjacency (2 for the least-squares reconstruction,
    /root/reference/plugins/FiniteVolume/LeastSquareP1Setup.cxx:60-63),
  * the physical boundary TRSs restricted to local cells, plus a ``PartitionFaces`` TRS for the faces whose other
    cell is not local (they are never evaluated: their inner cell is not updatable),
  * send / receive lists for the state synchronisation after every update
    (DataHandle<State*,GLOBAL>::synchronize, /root/reference/src/Common/MPI/MPICommPattern.hh:1735-1782).

Local cell order: [updatable cells whose node neighbourhood is entirely updatable | other updatable cells | overlap
cells], each group in ascending global ID, so that work on the first group can start before the halo has landed.
"""
from __future__ import annotations
import numpy as np

from .meshgen import ElementMesh, SH_TRIAG, SH_QUAD, SH_TETRA, SH_HEXA, _pack_trs

# local face -> local nodes (node sets only matter: the host mesh builder matches faces by their sorted node tuple)
_FACES = {
    SH_TRIAG: [[0, 1], [1, 2], [2, 0]],
    SH_QUAD: [[0, 1], [1, 2], [2, 3], [3, 0]],
    SH_TETRA: [[0, 2, 1], [0, 1, 3], [1, 2, 3], [0, 3, 2]],
    SH_HEXA: [[0, 3, 2, 1], [4, 5, 6, 7], [0, 1, 5, 4], [1, 2, 6, 5], [3, 7, 6, 2], [0, 4, 7, 3]],
}


def cell_centroids(em: ElementMesh) -> np.ndarray:
    npc = np.diff(em.cell_node_ptr)
    sums = np.zeros((em.n_cells, em.dim))
    owner = np.repeat(np.arange(em.n_cells), npc)
    np.add.at(sums, owner, em.nodes[em.cell_nodes])
    return sums/npc[:, None]


def rcb_partition(em: ElementMesh, nparts: int) -> np.ndarray:
    """Recursive coordinate bisection of the cell centroids (stand-in for ParMETIS in the harness)."""
    xc = cell_centroids(em)
    part = np.zeros(em.n_cells, dtype=np.int32)

    def split(ids, p0, n):
        if n == 1:
            part[ids] = p0
            return
        nl = n//2
        ext = xc[ids].max(axis=0) - xc[ids].min(axis=0)
        ax = int(np.argmax(ext))
        order = ids[np.argsort(xc[ids, ax], kind="stable")]
        cut = (len(ids)*nl)//n
        split(order[:cut], p0, nl)
        split(order[cut:], p0 + nl, n - nl)

    split(np.arange(em.n_cells), 0, int(nparts))
    return part


def _node_to_cells(em: ElementMesh):
    npc = np.diff(em.cell_node_ptr)
    owner = np.repeat(np.arange(em.n_cells, dtype=np.int64), npc)
    order = np.argsort(em.cell_nodes, kind="stable")
    ptr = np.zeros(em.n_nodes + 1, dtype=np.int64)
    np.add.at(ptr, em.cell_nodes.astype(np.int64) + 1, 1)
    return np.cumsum(ptr), owner[order]


def _grow(em, n2c_ptr, n2c, mask):
    """cells sharing at least one node with a cell of ``mask`` (including those cells)."""
    npc = np.diff(em.cell_node_ptr)
    sel = np.repeat(mask, npc)
    nodes = np.unique(em.cell_nodes[sel])
    out = mask.copy()
    starts, ends = n2c_ptr[nodes], n2c_ptr[nodes + 1]
    idx = np.concatenate([np.arange(s, e) for s, e in zip(starts, ends)]) if len(nodes) else np.zeros(0, dtype=np.int64)
    out[n2c[idx]] = True
    return out


class Partition:
    """One rank's view: local element mesh + maps + halo lists (all cell ids local unless named global)."""

    def __init__(self):
        self.rank = 0
        self.nranks = 1
        self.em = None              # local ElementMesh (TRS: the global ones restricted + "PartitionFaces")
        self.cell_l2g = None        # [nLocal] global cell id
        self.updatable = None       # [nLocal] uint8
        self.n_interior = 0         # cells whose whole node neighbourhood is updatable (leading range when order = "grouped")
        self.indep = None           # the same cells as a list of local ids (order = "global")
        self.n_updatable = 0
        self.send = {}              # peer -> local ids (this rank's updatable cells that are overlap cells on peer)
        self.recv = {}              # peer -> local ids of overlap cells owned by peer

    def halo_csr(self):
        """(sendPtr, sendCells, recvPtr, recvCells): CSR over ranks, as cfb200_set_halo takes them."""
        sp, sc, rp, rc = [0], [], [0], []
        for r in range(self.nranks):
            s = self.send.get(r, np.zeros(0, dtype=np.int32)); q = self.recv.get(r, np.zeros(0, dtype=np.int32))
            sc.append(s); rc.append(q)
            sp.append(sp[-1] + len(s)); rp.append(rp[-1] + len(q))
        cat = lambda xs: np.ascontiguousarray(np.concatenate(xs) if xs else np.zeros(0), dtype=np.int32)
        return (np.array(sp, dtype=np.int32), cat(sc), np.array(rp, dtype=np.int32), cat(rc))


def _face_keys(cell_nodes_2d, shape):
    """[n, nf, 4] sorted (descending, -1 padded) node keys of the faces of n cells of one shape."""
    lf = np.array(_FACES[shape])
    f = cell_nodes_2d[:, lf]                                   # [n, nf, k]
    if f.shape[2] < 4:
        f = np.concatenate([f, np.full(f.shape[:2] + (4 - f.shape[2],), -1, dtype=f.dtype)], axis=2)
    return -np.sort(-f, axis=2), lf


def extract(em: ElementMesh, part: np.ndarray, rank: int, nranks: int | None = None, overlap: int = 2,
            order: str = "global") -> Partition:
    """Local mesh, maps and halo lists of ``rank`` for the element -> rank map ``part``.

    order = "global" (default): local cells in ascending global ID.  Face creation order, face orientation (L = the
    lower-ID cell), least-squares edge order and stencil order of every updatable cell are then those of the
    single-domain mesh restricted to the partition, so the partitioned evaluation is BIT-IDENTICAL to the single-domain
    one; the cells whose gradients may run while the halo is in flight are given as the list ``indep``.
    order = "grouped": [interior | next to the overlap | overlap], the interior ones as a leading range
    (``n_interior``); same values up to the summation order."""
    part = np.asarray(part)
    nranks = int(nranks if nranks is not None else part.max() + 1)
    n2c_ptr, n2c = _node_to_cells(em)
    own = part == rank
    local = own.copy()
    for _ in range(overlap):
        local = _grow(em, n2c_ptr, n2c, local)
    near_other = _grow(em, n2c_ptr, n2c, ~own) if not own.all() else np.zeros_like(own)
    g_int = np.nonzero(own & ~near_other)[0]
    g_bnd = np.nonzero(own & near_other)[0]
    g_ovl = np.nonzero(local & ~own)[0]
    if order == "global":
        l2g = np.sort(np.concatenate([g_int, g_bnd, g_ovl])).astype(np.int64)
    else:
        l2g = np.concatenate([g_int, g_bnd, g_ovl]).astype(np.int64)
    nl = len(l2g)
    g2l = np.full(em.n_cells, -1, dtype=np.int64)
    g2l[l2g] = np.arange(nl)

    # local element-node table (node ids compacted in ascending global order)
    npc = np.diff(em.cell_node_ptr).astype(np.int64)
    lnpc = npc[l2g]
    lptr = np.concatenate([[0], np.cumsum(lnpc)])
    src = np.repeat(em.cell_node_ptr[l2g].astype(np.int64), lnpc) + (np.arange(lptr[-1]) - np.repeat(lptr[:-1], lnpc))
    gnodes = em.cell_nodes[src].astype(np.int64)
    used = np.unique(gnodes)
    n_g2l = np.full(em.n_nodes, -1, dtype=np.int64)
    n_g2l[used] = np.arange(len(used))
    lnodes = n_g2l[gnodes]
    shape = em.cell_shape[l2g]

    # physical boundary TRSs restricted to local cells
    trs, trs_key_rows = [], []
    for t in range(em.n_trs):
        b, e = em.trs_face_ptr[t], em.trs_face_ptr[t+1]
        cells_g = em.trs_face_cell[b:e].astype(np.int64)
        keep = np.nonzero(g2l[cells_g] >= 0)[0] + b
        fcells, fnodes = [], []
        for gf in keep:
            nn = em.trs_face_nodes[em.trs_face_node_ptr[gf]:em.trs_face_node_ptr[gf+1]].astype(np.int64)
            fcells.append(g2l[em.trs_face_cell[gf]])
            fnodes.append(n_g2l[nn])
        k = len(fnodes[0]) if fnodes else (em.dim if em.dim == 2 else 3)
        same = all(len(x) == k for x in fnodes)
        if not same:
            raise NotImplementedError("mixed face shapes inside one TRS")
        fn = np.array(fnodes, dtype=np.int64).reshape(len(fcells), k)
        trs.append((np.array(fcells, dtype=np.int64), fn))
        if len(fcells):
            pad = np.concatenate([fn, np.full((len(fn), 4 - k), -1, dtype=np.int64)], axis=1)
            trs_key_rows.append(-np.sort(-pad, axis=1))

    # partition faces: cell faces that appear once among the local cells and in no physical TRS
    keys, owner_cell, raw_nodes = [], [], []
    for sh in np.unique(shape):
        ids = np.nonzero(shape == sh)[0]
        k = int(lnpc[ids[0]])
        cn = lnodes[(lptr[ids][:, None] + np.arange(k)[None, :])]
        fk, lf = _face_keys(cn, int(sh))
        nf = fk.shape[1]
        keys.append(fk.reshape(-1, 4))
        owner_cell.append(np.repeat(ids, nf))
        rn = cn[:, lf]                                           # [n, nf, kf] unsorted local nodes
        kf = rn.shape[2]
        if kf < 4:
            rn = np.concatenate([rn, np.full(rn.shape[:2] + (4 - kf,), -1, dtype=rn.dtype)], axis=2)
        raw_nodes.append(rn.reshape(-1, 4))
    keys = np.concatenate(keys); owner_cell = np.concatenate(owner_cell); raw_nodes = np.concatenate(raw_nodes)
    allk = np.concatenate([keys] + trs_key_rows) if trs_key_rows else keys
    _, inv, counts = np.unique(allk, axis=0, return_inverse=True, return_counts=True)
    inv = inv.reshape(-1)
    single = np.nonzero(counts[inv[:len(keys)]] == 1)[0]
    single = single[np.argsort(owner_cell[single], kind="stable")]
    pf_cells = owner_cell[single]
    pf_nodes = raw_nodes[single]
    names = list(em.trs_names)
    if len(single):
        kf = int((pf_nodes[0] >= 0).sum())
        if not ((pf_nodes >= 0).sum(axis=1) == kf).all():
            raise NotImplementedError("mixed face shapes among the partition faces")
        trs.append((pf_cells, pf_nodes[:, :kf]))
        names.append("PartitionFaces")
    ptr, fc, fnp, fn = _pack_trs([(c, n) if len(c) else (np.zeros(0, dtype=np.int64), np.zeros((0, 1), dtype=np.int64))
                                  for c, n in trs])
    lem = ElementMesh(em.dim, em.nodes[used], lptr.astype(np.int32), lnodes.astype(np.int32), shape, names, ptr, fc, fnp, fn)
    if hasattr(em, "stencil_kind"):          # Setup1.stencil of the global mesh (lib.build_mesh reads it)
        lem.stencil_kind = em.stencil_kind

    p = Partition()
    p.rank, p.nranks, p.em, p.cell_l2g = int(rank), nranks, lem, l2g
    p.node_l2g = used
    p.n_interior, p.n_updatable = len(g_int), len(g_int) + len(g_bnd)
    p.updatable = np.ascontiguousarray(own[l2g].astype(np.uint8))
    if order == "global":
        p.indep = g2l[g_int].astype(np.int32)          # ascending
    # halo lists (both sides sort by global id, so the packed orders agree without any handshake)
    for peer in np.unique(part[g_ovl]):
        peer = int(peer)
        p.recv[peer] = g2l[g_ovl[part[g_ovl] == peer]].astype(np.int32)
        theirs = part == peer
        for _ in range(overlap):
            theirs = _grow(em, n2c_ptr, n2c, theirs)
        p.send[peer] = g2l[np.nonzero(theirs & own)[0]].astype(np.int32)
    return p
