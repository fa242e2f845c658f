"""FlatMesh: the flat-array form of the FVMCC mesh data (what the reference keeps in the sockets
``nodes, states(coords), normals, faceAreas, volumes, stencil`` and the connectivity tables
``cellFaces``, ``InnerFacesStates`` ... — SURVEY.md §8b "Data ownership").

Faces are stored in the reference's LOOP order (InnerFaces TRS first, then each boundary TRS);
``faceR`` of a boundary face is ``nCells + ghostIndex``; state-indexed arrays (centres, U) hold the
cells first, then the ghost states.
"""
from __future__ import annotations
import numpy as np

INT_FIELDS = ["faceL", "faceR", "faceNodePtr", "faceNodes", "faceGeoID", "faceBC", "cellFacePtr", "cellFaces",
              "stencilPtr", "stencil", "nodeNbrPtr", "nodeNbr", "trsPtr", "bsrRowPtr", "bsrColInd"]
F64_FIELDS = ["centres", "normals", "areas", "volumes", "nodeW", "cvol", "faceCentres"]
U8_FIELDS = ["cvSign", "updatable"]


class FlatMesh:
    def __init__(self, dim, nodes, cell_node_ptr, cell_nodes, cell_shape, trs_names):
        self.dim = int(dim)
        self.nodes = np.ascontiguousarray(nodes, dtype=np.float64)
        self.cellNodePtr = np.ascontiguousarray(cell_node_ptr, dtype=np.int32)
        self.cellNodes = np.ascontiguousarray(cell_nodes, dtype=np.int32)
        self.cellShape = np.ascontiguousarray(cell_shape, dtype=np.int32)
        self.trs_names = list(trs_names)
        self.nCells = len(self.cellShape)
        self.nNodes = self.nodes.shape[0]
        self.nFaces = self.nInner = self.nGhosts = 0

    @property
    def nStates(self):
        return self.nCells + self.nGhosts

    def sizes_for(self, name):
        d, nc, ng, nf, nn = self.dim, self.nCells, self.nGhosts, self.nFaces, self.nNodes
        return {"centres": (nc + ng)*d, "normals": nf*d, "areas": nf, "volumes": nc, "cvol": nf,
                "faceCentres": nf*d, "faceL": nf, "faceR": nf, "faceNodePtr": nf + 1, "faceGeoID": nf,
                "faceBC": nf, "cellFacePtr": nc + 1, "stencilPtr": nc + 1, "nodeNbrPtr": nn + 1,
                "trsPtr": len(self.trs_names) + 1, "cvSign": nf, "updatable": nc, "bsrRowPtr": nc + 1}[name]

    def equal_to(self, other, exact=True):
        """Field-by-field comparison (used by the parity tests of the host mesh builder)."""
        bad = []
        for f in INT_FIELDS + U8_FIELDS + F64_FIELDS:
            a, b = getattr(self, f), getattr(other, f)
            if a.shape != b.shape or not np.array_equal(a, b):
                bad.append(f)
        return bad
