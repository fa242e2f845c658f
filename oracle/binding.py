"""ORACLE — TEST INFRASTRUCTURE ONLY.  ctypes binding of oracle/liboracle*.so (the CPU restatement
of the reference's FVMCC path) and of oracle/_ref/libcfref.so (the reference's own MathTools headers).
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this."""
from __future__ import annotations
import ctypes as C
import os
import subprocess
import numpy as np

from coolfluid_b200.config import ConfigStruct, BCStruct, bc_array
from coolfluid_b200.flatmesh import FlatMesh, INT_FIELDS, F64_FIELDS, U8_FIELDS

HERE = os.path.dirname(os.path.abspath(__file__))
P_D = C.POINTER(C.c_double)
P_I = C.POINTER(C.c_int32)
P_U8 = C.POINTER(C.c_uint8)


class MeshView(C.Structure):
    _fields_ = [("dim", C.c_int32), ("nCells", C.c_int32), ("nGhosts", C.c_int32), ("nNodes", C.c_int32),
                ("nFaces", C.c_int32), ("nInner", C.c_int32),
                ("nodes", P_D), ("centres", P_D), ("faceL", P_I), ("faceR", P_I), ("faceNodePtr", P_I),
                ("faceNodes", P_I), ("normals", P_D), ("areas", P_D), ("volumes", P_D),
                ("cellFacePtr", P_I), ("cellFaces", P_I), ("stencilPtr", P_I), ("stencil", P_I),
                ("nodeNbrPtr", P_I), ("nodeNbr", P_I), ("nodeW", P_D), ("cvol", P_D), ("cvSign", P_U8),
                ("faceBC", P_I), ("updatable", P_U8), ("cellNodePtr", P_I), ("cellNodes", P_I),
                ("bcFaceVal", P_D), ("bcFaceFlag", P_U8),
                ("nodePresc", P_U8), ("nodePrescIdx", P_I), ("nodePrescVal", P_D), ("nNodePresc", C.c_int32)]


class Outputs(C.Structure):
    _fields_ = [("rhs", P_D), ("updateCoeff", P_D), ("uX", P_D), ("uY", P_D), ("uZ", P_D), ("limiter", P_D),
                ("nstates", P_D), ("bsrRowPtr", P_I), ("bsrColInd", P_I), ("bsrVal", P_D), ("status", C.c_int32), ("divB", P_D)]


def build(omp=False, ref=True):
    """Compile the oracle (and, when /root/reference exists, oracle/_ref) with the committed Makefile."""
    targets = ["liboracle.so", "liboracle_fast.so"] + (["ref"] if ref else [])
    subprocess.check_call(["make", "-s", "-C", HERE] + targets)


_libs = {}


def lib(omp=False):
    name = "liboracle_fast.so" if omp else "liboracle.so"
    if name not in _libs:
        path = os.path.join(HERE, name)
        if not os.path.exists(path):
            build()
        L = C.CDLL(path)
        L.cfo_mesh_build.restype = C.c_void_p
        L.cfo_mesh_build.argtypes = [C.c_int, C.c_int, P_D, C.c_int, P_I, P_I, P_I, C.c_int, P_I, P_I, P_I, P_I]
        L.cfo_mesh_sizes.argtypes = [C.c_void_p, C.POINTER(C.c_int64)]
        L.cfo_mesh_get.argtypes = [C.c_void_p, C.c_char_p, C.c_void_p]
        L.cfo_mesh_free.argtypes = [C.c_void_p]
        L.cfo_mesh_set_stencil.restype = C.c_int
        L.cfo_mesh_set_stencil.argtypes = [C.c_void_p, C.c_int]
        L.cfo_apply_bc.argtypes = [C.POINTER(ConfigStruct), C.POINTER(MeshView), C.POINTER(BCStruct), P_D]
        L.cfo_compute.argtypes = [C.POINTER(ConfigStruct), C.POINTER(MeshView), C.POINTER(BCStruct), P_D, C.POINTER(Outputs)]
        L.cfo_time_rhs.argtypes = [C.POINTER(ConfigStruct), C.POINTER(MeshView), C.POINTER(BCStruct), P_D, P_D, C.POINTER(Outputs)]
        L.cfo_run_fwd_euler.argtypes = [C.POINTER(ConfigStruct), C.POINTER(MeshView), C.POINTER(BCStruct), P_D,
                                        C.POINTER(Outputs), C.c_double, C.c_int, C.c_double, C.c_int, P_D]
        L.cfo_roe_dissipation.argtypes = [C.c_int, P_D, P_D, P_D, C.c_double, P_D, P_D, P_D, P_D]
        L.cfo_det3.restype = C.c_double; L.cfo_det3.argtypes = [P_D]
        L.cfo_det4.restype = C.c_double; L.cfo_det4.argtypes = [P_D]
        L.cfo_pow15.restype = C.c_double; L.cfo_pow15.argtypes = [C.c_double]
        L.cfo_pow3.restype = C.c_double; L.cfo_pow3.argtypes = [C.c_double]
        assert L.cfo_sizeof_config() == C.sizeof(ConfigStruct), "Config layout mismatch"
        assert L.cfo_sizeof_bc() == C.sizeof(BCStruct), "BC layout mismatch"
        assert L.cfo_sizeof_meshview() == C.sizeof(MeshView), "MeshView layout mismatch"
        _libs[name] = L
    return _libs[name]


def reflib():
    """oracle/_ref/libcfref.so: the reference's MathTools headers compiled in place (None if absent)."""
    if "ref" not in _libs:
        path = os.path.join(HERE, "_ref", "libcfref.so")
        if not os.path.exists(path):
            _libs["ref"] = None
        else:
            L = C.CDLL(path)
            L.ref_roe_dissipation.argtypes = [C.c_int, P_D, P_D, P_D, C.c_double, P_D, P_D, P_D, P_D]
            L.ref_matvec.argtypes = [C.c_int, P_D, P_D, P_D]
            for f in ("ref_det3", "ref_det4"):
                getattr(L, f).restype = C.c_double; getattr(L, f).argtypes = [P_D]
            L.ref_inner_prod.restype = C.c_double; L.ref_inner_prod.argtypes = [C.c_int, P_D, P_D]
            L.ref_distance.restype = C.c_double; L.ref_distance.argtypes = [C.c_int, P_D, P_D]
            L.ref_cross.argtypes = [P_D, P_D, P_D]
            L.ref_norm2.restype = C.c_double; L.ref_norm2.argtypes = [C.c_int, P_D]
            _libs["ref"] = L
    return _libs["ref"]


def pd(a):
    return a.ctypes.data_as(P_D)


def pi(a):
    return a.ctypes.data_as(P_I)


def pu8(a):
    return a.ctypes.data_as(P_U8)


STENCIL_KINDS = {"FaceVertexPlusGhost": 0, "FaceVertex": 1, "FaceEdge": 2, "Face": 3}     # Setup1.stencil


def build_mesh(em, stencil=None) -> FlatMesh:
    """Literal (slow) reference algorithm: FVMCC_MeshDataBuilder + geometry + stencils (oracle/cfo_mesh.hpp)."""
    stencil = stencil or getattr(em, "stencil_kind", "FaceVertexPlusGhost")   # Setup1.stencil of the case (tests/cases.with_stencil)
    L = lib()
    h = L.cfo_mesh_build(em.dim, em.n_nodes, pd(em.nodes), em.n_cells, pi(em.cell_node_ptr), pi(em.cell_nodes),
                         pi(em.cell_shape), em.n_trs, pi(em.trs_face_ptr), pi(em.trs_face_cell),
                         pi(em.trs_face_node_ptr), pi(em.trs_face_nodes))
    if not h:
        raise RuntimeError("oracle mesh build failed")
    if STENCIL_KINDS[stencil] and L.cfo_mesh_set_stencil(h, STENCIL_KINDS[stencil]):
        raise RuntimeError("oracle stencil rebuild failed")
    sz = (C.c_int64*8)()
    L.cfo_mesh_sizes(h, sz)
    fm = FlatMesh(em.dim, em.nodes, em.cell_node_ptr, em.cell_nodes, em.cell_shape, em.trs_names)
    fm.nFaces, fm.nInner, fm.nGhosts = int(sz[0]), int(sz[1]), int(sz[2])
    var = {"faceNodes": sz[3], "cellFaces": sz[4], "stencil": sz[5], "nodeNbr": sz[6], "nodeW": sz[6], "bsrColInd": sz[7]}
    for name in INT_FIELDS + F64_FIELDS + U8_FIELDS:
        n = int(var[name]) if name in var else fm.sizes_for(name)
        dt = np.int32 if name in INT_FIELDS else (np.float64 if name in F64_FIELDS else np.uint8)
        a = np.zeros(n, dtype=dt)
        if L.cfo_mesh_get(h, name.encode(), a.ctypes.data_as(C.c_void_p)):
            raise KeyError(name)
        setattr(fm, name, a)
    L.cfo_mesh_free(h)
    return fm


def mesh_view(fm: FlatMesh) -> MeshView:
    mv = MeshView()
    mv.dim, mv.nCells, mv.nGhosts, mv.nNodes, mv.nFaces, mv.nInner = fm.dim, fm.nCells, fm.nGhosts, fm.nNodes, fm.nFaces, fm.nInner
    for name, _t in MeshView._fields_[6:]:
        a = getattr(fm, name, None)
        if name == "nNodePresc":
            mv.nNodePresc = len(getattr(fm, "nodePrescIdx", ()))
            continue
        if a is None:                       # optional: bcFaceVal / bcFaceFlag / nodePresc*
            continue
        setattr(mv, name, pd(a) if a.dtype == np.float64 else (pi(a) if a.dtype == np.int32 else pu8(a)))
    return mv


class Result:
    pass


def apply_bc(cfg, fm, bcs, U, omp=False):
    mv = mesh_view(fm)
    rc = lib(omp).cfo_apply_bc(C.byref(cfg), C.byref(mv), bc_array(bcs), pd(U))
    if rc:
        raise RuntimeError("oracle apply_bc failed")


def alloc_outputs(cfg, fm, jacobian, limiter_init=None):
    q = cfg.nEqs
    r = Result()
    r.rhs = np.zeros(fm.nCells*q); r.updateCoeff = np.zeros(fm.nCells)
    r.uX = np.zeros(fm.nCells*q); r.uY = np.zeros(fm.nCells*q); r.uZ = np.zeros(fm.nCells*q)
    r.limiter = np.ones(fm.nCells*q) if limiter_init is None else np.array(limiter_init, dtype=np.float64).copy()
    r.nstates = np.zeros(fm.nNodes*q)
    r.bsrVal = np.zeros(len(fm.bsrColInd)*q*q) if jacobian else None
    o = Outputs()
    o.rhs, o.updateCoeff, o.uX, o.uY, o.uZ = pd(r.rhs), pd(r.updateCoeff), pd(r.uX), pd(r.uY), pd(r.uZ)
    o.limiter, o.nstates = pd(r.limiter), pd(r.nstates)
    o.bsrRowPtr, o.bsrColInd = pi(fm.bsrRowPtr), pi(fm.bsrColInd)
    o.bsrVal = pd(r.bsrVal) if jacobian else None
    r.divB = np.zeros(fm.nCells) if cfg.nEqs == 9 else None
    o.divB = pd(r.divB) if r.divB is not None else None
    r._o = o
    return r


def compute(cfg, fm, bcs, U, jacobian=False, limiter_init=None, update_coeff_init=None, omp=False, out=None):
    """applyBC + one ComputeRHS(+Jacobian) evaluation.  U: [(nCells+nGhosts)*q] AoS, modified in place
    only in its ghost part."""
    r = out if out is not None else alloc_outputs(cfg, fm, jacobian, limiter_init)
    if update_coeff_init is not None:
        r.updateCoeff[:] = update_coeff_init
    mv = mesh_view(fm)
    bca = bc_array(bcs)
    L = lib(omp)
    if L.cfo_apply_bc(C.byref(cfg), C.byref(mv), bca, pd(U)):
        raise RuntimeError("oracle apply_bc failed")
    rc = L.cfo_compute(C.byref(cfg), C.byref(mv), bca, pd(U), C.byref(r._o))
    r.status = rc
    return r


def time_rhs(cfg, fm, bcs, U, pastU, r, omp=False):
    mv = mesh_view(fm)
    rc = lib(omp).cfo_time_rhs(C.byref(cfg), C.byref(mv), bc_array(bcs), pd(U), pd(pastU), C.byref(r._o))
    if rc:
        raise RuntimeError("oracle time_rhs failed")


def run_fwd_euler(cfg, fm, bcs, U, cfl, norm_target, max_iter, monitor_var=0, omp=False):
    r = alloc_outputs(cfg, fm, False)
    hist = np.zeros(max_iter)
    mv = mesh_view(fm)
    n = lib(omp).cfo_run_fwd_euler(C.byref(cfg), C.byref(mv), bc_array(bcs), pd(U), C.byref(r._o),
                                   cfl, monitor_var, norm_target, max_iter, pd(hist))
    if n < 0:
        raise RuntimeError("oracle fwd euler failed: %d" % n)
    r.history = hist[:n]
    return r
