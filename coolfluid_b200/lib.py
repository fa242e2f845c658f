"""ctypes binding of libcfb200.so (the C ABI declared in include/cfb200.h).

This is plumbing only: it loads the CUDA library built in-tree by ``coolfluid_b200/csrc/Makefile`` and exposes the
entry points with numpy arrays.  It never computes anything itself and has no fallback: if the shared library is
missing, or no CUDA device is usable, it raises.
"""
from __future__ import annotations
import ctypes as C
import os
import numpy as np

from .config import ConfigStruct, BCStruct, bc_array
from .flatmesh import FlatMesh, INT_FIELDS, F64_FIELDS, U8_FIELDS

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libcfb200.so")
P_D = C.POINTER(C.c_double)
P_I = C.POINTER(C.c_int32)
P_U8 = C.POINTER(C.c_uint8)

ERRORS = {1: "BAD_VALUE", 2: "INVALID_ARG", 3: "NO_DEVICE", 4: "CUDA", 5: "STATE", 6: "UNSUPPORTED", 7: "NCCL",
          8: "GHOST_REPOSITION"}


class Cfb200Error(RuntimeError):
    def __init__(self, code, text=""):
        super().__init__("cfb200 error %d (%s) %s" % (code, ERRORS.get(code, "?"), text))
        self.code = code


class MeshStruct(C.Structure):   # cfb200_mesh
    _fields_ = [("dim", C.c_int32), ("nCells", C.c_int32), ("nGhosts", C.c_int32), ("nNodes", C.c_int32),
                ("nFaces", C.c_int32), ("nInner", C.c_int32),
                ("nodes", P_D), ("centres", P_D), ("faceL", P_I), ("faceR", P_I), ("faceNodePtr", P_I),
                ("faceNodes", P_I), ("normals", P_D), ("areas", P_D), ("volumes", P_D),
                ("cellFacePtr", P_I), ("cellFaces", P_I), ("stencilPtr", P_I), ("stencil", P_I),
                ("nodeNbrPtr", P_I), ("nodeNbr", P_I), ("nodeW", P_D), ("cvol", P_D), ("cvSign", P_U8),
                ("faceBC", P_I), ("updatable", P_U8), ("cellNodePtr", P_I), ("cellNodes", P_I)]


class GmresParams(C.Structure):  # cfb200_gmres_params
    _fields_ = [("restart", C.c_int32), ("maxIter", C.c_int32), ("precond", C.c_int32), ("reserved", C.c_int32),
                ("relTol", C.c_double), ("absTol", C.c_double)]


class BsrStruct(C.Structure):    # cfb200_bsr
    _fields_ = [("nRows", C.c_int32), ("nnzb", C.c_int32), ("blockSize", C.c_int32),
                ("rowPtr", C.c_void_p), ("colInd", C.c_void_p), ("val", C.c_void_p)]


# every symbol include/cfb200.h declares (tests check that the library exports all of them)
ABI_SYMBOLS = [
    "abi_version", "device_count", "create", "destroy", "last_error", "set_mesh", "set_step", "set_states",
    "get_states", "set_past_states", "set_bc_face_values", "set_nodal_prescribed", "backup_past_states", "set_limiter", "set_update_coeff", "apply_bc",
    "compute_rhs", "time_rhs", "update_sol_fwd_euler", "update_sol_newton", "bsr_view", "bsr_pattern", "bsr_nnzb",
    "get_bsr_values",
    "export_csr", "get_field", "device_ptr", "status", "enable_timing", "last_timings", "stream",
    "nccl_unique_id", "comm_init", "set_halo", "set_independent_cells", "halo_begin", "halo_exchange", "allreduce_sum",
    "mesh_build", "mesh_build_error", "mesh_sizes", "mesh_get", "mesh_ptr", "mesh_view", "mesh_set_updatable", "mesh_set_stencil", "mesh_free", "mesh_diamond_volumes",
    "selftest_divsqrt", "export_csr_device", "solve_gmres", "l2_norm_noalloc",
]
# symbols the engine part needs (a test-only host build of the kernel functors exports just these)
ENGINE_SYMBOLS = ABI_SYMBOLS[2:30]

_libs = {}


def load(path=None, prefix="cfb200_"):
    """Load the shared library; raises OSError when it has not been built."""
    path = path or LIB_PATH
    key = (path, prefix)
    if key in _libs:
        return _libs[key]
    if not os.path.exists(path):
        raise OSError("%s not found: build it with `make -C coolfluid_b200/csrc` (or __graft_entry__.build())" % path)
    L = C.CDLL(path)
    ns = type("Lib", (), {})()
    ns._cdll = L
    ns.prefix = prefix
    for name in ABI_SYMBOLS:
        try:
            setattr(ns, name, getattr(L, prefix + name))
        except AttributeError:
            if prefix == "cfb200_" or name in ENGINE_SYMBOLS:
                raise
    if hasattr(ns, "l2_norm_noalloc"):
        ns.l2_norm_noalloc.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_int32, C.c_void_p,
                                       C.c_void_p, C.c_char_p, C.c_int32]
    if hasattr(ns, "selftest_divsqrt"):
        ns.selftest_divsqrt.argtypes = [C.c_int, C.c_int64, C.c_uint64, C.c_int, C.POINTER(C.c_int64)]
    H = C.c_void_p
    ns.create.argtypes = [C.POINTER(ConfigStruct), C.c_int, C.POINTER(H)]
    ns.destroy.argtypes = [H]; ns.destroy.restype = None
    ns.last_error.argtypes = [H]; ns.last_error.restype = C.c_char_p
    ns.set_mesh.argtypes = [H, C.POINTER(MeshStruct), C.c_int, C.POINTER(BCStruct)]
    ns.set_step.argtypes = [H, C.POINTER(ConfigStruct)]
    for f in (ns.set_states, ns.get_states, ns.set_past_states):
        f.argtypes = [H, P_D, C.c_int64]
    ns.backup_past_states.argtypes = [H]
    ns.set_bc_face_values.argtypes = [H, C.c_int, P_D]
    ns.set_nodal_prescribed.argtypes = [H, P_U8, C.c_int32, P_I, P_D]
    ns.set_limiter.argtypes = [H, P_D]
    ns.set_update_coeff.argtypes = [H, P_D]
    ns.apply_bc.argtypes = [H]
    ns.compute_rhs.argtypes = [H, C.c_int, P_D, P_D]
    ns.time_rhs.argtypes = [H, P_D]
    ns.update_sol_fwd_euler.argtypes = [H, C.c_double, C.c_int, P_D]
    ns.update_sol_newton.argtypes = [H, P_D, P_D, C.c_int, P_D]
    ns.bsr_view.argtypes = [H, C.POINTER(BsrStruct)]
    ns.bsr_pattern.argtypes = [H, P_I, P_I]
    ns.bsr_nnzb.argtypes = [H]; ns.bsr_nnzb.restype = C.c_int64
    ns.get_bsr_values.argtypes = [H, P_D]
    ns.export_csr.argtypes = [H, P_I, P_I, P_D]
    ns.get_field.argtypes = [H, C.c_char_p, P_D]
    ns.device_ptr.argtypes = [H, C.c_char_p]; ns.device_ptr.restype = C.c_void_p
    ns.status.argtypes = [H]
    ns.enable_timing.argtypes = [H, C.c_int]
    ns.last_timings.argtypes = [H, C.POINTER(C.c_char_p), C.POINTER(C.c_float), C.c_int]
    ns.stream.argtypes = [H]; ns.stream.restype = C.c_void_p
    if hasattr(ns, "mesh_build"):
        ns.device_count.restype = C.c_int
        ns.mesh_build.restype = C.c_void_p
        ns.mesh_build.argtypes = [C.c_int, C.c_int, P_D, C.c_int, P_I, P_I, P_I, C.c_int, P_I, P_I, P_I, P_I]
        ns.mesh_build_error.argtypes = []; ns.mesh_build_error.restype = C.c_char_p
        ns.mesh_sizes.argtypes = [C.c_void_p, C.POINTER(C.c_int64)]; ns.mesh_sizes.restype = None
        ns.mesh_get.argtypes = [C.c_void_p, C.c_char_p, C.c_void_p]
        ns.mesh_ptr.argtypes = [C.c_void_p, C.c_char_p]; ns.mesh_ptr.restype = C.c_void_p
        ns.mesh_view.argtypes = [C.c_void_p, C.POINTER(MeshStruct)]; ns.mesh_view.restype = None
        ns.mesh_set_updatable.argtypes = [C.c_void_p, P_U8]; ns.mesh_set_updatable.restype = None
        ns.mesh_set_stencil.argtypes = [C.c_void_p, C.c_int]; ns.mesh_set_stencil.restype = C.c_int
        ns.mesh_free.argtypes = [C.c_void_p]; ns.mesh_free.restype = None
        ns.mesh_diamond_volumes.argtypes = [C.POINTER(MeshStruct), P_D, P_U8]; ns.mesh_diamond_volumes.restype = None
        ns.export_csr_device.argtypes = [H, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
        ns.solve_gmres.argtypes = [H, C.POINTER(GmresParams), C.POINTER(C.c_int32), P_D]
        ns.nccl_unique_id.argtypes = [C.c_void_p]
        ns.comm_init.argtypes = [H, C.c_int, C.c_int, C.c_void_p]
        ns.set_halo.argtypes = [H, P_I, P_I, P_I, P_I, C.c_int32]
        ns.set_independent_cells.argtypes = [H, P_I, C.c_int32]
        ns.halo_begin.argtypes = [H]
        ns.halo_exchange.argtypes = [H]
        ns.allreduce_sum.argtypes = [H, P_D, C.c_int]
    _libs[key] = ns
    return ns


def _pd(a):
    return a.ctypes.data_as(P_D)


def _pi(a):
    return a.ctypes.data_as(P_I)


def _pu8(a):
    return a.ctypes.data_as(P_U8)


STENCIL_KINDS = {"FaceVertexPlusGhost": 0, "FaceVertex": 1, "FaceEdge": 2, "Face": 3}     # Setup1.stencil


def build_mesh(em, lib=None, stencil=None) -> FlatMesh:
    """Host mesh-data construction (cfb200_mesh_build): element description -> the flat arrays of the FVMCC sockets."""
    stencil = stencil or getattr(em, "stencil_kind", "FaceVertexPlusGhost")   # Setup1.stencil of the case (tests/cases.with_stencil)
    L = lib or load()
    h = L.mesh_build(em.dim, em.n_nodes, _pd(em.nodes), em.n_cells, _pi(em.cell_node_ptr), _pi(em.cell_nodes),
                     _pi(em.cell_shape), em.n_trs, _pi(em.trs_face_ptr), _pi(em.trs_face_cell),
                     _pi(em.trs_face_node_ptr), _pi(em.trs_face_nodes))
    if not h:
        raise Cfb200Error(2, "mesh build failed: " + (L.mesh_build_error() or b"").decode())
    try:
        if STENCIL_KINDS[stencil] and L.mesh_set_stencil(h, STENCIL_KINDS[stencil]):
            raise Cfb200Error(2, "stencil rebuild failed")
        sz = (C.c_int64*8)()
        L.mesh_sizes(h, sz)
        fm = FlatMesh(em.dim, em.nodes, em.cell_node_ptr, em.cell_nodes, em.cell_shape, em.trs_names)
        fm.nFaces, fm.nInner, fm.nGhosts = int(sz[0]), int(sz[1]), int(sz[2])
        var = {"faceNodes": sz[3], "cellFaces": sz[4], "stencil": sz[5], "nodeNbr": sz[6], "nodeW": sz[6], "bsrColInd": sz[7]}
        for name in INT_FIELDS + F64_FIELDS + U8_FIELDS:
            n = int(var[name]) if name in var else fm.sizes_for(name)
            dt = np.int32 if name in INT_FIELDS else (np.float64 if name in F64_FIELDS else np.uint8)
            a = np.zeros(n, dtype=dt)
            if L.mesh_get(h, name.encode(), a.ctypes.data_as(C.c_void_p)):
                raise KeyError(name)
            setattr(fm, name, a)
    finally:
        L.mesh_free(h)
    return fm


def mesh_struct(fm: FlatMesh) -> MeshStruct:
    ms = MeshStruct()
    ms.dim, ms.nCells, ms.nGhosts, ms.nNodes, ms.nFaces, ms.nInner = fm.dim, fm.nCells, fm.nGhosts, fm.nNodes, fm.nFaces, fm.nInner
    for name, _t in MeshStruct._fields_[6:]:
        a = getattr(fm, name)
        setattr(ms, name, _pd(a) if a.dtype == np.float64 else (_pi(a) if a.dtype == np.int32 else _pu8(a)))
    return ms


class Engine:
    """One cfb200 handle: device-resident mesh, states, residual and block Jacobian of one partition."""

    def __init__(self, cfg: ConfigStruct, device=0, lib=None):
        self.L = lib or load()
        self.cfg = cfg
        self.q = cfg.nEqs
        self.h = C.c_void_p()
        rc = self.L.create(C.byref(cfg), device, C.byref(self.h))
        if rc:
            self.h = None
            raise Cfb200Error(rc, "cfb200_create")
        self.fm = None

    def _chk(self, rc, what=""):
        if rc:
            raise Cfb200Error(rc, what + ": " + (self.L.last_error(self.h) or b"").decode())

    def close(self):
        if getattr(self, "h", None):
            self.L.destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_mesh(self, fm: FlatMesh, bcs):
        self.fm = fm
        ms = mesh_struct(fm)
        self._chk(self.L.set_mesh(self.h, C.byref(ms), len(bcs), bc_array(bcs)), "set_mesh")
        self.nnzb = int(self.L.bsr_nnzb(self.h))

    def set_step(self, cfg=None):
        if cfg is not None:
            self.cfg = cfg
        self._chk(self.L.set_step(self.h, C.byref(self.cfg)), "set_step")

    def set_states(self, U):
        U = np.ascontiguousarray(U, dtype=np.float64).reshape(-1)
        self._chk(self.L.set_states(self.h, _pd(U), self.fm.nCells), "set_states")

    def get_states(self):
        U = np.empty(self.fm.nCells*self.q)
        self._chk(self.L.get_states(self.h, _pd(U), self.fm.nCells), "get_states")
        return U

    def set_past_states(self, U):
        U = np.ascontiguousarray(U, dtype=np.float64).reshape(-1)
        self._chk(self.L.set_past_states(self.h, _pd(U), self.fm.nCells), "set_past_states")

    def backup_past_states(self):
        self._chk(self.L.backup_past_states(self.h), "backup_past_states")

    def set_limiter(self, lim):
        lim = np.ascontiguousarray(lim, dtype=np.float64).reshape(-1)
        self._chk(self.L.set_limiter(self.h, _pd(lim)), "set_limiter")

    def set_bc_face_values(self, bc_index, values=None):
        pv = None if values is None else _pd(np.ascontiguousarray(values, dtype=np.float64))
        self._chk(self.L.set_bc_face_values(self.h, bc_index, pv), "set_bc_face_values")

    def set_nodal_prescribed(self, node_flag=None, idx=(), values=()):
        """NodalExtrapolation = DistanceBasedGMove: ValuesIdx / Values imposed on the flagged nodes (None reverts)."""
        if node_flag is None:
            self._chk(self.L.set_nodal_prescribed(self.h, None, 0, None, None), "set_nodal_prescribed")
            return
        fl = np.ascontiguousarray(node_flag, dtype=np.uint8)
        ix = np.ascontiguousarray(idx, dtype=np.int32)
        vv = np.ascontiguousarray(values, dtype=np.float64)
        self._chk(self.L.set_nodal_prescribed(self.h, _pu8(fl), len(ix), _pi(ix), _pd(vv)), "set_nodal_prescribed")

    def set_update_coeff(self, uc=None):
        if uc is None:
            self._chk(self.L.set_update_coeff(self.h, None), "set_update_coeff")
        else:
            uc = np.ascontiguousarray(uc, dtype=np.float64)
            self._chk(self.L.set_update_coeff(self.h, _pd(uc)), "set_update_coeff")

    def apply_bc(self):
        self._chk(self.L.apply_bc(self.h), "apply_bc")

    def compute_rhs(self, apply_bc=True, download=True, out_rhs=None, out_uc=None):
        """Returns (rhs [nCells*q], updateCoeff [nCells]) when download, else (None, None); raises on BAD_VALUE."""
        if download:
            rhs = out_rhs if out_rhs is not None else np.empty(self.fm.nCells*self.q)
            uc = out_uc if out_uc is not None else np.empty(self.fm.nCells)
            self._chk(self.L.compute_rhs(self.h, int(apply_bc), _pd(rhs), _pd(uc)), "compute_rhs")
            return rhs, uc
        self._chk(self.L.compute_rhs(self.h, int(apply_bc), None, None), "compute_rhs")
        return None, None

    def time_rhs(self, download=True):
        if download:
            rhs = np.empty(self.fm.nCells*self.q)
            self._chk(self.L.time_rhs(self.h, _pd(rhs)), "time_rhs")
            return rhs
        self._chk(self.L.time_rhs(self.h, None), "time_rhs")

    def update_sol_fwd_euler(self, cfl, monitor_var=0):
        r = C.c_double()
        self._chk(self.L.update_sol_fwd_euler(self.h, cfl, monitor_var, C.byref(r)), "update_sol")
        return r.value

    def update_sol_newton(self, dU=None, alpha=None, monitor_var=0):
        """StdUpdateSol on the device: U += alpha*dU (dU = None: the increment already in the device rhs buffer)."""
        r = C.c_double()
        pdu = None if dU is None else _pd(np.ascontiguousarray(dU, dtype=np.float64))
        pal = None if alpha is None else _pd(np.ascontiguousarray(alpha, dtype=np.float64))
        self._chk(self.L.update_sol_newton(self.h, pdu, pal, monitor_var, C.byref(r)), "update_sol_newton")
        return r.value

    def bsr_pattern(self):
        rp = np.empty(self.fm.nCells + 1, dtype=np.int32)
        ci = np.empty(self.nnzb, dtype=np.int32)
        self._chk(self.L.bsr_pattern(self.h, _pi(rp), _pi(ci)), "bsr_pattern")
        return rp, ci

    def bsr_values(self):
        v = np.empty(self.nnzb*self.q*self.q)
        self._chk(self.L.get_bsr_values(self.h, _pd(v)), "get_bsr_values")
        return v

    def bsr_view(self):
        b = BsrStruct()
        self._chk(self.L.bsr_view(self.h, C.byref(b)), "bsr_view")
        return b

    def export_csr(self):
        n = self.fm.nCells*self.q
        ro = np.empty(n + 1, dtype=np.int32)
        col = np.empty(self.nnzb*self.q*self.q, dtype=np.int32)
        val = np.empty(self.nnzb*self.q*self.q)
        self._chk(self.L.export_csr(self.h, _pi(ro), _pi(col), _pd(val)), "export_csr")
        return ro, col, val

    def solve_gmres(self, restart=30, max_iter=50, rel_tol=1e-5, abs_tol=1e-30, precond=1):
        """J x = rhs on the device (block-Jacobi GMRES); the solution replaces the device rhs buffer.
        Returns (iterations, final relative residual)."""
        p = GmresParams(restart, max_iter, precond, 0, rel_tol, abs_tol)
        it = C.c_int32()
        rr = C.c_double()
        self._chk(self.L.solve_gmres(self.h, C.byref(p), C.byref(it), C.byref(rr)), "solve_gmres")
        return it.value, rr.value

    def export_csr_device(self, row_off_ptr, col_ptr, val_ptr, index64=False):
        self._chk(self.L.export_csr_device(self.h, row_off_ptr, col_ptr, val_ptr, int(index64)), "export_csr_device")

    def field(self, name):
        fm, q = self.fm, self.q
        n = {"rhs": fm.nCells*q, "updateCoeff": fm.nCells, "uX": fm.nCells*q, "uY": fm.nCells*q, "uZ": fm.nCells*q,
             "limiter": fm.nCells*q, "nstates": fm.nNodes*q, "states": (fm.nCells + fm.nGhosts)*q,
             "gstates": fm.nGhosts*q, "faceFlux": fm.nFaces*q, "faceCentres": fm.nFaces*fm.dim,
             "bndBlocks": fm.nGhosts*q*q, "divB": fm.nCells}[name]
        a = np.empty(n)
        self._chk(self.L.get_field(self.h, name.encode(), _pd(a)), "get_field " + name)
        return a

    def enable_timing(self, on=True):
        self._chk(self.L.enable_timing(self.h, int(on)), "enable_timing")

    def last_timings(self):
        names = (C.c_char_p*64)()
        ms = (C.c_float*64)()
        n = self.L.last_timings(self.h, names, ms, 64)
        return [(names[i].decode(), float(ms[i])) for i in range(n)]

    def stream(self):
        return self.L.stream(self.h)

    # ---- multi-GPU (one partition per GPU) ----
    def comm_init(self, rank, nranks, unique_id: bytes):
        buf = C.create_string_buffer(unique_id, 128)
        self._chk(self.L.comm_init(self.h, rank, nranks, buf), "comm_init")

    def set_halo(self, part):
        """part: coolfluid_b200.partition.Partition of this rank"""
        sp, sc, rp, rc = part.halo_csr()
        self._halo_keep = (sp, sc, rp, rc)
        indep = getattr(part, "indep", None)
        self._chk(self.L.set_halo(self.h, _pi(sp), _pi(sc), _pi(rp), _pi(rc), 0 if indep is not None else int(part.n_interior)),
                  "set_halo")
        if indep is not None:
            indep = np.ascontiguousarray(indep, dtype=np.int32)
            self._halo_keep += (indep,)
            self._chk(self.L.set_independent_cells(self.h, _pi(indep), len(indep)), "set_independent_cells")

    def halo_begin(self):
        self._chk(self.L.halo_begin(self.h), "halo_begin")

    def halo_exchange(self):
        self._chk(self.L.halo_exchange(self.h), "halo_exchange")

    def allreduce_sum(self, values):
        v = np.ascontiguousarray(values, dtype=np.float64).copy()
        self._chk(self.L.allreduce_sum(self.h, _pd(v), len(v)), "allreduce_sum")
        return v


def nccl_unique_id(lib=None) -> bytes:
    L = lib or load()
    buf = C.create_string_buffer(128)
    rc = L.nccl_unique_id(buf)
    if rc:
        raise Cfb200Error(rc, "nccl_unique_id")
    return buf.raw
