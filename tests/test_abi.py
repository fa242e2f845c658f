"""The C-ABI shared library loads and exports every symbol include/cfb200.h declares; without a GPU the
compute entry points fail loudly (no CPU fallback in the product)."""
import ctypes as C
import os
import re

import pytest

from conftest import ROOT, has_gpu
from coolfluid_b200 import lib as cl, config as B


def test_header_symbols_exported(product_lib):
    hdr = open(os.path.join(ROOT, "include", "cfb200.h")).read()
    declared = sorted(set(re.findall(r"\b(cfb200_[a-z0-9_]+)\s*\(", hdr)))
    assert len(declared) >= 40
    dll = product_lib._cdll
    missing = [s for s in declared if not hasattr(dll, s)]
    assert missing == []
    assert sorted("cfb200_" + s for s in cl.ABI_SYMBOLS) == declared
    assert dll.cfb200_abi_version() == B.ABI_VERSION == 3


def test_struct_layouts_match_the_oracle(oracle):
    """cfb200_config / cfb200_bc / cfb200_mesh are shared by value with the oracle's structs."""
    L = oracle.lib()
    assert L.cfo_sizeof_config() == C.sizeof(B.ConfigStruct)
    assert L.cfo_sizeof_bc() == C.sizeof(B.BCStruct)
    # the oracle's view = cfb200_mesh + trailing optional entries: per-face Dirichlet data (two pointers; the product
    # receives them through cfb200_set_bc_face_values) and the prescribed nodal values of DistanceBasedGMove (three
    # pointers and a count, padded; the product: cfb200_set_nodal_prescribed)
    assert L.cfo_sizeof_meshview() == C.sizeof(cl.MeshStruct) + 6*C.sizeof(C.c_void_p)


def test_product_does_not_link_the_oracle():
    out = os.popen("ldd %s" % cl.LIB_PATH).read()
    assert "oracle" not in out and "hostsim" not in out
    src = os.path.join(ROOT, "coolfluid_b200")
    for dp, _dn, fns in os.walk(src):
        for fn in fns:
            if fn.endswith((".py", ".cu", ".cuh", ".cpp", ".hpp", ".inc", ".h")):
                text = open(os.path.join(dp, fn)).read()
                assert "import oracle" not in text and "from oracle" not in text and "liboracle" not in text, fn


@pytest.mark.skipif(has_gpu(), reason="a GPU is present")
def test_no_device_fails_loudly(product_lib):
    cfg = B.make_config(2)
    with pytest.raises(cl.Cfb200Error) as ei:
        cl.Engine(cfg, 0, lib=product_lib)
    assert ei.value.code == 3      # CFB200_ERR_NO_DEVICE
    assert product_lib.device_count() == 0


def test_invalid_arguments(product_lib):
    h = C.c_void_p()
    cfg = B.make_config(2)
    cfg.nEqs = 7
    assert product_lib.create(C.byref(cfg), 0, C.byref(h)) == 2
    cfg = B.make_config(2, flux=B.FS_AUSMPLUS, choice_a12=5)      # TCNEQ interface sound speed: unsupported
    assert product_lib.create(C.byref(cfg), 0, C.byref(h)) == 6
    cfg = B.make_config(2, flux=B.FS_AUSMPLUSUP)                  # AUSM+up needs machInf > 0 (AUSMPlusUpFlux.ci:55-57)
    assert product_lib.create(C.byref(cfg), 0, C.byref(h)) == 6
    cfg = B.make_config(2, diffusive=B.DIFF_NS, update_var=B.UV_PVT, derivative="corrected")   # needs the LSQ gradients
    assert product_lib.create(C.byref(cfg), 0, C.byref(h)) == 6
    cfg = B.make_config(2, diffusive=B.DIFF_NS, update_var=B.UV_PVT, poly_rec=B.PR_LINEARLS)
    cfg.derivativeStrategy = 4                                    # not a derivative computer
    assert product_lib.create(C.byref(cfg), 0, C.byref(h)) == 6
    cfg = B.make_config(2, flux=B.FS_AUSMPLUSUP, mach_inf=0.3, diffusive=B.DIFF_NS, update_var=B.UV_PVT,
                        jacobian="analytic", derivative="correctedgg")      # getGradientsJacob of CorrectedGG is a stub
    assert product_lib.create(C.byref(cfg), 0, C.byref(h)) == 6
    assert product_lib.compute_rhs(None, 1, None, None) == 2


def test_graft_entry_build_passes(product_lib):
    """__graft_entry__.build() is what the driver runs first: its own consistency checks (ABI version of the built library
    against the Python mirrors) must hold for the tree as it is."""
    import importlib
    import sys
    sys.path.insert(0, ROOT)
    ge = importlib.import_module("__graft_entry__")
    ge.build()
