"""CPU tier, world_size 2 and 3 over `gloo`: one process per rank, each with its partition (tests/multirank.py).  After K
ForwardEuler iterations with a state halo exchange after every update, the updatable states of all ranks together
equal the single-domain run (to summation order) and the summed residual history matches to 1e-10 (BASELINE.json)."""
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

import multirank

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def launch(world, mode, name, steps, out_dir, timeout=600, renumber=0):
    port = free_port()
    procs = []
    for r in range(world):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE=str(world), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port),
                   LOCAL_RANK=str(r), CFB_TEST_RENUMBER=str(renumber))
        procs.append(subprocess.Popen([sys.executable, os.path.join(ROOT, "tests", "multirank.py"), mode, name, str(steps),
                                       str(out_dir)], env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True))
    outs = []
    for pr in procs:
        try:
            o, _ = pr.communicate(timeout=timeout)
        except subprocess.TimeoutExpired:
            for q in procs:
                q.kill()
            raise
        outs.append(o)
    for r, pr in enumerate(procs):
        assert pr.returncode == 0, "rank %d failed:\n%s" % (r, outs[r][-3000:])


def check_against_single_domain(world, name, steps, out_dir, lib, device=0):
    """The partitions use the order-preserving local numbering (ascending global ID), so everything an updatable cell
    computes is BIT-IDENTICAL to the single-domain run: states after `steps` explicit iterations with a halo exchange
    after each, then the residual, update coefficient and every Jacobian block row of one implicit evaluation
    (fused pseudo-steady term).  Only the residual norm, a sum over ranks, is compared to 1e-10 (BASELINE.json)."""
    Ug, hist, g = multirank.single_domain(name, steps, lib, device=device)
    seen = np.zeros(len(Ug), dtype=bool)
    gslot = {}
    rp, ci = g["rowPtr"], g["colInd"]
    for c in range(len(Ug)):
        for k in range(rp[c], rp[c+1]):
            gslot[(c, int(ci[k]))] = k
    nblocks = 0
    for r in range(world):
        z = np.load(os.path.join(out_dir, "rank%d.npz" % r))
        assert not seen[z["gid"]].any()
        seen[z["gid"]] = True
        assert np.array_equal(z["U"], Ug[z["gid"]]), "rank %d states differ" % r
        # overlap cells hold their owners' latest states as well (the last exchange has landed)
        assert np.array_equal(z["Uall"], Ug[z["gall"]]), "rank %d overlap states are stale" % r
        assert np.abs(z["hist"] - hist).max() <= 1e-10
        assert np.array_equal(z["rhs"], g["rhs"][z["gid"]]), "rank %d residual differs" % r
        assert np.array_equal(z["uc"], g["uc"][z["gid"]]), "rank %d updateCoeff differs" % r
        ks = np.array([gslot[(int(a), int(b))] for a, b in zip(z["jrow"], z["jcol"])])
        assert np.array_equal(z["jac"], g["jac"][ks]), "rank %d Jacobian rows differ" % r
        nblocks += len(ks)
    assert seen.all()
    assert nblocks == len(ci)          # every block of the global matrix was produced by exactly one rank


@pytest.mark.parametrize("name,world,renumber", [("ns3d_hex", 2, 0), ("euler2d_2nd_tri_fullstencil", 3, 0), ("ns3d_tet", 2, 1)])
def test_forward_euler_with_halo_gloo(name, world, renumber, tmp_path, hostsim_lib, oracle):
    steps = 5
    launch(world, "hostsim", name, steps, tmp_path, renumber=renumber)
    check_against_single_domain(world, name, steps, tmp_path, hostsim_lib)
