"""In-tree build of libcfb200.so (nvcc, sm_100a) with a source hash beside it, so that tests never run against a
stale binary: the hash, not file times, decides (the GPU box receives a snapshot whose times are not the build's)."""
import glob
import hashlib
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libcfb200.so")
STAMP = os.path.join(HERE, "libcfb200.srchash")      # travels with the .so (not git-tracked)


def source_hash():
    files = sorted(glob.glob(os.path.join(CSRC, "*.cu")) + glob.glob(os.path.join(CSRC, "*.cuh")) +
                   glob.glob(os.path.join(CSRC, "*.hpp")) + glob.glob(os.path.join(CSRC, "*.inc")) +
                   glob.glob(os.path.join(CSRC, "*.cpp")) + [os.path.join(CSRC, "Makefile"),
                                                            os.path.join(HERE, "..", "include", "cfb200.h")])
    h = hashlib.sha1()
    for f in files:
        h.update(os.path.basename(f).encode())
        h.update(open(f, "rb").read())
    return h.hexdigest()


def up_to_date():
    return os.path.exists(LIB) and os.path.exists(STAMP) and open(STAMP).read().strip() == source_hash()


def have_nvcc():
    import shutil
    return bool(shutil.which("nvcc")) or os.path.exists("/usr/local/cuda/bin/nvcc")


def ensure(force=False):
    """Builds the library when its sources changed since the last build (or when it is missing)."""
    if up_to_date() and not force:
        return LIB
    if not have_nvcc():
        if os.path.exists(LIB):
            raise OSError("libcfb200.so is older than its sources and nvcc is not available to rebuild it")
        raise OSError("libcfb200.so is missing and nvcc is not available to build it")
    subprocess.check_call(["make", "-s", "-B", "-C", CSRC])
    with open(STAMP, "w") as f:
        f.write(source_hash() + "\n")
    return LIB
