import sys
import os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import cases
from oracle import binding as oracle
from coolfluid_b200 import lib as cl
L = cl.load("/tmp/libhostsim_asan.so", "hostsim_")
names = sys.argv[1:] or sorted(cases.CASES)
bad = 0
for name in names:
    out = cases.run_case(name, oracle, lambda cfg: cl.Engine(cfg, 0, lib=L))
    ok = all(np.array_equal(a, b) for k, (a, b) in ((k, v) for k, v in out.items() if not k.startswith("_") and v[1] is not None))
    print(name, "ok" if ok else "DIFF"); bad += (not ok)
print("done, diffs:", bad)
