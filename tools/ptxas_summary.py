"""Summarise `nvcc -Xptxas -v` output (coolfluid_b200/csrc/ptxas_capi.log): registers, stack, spills per kernel."""
import re, subprocess, sys
s = open(sys.argv[1] if len(sys.argv) > 1 else "coolfluid_b200/csrc/ptxas_capi.log").read()
seen = set()
for m in re.finditer(r"Compiling entry function '(\S+)' for 'sm_100a'\nptxas info\s+: Function properties for \S+\n\s+(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads\nptxas info\s+: Used (\d+) registers", s):
    dn = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
    k = re.search(r"k_run<cfb::(\w+(<[^>]*>)?), (\d+)>", dn)
    name = (k.group(1) + " bs" + k.group(3)) if k else dn[:50]
    if name in seen: continue
    seen.add(name)
    print(name.ljust(48), "regs", m.group(5).rjust(3), " stack", m.group(2).rjust(5), " spill st/ld", m.group(3), m.group(4))
