#!/usr/bin/env python
"""Per-source-region profile of one kernel from `ncu -i rep --page source --csv --print-source cuda,sass`:
warp instructions executed, FP64 instructions and stall samples attributed to source regions (file:line ranges
named in REGIONS) of the face kernel.  usage: ncu_src_regions.py dump.csv [n_items]  -> markdown table"""
import csv, sys, re, collections

REGIONS = [  # (name, file suffix, first line, last line) -- lines of coolfluid_b200/csrc at the profiled commit
]

def load(path):
    rows = list(csv.reader(open(path)))
    out = []      # (file, line, sass, samples, executed, stalls{})
    cur_file = None; hdr = None; cur_line = None
    for r in rows:
        if not r: continue
        if r[0] == "File Path": cur_file = r[1]; continue
        if r[0] == "Line No": hdr = r; continue
        if r[0] == "Function Name" or hdr is None: continue
        ix = {h: i for i, h in enumerate(hdr)}
        if r[0] != "":
            cur_line = int(r[0]); continue      # aggregate row of the source line; SASS rows follow
        sass = r[3]
        d = {h: r[i] for h, i in ix.items() if h.startswith("stall_") and "Not Issued" not in h}
        # hdr has two 'Source' columns; index explicitly
        num = lambda v: int(float(v)) if v not in ("", "-") else 0
        samples = num(r[6]); executed = num(r[7])
        out.append((cur_file, cur_line, sass, samples, executed, {k: num(v) for k, v in d.items()}))
    return out

if __name__ == "__main__":
    data = load(sys.argv[1])
    items = float(sys.argv[2]) if len(sys.argv) > 2 else None
    tot_s = sum(d[3] for d in data); tot_e = sum(d[4] for d in data)
    by = collections.OrderedDict()
    for f, line, sass, s, e, st in data:
        key = (f.split("/")[-1], line)
        a = by.setdefault(key, [0, 0, 0, collections.Counter()])
        a[0] += s; a[1] += e
        if re.match(r"\s*(@!?U?P\d+\s+)?D(FMA|MUL|ADD|SETP)", sass): a[2] += e
        a[3].update(st)
    print("total samples %d, warp instructions %d" % (tot_s, tot_e))
    print("| file:line | samples % | warp instr % | fp64 instr | top stalls |")
    print("|---|---:|---:|---:|---|")
    for (f, line), (s, e, f64, st) in sorted(by.items(), key=lambda kv: -kv[1][0])[:int(sys.argv[3]) if len(sys.argv) > 3 else 45]:
        top = ", ".join("%s %.0f%%" % (k.replace("stall_", ""), 100.0*v/max(s, 1)) for k, v in st.most_common(3))
        print("| %s:%d | %.1f | %.1f | %d | %s |" % (f, line, 100.0*s/tot_s, 100.0*e/tot_e, f64, top))
