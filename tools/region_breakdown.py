"""Region-level instruction breakdown of a kernel from `ncu --page source --csv --print-source cuda,sass`."""
import csv, collections, sys
rows = list(csv.reader(open(sys.argv[1]))); srcfile = sys.argv[2]; nwarps = float(sys.argv[3])
agg = collections.Counter(); smp = collections.Counter(); cur = ""; hdr = None; seen = set(); active = False
for r in rows:
    if not r: continue
    if r[0] == "File Path": cur = r[1].split("/")[-1]; continue
    if r[0] == "Function Name":
        key = (r[1], cur); active = key not in seen; seen.add(key); continue
    if r[0] == "Line No": hdr = r; continue
    if not active or hdr is None or r[2] != "-": continue
    try:
        ie = hdr.index("Instructions Executed"); ss = hdr.index("# Samples")
        k = int(r[0]) if cur == srcfile.split("/")[-1] else cur
        agg[k] += float(r[ie]); smp[k] += float(r[ss])
    except Exception:
        pass
src = open(srcfile).read().split("\n")
marks = [(i + 1, l.strip()) for i, l in enumerate(src) if l.strip().startswith("// ----") or l.strip().startswith("// ====")]
first_kernel_line = next(i + 1 for i, l in enumerate(src) if "__global__" in l)
tot = sum(agg.values()); ts = sum(smp.values())
print(f"total warp-instructions {tot:.3e}; per warp {tot/nwarps:.0f}")
reg = {}
for k, v in agg.items():
    if isinstance(k, int) and k >= first_kernel_line:
        name = "(kernel prologue)"
        for m, l in marks:
            if m <= k and m >= first_kernel_line: name = l[:72]
    else:
        name = "inlined helpers: " + (k if not isinstance(k, int) else "vec/mat fns of this file")
    a = reg.setdefault(name, [0, 0]); a[0] += v; a[1] += smp[k]
for n, (v, s_) in sorted(reg.items(), key=lambda kv: -kv[1][0]):
    print(f"{v/tot*100:5.1f}% inst ({v/nwarps:6.0f}/warp) {s_/max(ts,1)*100:5.1f}% smp  {n}")
