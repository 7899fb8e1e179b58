"""Aggregate an `ncu --page source --csv --print-source cuda,sass` dump by CUDA source line."""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
kernel_filter = sys.argv[2] if len(sys.argv) > 2 else ""
agg = collections.OrderedDict(); cur_file = ""; fn = ""; seen_fn = set(); active = False
hdr = None
for r in rows:
    if not r: continue
    if r[0] == "File Path": cur_file = r[1].split("/")[-1]; continue
    if r[0] == "Function Name":
        fn = r[1]; active = (kernel_filter in fn) and (fn, cur_file) not in seen_fn; seen_fn.add((fn, cur_file)); continue
    if r[0] == "Line No": hdr = r; continue
    if not active or hdr is None: continue
    if r[2] != "-": continue  # only the per-line summary rows
    try:
        ie = hdr.index("Instructions Executed"); ss = hdr.index("# Samples")
        key = (cur_file, int(r[0]), r[1].strip())
        a = agg.setdefault(key, [0.0, 0.0])
        a[0] += float(r[ie]); a[1] += float(r[ss])
    except (ValueError, IndexError):
        pass
ti = sum(a[0] for a in agg.values()); ts = sum(a[1] for a in agg.values())
print(f"total warp-instructions {ti:.3e}, samples {ts:.0f}")
for (f, ln, src), (i, s) in sorted(agg.items(), key=lambda kv: -kv[1][0])[: int(sys.argv[3]) if len(sys.argv) > 3 else 40]:
    print(f"{i/ti*100:5.1f}% inst {s/max(ts,1)*100:5.1f}% smp  {f}:{ln:<4} {src[:100]}")
