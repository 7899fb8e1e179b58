"""Instructions / samples of an `ncu --page source --csv --print-source cuda,sass` dump summed over source line ranges.
argv: dump.csv kernel-substring file:lo-hi [file:lo-hi ...]"""
import csv, collections, sys
rows = list(csv.reader(open(sys.argv[1])))
agg = collections.OrderedDict(); cur_file = ""; active = False; hdr = None; seen = set()
for r in rows:
    if not r: continue
    if r[0] == "File Path": cur_file = r[1].split("/")[-1]; continue
    if r[0] == "Function Name":
        fn = r[1]; active = (sys.argv[2] in fn) and (fn, cur_file) not in seen; seen.add((fn, cur_file)); continue
    if r[0] == "Line No": hdr = r; continue
    if not active or hdr is None or r[2] != "-": continue
    try:
        ie = hdr.index("Instructions Executed"); ss = hdr.index("# Samples")
        a = agg.setdefault((cur_file, int(r[0])), [0.0, 0.0]); a[0] += float(r[ie]); a[1] += float(r[ss])
    except (ValueError, IndexError): pass
ti = sum(a[0] for a in agg.values()); ts = sum(a[1] for a in agg.values())
print(f"total {ti/1e6:.2f} M warp-instructions, {ts:.0f} samples")
for spec in sys.argv[3:]:
    f, r2 = spec.split(":"); lo, hi = map(int, r2.split("-"))
    i = sum(a[0] for (ff, l), a in agg.items() if ff == f and lo <= l <= hi)
    s = sum(a[1] for (ff, l), a in agg.items() if ff == f and lo <= l <= hi)
    print(f"{spec:40s} {i/1e6:7.2f} M inst ({i/ti*100:5.1f} %), samples {s/max(ts,1)*100:5.1f} %")
