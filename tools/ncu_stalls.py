"""Per-source-line stall samples / local-memory sectors of an `ncu --page source --csv --print-source cuda,sass` dump.
argv: dump.csv kernel-substring source-file [top]"""
import csv, collections, sys
rows = list(csv.reader(open(sys.argv[1]))); top = int(sys.argv[4]) if len(sys.argv) > 4 else 14
src = open(sys.argv[3]).read().split("\n"); sname = sys.argv[3].split("/")[-1]
cur = ""; hdr = None; active = False; seen = set()
cols = {"long_sb": "stall_long_sb", "barrier": "stall_barrier", "short_sb": "stall_short_sb", "wait": "stall_wait", "no_inst": "stall_no_inst",
        "local": "L2 Theoretical Sectors Local", "samples": "# Samples"}
agg = {k: collections.Counter() for k in cols}
for r in rows:
    if not r: continue
    if r[0] == "File Path": cur = r[1].split("/")[-1]; continue
    if r[0] == "Function Name":
        fn = r[1]; active = (sys.argv[2] in fn) and (fn, cur) not in seen; seen.add((fn, cur)); continue
    if r[0] == "Line No": hdr = r; continue
    if not active or hdr is None or r[2] != "-": continue
    for k, c in cols.items():
        try: agg[k][(cur, int(r[0]))] += float(r[hdr.index(c)] or 0)
        except (ValueError, IndexError): pass
for k in cols:
    print(f"== {k}: total {sum(agg[k].values()):.0f}")
    for (f, l), v in agg[k].most_common(top):
        text = src[l - 1].strip()[:90] if f == sname and l <= len(src) else ""
        print(f"  {v:9.0f}  {f}:{l}  {text}")
