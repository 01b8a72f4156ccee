"""Developer tool: per-source-line executed-instruction shares of one kernel in an ncu report (--import-source on),
sorted by instruction count.  usage: python tools/ncu_lines.py report.ncu-rep [launch_index] [top_n]"""
import collections, csv, io, subprocess, sys
rep = sys.argv[1]; kid = sys.argv[2] if len(sys.argv) > 2 else "0"; topn = int(sys.argv[3]) if len(sys.argv) > 3 else 60
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv", "--launch-skip", kid,
                      "--launch-count", "1"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
cur = None; agg = collections.OrderedDict(); ti = 0
for r in rows:
    if len(r) == 2 and r[0] == 'File Path': cur = r[1].split('/')[-1]; continue
    if len(r) < 10 or r[0] == '' or r[0] == 'Line No': continue
    try: line = int(r[0]); samp = int(r[6]); inst = int(r[7])
    except ValueError: continue
    a = agg.setdefault((cur, line), [0, 0, r[1]]); a[0] += inst; a[1] += samp; ti += inst
print("total warp instructions", ti)
for (f, l), (i, s, src) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:topn]:
    print(f"{f}:{l:4d} inst {100*i/ti:5.1f}% samp {s:5d} {src.strip()[:100]}")
if len(sys.argv) > 4:      # region sums: name=file:lo-hi,...
    for spec in sys.argv[4:]:
        name, rng = spec.split("=")
        tot = 0; sm = 0
        for part in rng.split(","):
            f, lh = part.split(":"); lo, hi = map(int, lh.split("-"))
            for (ff, l), (i, s, src) in agg.items():
                if ff == f and lo <= l <= hi: tot += i; sm += s
        print(f"region {name:12s} inst {100*tot/ti:5.1f}%  samples {sm}")
