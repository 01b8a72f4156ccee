"""Developer tool: per-source-line instruction / stall-sample shares from an ncu report (--import-source on).
usage: python tools/ncu_regions.py report.ncu-rep [top_n]"""
import collections, csv, io, subprocess, sys
rep = sys.argv[1]; topn = int(sys.argv[2]) if len(sys.argv) > 2 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
cur = None; agg = collections.OrderedDict(); ti = ts = 0; hdr = None
stall_cols = {}
for r in rows:
    if len(r) == 2 and r[0] == 'File Path': cur = r[1].split('/')[-1]; continue
    if len(r) > 10 and r[0] == 'Line No':
        hdr = r; stall_cols = {h: i for i, h in enumerate(r) if h.startswith('stall_') and 'Not Issued' not in h}; continue
    if len(r) < 10 or r[0] == '': continue
    try: line = int(r[0]); samp = int(r[6]); inst = int(r[7])
    except ValueError: continue
    a = agg.setdefault((cur, line), [0, 0, r[1], collections.Counter()])
    a[0] += inst; a[1] += samp; ti += inst; ts += samp
    for h, i in stall_cols.items():
        try: a[3][h] += int(r[i])
        except ValueError: pass
print("total warp inst", ti, "samples", ts)
tot_st = collections.Counter()
for v in agg.values(): tot_st.update(v[3])
print("stalls:", ", ".join(f"{k[6:]} {100*v/max(ts,1):.1f}%" for k, v in tot_st.most_common(10)))
for (f, l), (i, s, src, st) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:topn]:
    top = ",".join(f"{k[6:]}:{v}" for k, v in st.most_common(3))
    print(f"{f}:{l:4d} inst {100*i/ti:5.1f}% samp {100*s/ts:5.1f}%  [{top}] {src.strip()[:80]}")
