"""Refresh profiles/linearize_traffic.json from an `ncu --set full` capture of ONE LM iteration's linearize launches
(bench.py reads it for roofline.traffic).
usage: python tools/ncu_traffic.py report.ncu-rep workload@ngpus [LM iterations in the capture, default 1]"""
import csv, io, json, os, subprocess, sys
rep, key = sys.argv[1], sys.argv[2]
niter = int(sys.argv[3]) if len(sys.argv) > 3 else 1
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
h, units = rows[0], rows[1]
tot = 0.0
for r in rows[2:]:
    d = dict(zip(h, r)); u = dict(zip(h, units))
    if "k_linearize" not in d.get("Kernel Name", ""):
        continue
    for m in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
        v = float(d[m].replace(",", "")); unit = u[m].lower()
        v *= {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}.get(unit, 1)
        tot += v
    print(d["Kernel Name"][:70], d["gpu__time_duration.sum"], u["gpu__time_duration.sum"])
path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "profiles", "linearize_traffic.json")
rec = json.load(open(path)) if os.path.exists(path) else {}
tot /= niter
rec[key] = int(tot)
rec["_source"] = rec.get("_source", {})
rec["_source"][key] = os.path.basename(rep)
json.dump(rec, open(path, "w"), indent=1)
print(key, int(tot), "bytes")
